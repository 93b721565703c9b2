"""The C-ABI library loads and exports every symbol include/bspgemm.h declares (no compute
without a GPU), and it fails loudly rather than falling back when there is no CUDA device."""
import ctypes
import os
import re

import pytest

import bspgemm
import util

HEADER = os.path.join(util.ROOT, "include", "bspgemm.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bsp_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    lib = bspgemm.load_library()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "libbspgemm.so does not export " + n
        assert n in bspgemm._SIGNATURES, "bspgemm.py does not bind " + n
    assert b"sm_100a" in lib.bsp_version()


def test_stats_struct_matches_header():
    src = open(HEADER).read()
    body = re.search(r"typedef struct bsp_stats \{(.*?)\} bsp_stats;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            fields += [x.strip() for x in decl.split(None, 1)[1].split(",")]
    assert fields == [n for n, _ in bspgemm.Stats._fields_]


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(bspgemm.BspError) as e:
        bspgemm.Context(0)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)


def test_product_never_links_the_oracle():
    # the shipped library and binding must not reference anything under oracle/
    pkg = os.path.dirname(bspgemm.__file__)
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(root, f), errors="replace").read()
                assert "liborc" not in txt and "import orc" not in txt and "bsp_oracle" not in txt, f
