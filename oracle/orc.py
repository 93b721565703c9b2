"""ctypes binding of the CPU oracle (oracle/liborc.so) and a runner for the reference's
own host build (oracle/_ref/*).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product path.
"""
import ctypes as C
import os
import re
import shutil
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborc.so")
REF_DIR = os.path.join(HERE, "_ref")


class _OrcMat(C.Structure):
    _fields_ = [("rows", C.c_int32), ("cols", C.c_int32), ("k", C.c_int32), ("nnzb", C.c_int64),
                ("brow", C.POINTER(C.c_int32)), ("bcol", C.POINTER(C.c_int32)),
                ("vals", C.POINTER(C.c_uint64))]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            subprocess.check_call(["make", "-C", HERE, "liborc.so"])
        L = C.CDLL(LIB_PATH)
        P = C.POINTER(_OrcMat)
        L.orc_from_coo.restype = P
        L.orc_from_coo.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_free.argtypes = [P]
        L.orc_multiply.restype = P
        L.orc_multiply.argtypes = [P, P, C.POINTER(C.c_int64)]
        L.orc_multiply_mt.restype = P
        L.orc_multiply_mt.argtypes = [P, P, C.POINTER(C.c_int64), C.c_int]
        L.orc_chain.restype = P
        L.orc_chain.argtypes = [C.POINTER(P), C.c_int, C.POINTER(C.c_int64), C.c_char_p, C.c_int64]
        L.orc_chain_ranks.restype = P
        L.orc_chain_ranks.argtypes = [C.POINTER(P), C.c_int, C.c_int, C.POINTER(C.c_int64)]
        L.orc_prune_zero.argtypes = [P]
        L.orc_read.restype = P
        L.orc_read.argtypes = [C.c_char_p, C.c_int32]
        L.orc_write.argtypes = [P, C.c_char_p]
        L.orc_mulmod.restype = C.c_uint64
        L.orc_mulmod.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_addmod.restype = C.c_uint64
        L.orc_addmod.argtypes = [C.c_uint64, C.c_uint64]
        _lib = L
    return _lib


class OMat:
    """Owning handle of an oracle matrix (sorted coordinate list of k x k u64 tiles)."""

    def __init__(self, ptr):
        if not ptr:
            raise RuntimeError("oracle returned NULL")
        self.ptr = ptr

    def __del__(self):
        if getattr(self, "ptr", None) and _lib is not None:
            _lib.orc_free(self.ptr)
            self.ptr = None

    @classmethod
    def from_coo(cls, rows, cols, k, brow, bcol, vals):
        brow = np.ascontiguousarray(brow, dtype=np.int32)
        bcol = np.ascontiguousarray(bcol, dtype=np.int32)
        vals = np.ascontiguousarray(vals, dtype=np.uint64).reshape(-1)
        assert vals.size == brow.size * k * k
        return cls(lib().orc_from_coo(rows, cols, k, brow.size, brow.ctypes.data, bcol.ctypes.data, vals.ctypes.data))

    @classmethod
    def read(cls, path, k):
        return cls(lib().orc_read(path.encode(), k))

    rows = property(lambda s: s.ptr.contents.rows)
    cols = property(lambda s: s.ptr.contents.cols)
    k = property(lambda s: s.ptr.contents.k)
    nnzb = property(lambda s: s.ptr.contents.nnzb)

    def coo(self):
        m = self.ptr.contents
        n, kk = m.nnzb, m.k * m.k
        if n == 0:
            return np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros((0, m.k, m.k), np.uint64)
        br = np.ctypeslib.as_array(m.brow, shape=(n,)).copy()
        bc = np.ctypeslib.as_array(m.bcol, shape=(n,)).copy()
        v = np.ctypeslib.as_array(m.vals, shape=(n * kk,)).copy().reshape(n, m.k, m.k)
        return br, bc, v

    def write(self, path):
        if lib().orc_write(self.ptr, path.encode()) != 0:
            raise OSError("orc_write failed: " + path)

    def prune_zero(self):
        lib().orc_prune_zero(self.ptr)
        return self


def multiply(a, b, threads=1):
    np_ = C.c_int64(0)
    out = OMat(lib().orc_multiply_mt(a.ptr, b.ptr, C.byref(np_), threads))
    return out, np_.value


def chain(mats, ranks=1, want_log=False):
    """helper2 tree (ranks=1) or the P-rank split + rank-0 merge of main (ranks=P)."""
    P = C.POINTER(_OrcMat)
    arr = (P * len(mats))(*[m.ptr for m in mats])
    np_ = C.c_int64(0)
    if ranks == 1:
        buf = C.create_string_buffer(1 << 16) if want_log else None
        out = OMat(lib().orc_chain(arr, len(mats), C.byref(np_), buf, (1 << 16) if want_log else 0))
        if want_log:
            return out, np_.value, buf.value.decode()
        return out, np_.value
    out = OMat(lib().orc_chain_ranks(arr, len(mats), ranks, C.byref(np_)))
    return out, np_.value


# ---- the reference's own translation unit, built host-only (oracle/build_ref.sh) ----------

def have_ref(variant="ref_host_fixed"):
    return os.path.exists(os.path.join(REF_DIR, variant))


def write_folder(folder, mats, k):
    """mats: list of OMat. Writes folder/size and folder/matrix1..N in the reference format."""
    os.makedirs(folder, exist_ok=True)
    with open(os.path.join(folder, "size"), "w") as f:
        f.write("%d %d\n" % (len(mats), k))
    for i, m in enumerate(mats):
        m.write(os.path.join(folder, "matrix%d" % (i + 1)))


_TIMER_RE = re.compile(r"^(\d+) time taken(\d+) ([0-9.eE+-]+) seconds", re.M)


def run_ref(folder, ranks=1, variant="ref_host_fixed", threads=None, keep_output=True, timeout=3600):
    """Run the reference program on `folder`; returns dict(matrix=bytes|None, stdout=str,
    multiplying=[...], time_taken=[...], phases={tag: seconds summed over rank 0's calls})."""
    exe = os.path.join(REF_DIR, variant)
    if not os.path.exists(exe):
        raise FileNotFoundError(exe + " (run oracle/build_ref.sh where /root/reference exists)")
    cwd = tempfile.mkdtemp(prefix="bsp_ref_")
    try:
        env = dict(os.environ, BSP_REF_RANKS=str(ranks))
        if threads is not None:
            env["OMP_NUM_THREADS"] = str(threads)
        p = subprocess.run([exe, os.path.abspath(folder)], cwd=cwd, env=env, stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, timeout=timeout)
        out = p.stdout.decode(errors="replace")
        mpath = os.path.join(cwd, "matrix")
        matrix = None
        if keep_output and os.path.exists(mpath):
            with open(mpath, "rb") as f:
                matrix = f.read()
        phases = {}
        for rk, tag, sec in _TIMER_RE.findall(out):
            if rk == "0":
                phases[tag] = phases.get(tag, 0.0) + float(sec)
        return dict(returncode=p.returncode, matrix=matrix, stdout=out, stderr=p.stderr.decode(errors="replace"),
                    multiplying=re.findall(r"^multiplying (\d+) (\d+)$", out, re.M),
                    time_taken=[float(x) for x in re.findall(r"^time taken ([0-9.eE+-]+) seconds", out, re.M)],
                    phases=phases)
    finally:
        shutil.rmtree(cwd, ignore_errors=True)
