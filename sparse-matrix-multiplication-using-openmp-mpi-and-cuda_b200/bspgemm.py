"""ctypes binding of libbspgemm.so -- the host-side mirror of the reference's interface for
its one hot path (sparse_matrix_mult.cu, "cu:N"):

    one_matrix            cu:26-32   ->  class Matrix (device-resident CSR of blocks)
    helper(a, b, k, ...)  cu:97-286  ->  Context.helper(a, b)      one product A*B
    helper2(end, start, arr, k, ...) cu:287-327 -> Context.helper2(arr)   pairwise-tree chain
    zero-tile prune + writer cu:573-608 -> Matrix.prune_zero(), Matrix.write_text(path)
    reader                cu:340-391 ->  Context.read_text(path, k)

The product is the CUDA library; this module only marshals arguments.  There is no CPU
fallback: if libbspgemm.so is missing, or no CUDA device is present, calls raise.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libbspgemm.so")

BSP_OK = 0
ERR_NAMES = {1: "BSP_ERR_ARG", 2: "BSP_ERR_CUDA", 3: "BSP_ERR_NOMEM", 4: "BSP_ERR_LIMIT", 5: "BSP_ERR_IO"}


class BspError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s: %s" % (ERR_NAMES.get(code, str(code)), msg))
        self.code = code


class Stats(C.Structure):
    _fields_ = [("nnzb_a", C.c_int64), ("nnzb_b", C.c_int64), ("nnzb_c", C.c_int64), ("nprod", C.c_int64),
                ("ms_symbolic", C.c_double), ("ms_numeric", C.c_double), ("ms_total", C.c_double),
                ("launches", C.c_int64), ("exact_refolds", C.c_int64), ("numeric_path", C.c_int32),
                ("reserved", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


# every symbol include/bspgemm.h declares: name -> (restype, argtypes)
_VP = C.c_void_p
_SIGNATURES = {
    "bsp_version": (C.c_char_p, []),
    "bsp_ctx_create": (C.c_int, [C.c_int, _VP, C.POINTER(_VP)]),
    "bsp_ctx_destroy": (None, [_VP]),
    "bsp_last_error": (C.c_char_p, [_VP]),
    "bsp_ctx_sync": (C.c_int, [_VP]),
    "bsp_ctx_launch_count": (C.c_int64, [_VP]),
    "bsp_mat_from_coo": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_from_coo_f32": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_to_coo_f32": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "bsp_mat_is_f32": (C.c_int, [_VP]),
    "bsp_ctx_set_f32_path": (C.c_int, [_VP, C.c_int]),
    "bsp_mat_from_device_coo": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_info": (C.c_int, [_VP, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "bsp_mat_to_coo": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "bsp_mat_device_view": (C.c_int, [_VP, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP)]),
    "bsp_mat_export_device": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "bsp_mat_copy_to_ctx": (C.c_int, [_VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_from_device_bcsr": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_from_device_bcsr_f32": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, C.POINTER(_VP)]),
    "bsp_mat_free": (None, [_VP, _VP]),
    "bsp_spgemm": (C.c_int, [_VP, _VP, _VP, C.POINTER(_VP), C.POINTER(Stats)]),
    "bsp_spgemm_panel": (C.c_int, [_VP, _VP, _VP, C.c_int32, C.c_int32, C.POINTER(_VP), C.POINTER(Stats)]),
    "bsp_mat_concat_panels": (C.c_int, [_VP, C.POINTER(_VP), C.c_int32, C.POINTER(_VP)]),
    "bsp_chain": (C.c_int, [_VP, C.POINTER(_VP), C.c_int32, C.POINTER(_VP), C.POINTER(Stats), C.c_char_p, C.c_size_t, C.c_int32]),
    "bsp_prune_zero": (C.c_int, [_VP, _VP]),
    "bsp_multiply_host": (C.c_int, [_VP, C.c_int32,
                                    C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP,
                                    C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP,
                                    C.c_int64, C.POINTER(C.c_int64), _VP, _VP, _VP, C.POINTER(Stats)]),
    "bsp_prune_zero_noalloc": (C.c_int, [_VP, C.c_int32, C.c_int32, C.c_int64, _VP, _VP, _VP, _VP, _VP, _VP,
                                         C.POINTER(C.c_int64)]),
    "bsp_mat_read_text": (C.c_int, [_VP, C.c_char_p, C.c_int32, C.POINTER(_VP)]),
    "bsp_mat_write_text": (C.c_int, [_VP, _VP, C.c_char_p]),
    "bsp_mat_read_text_many": (C.c_int, [_VP, C.POINTER(C.c_char_p), C.c_int32, C.c_int32, C.POINTER(_VP)]),
    "bsp_mat_checksum": (C.c_int, [_VP, _VP, C.POINTER(C.c_uint64)]),
    "bsp_mat_checksum_body": (C.c_int, [_VP, _VP, C.POINTER(C.c_uint64)]),
    "bsp_checksum_finish": (C.c_uint64, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_uint64]),
    "bsp_comm_unique_id": (C.c_int, [_VP, C.c_size_t]),
    "bsp_comm_create": (C.c_int, [_VP, _VP, C.c_int, C.c_int, C.POINTER(_VP)]),
    "bsp_comm_adopt": (C.c_int, [_VP, _VP, C.c_int, C.c_int, C.POINTER(_VP)]),
    "bsp_comm_destroy": (None, [_VP, _VP]),
    "bsp_comm_info": (C.c_int, [_VP, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64)]),
    "bsp_mat_send": (C.c_int, [_VP, _VP, _VP, C.c_int]),
    "bsp_mat_recv": (C.c_int, [_VP, _VP, C.c_int, C.POINTER(_VP)]),
    "bsp_mat_bcast": (C.c_int, [_VP, _VP, C.c_int, _VP, C.POINTER(_VP)]),
    "bsp_gather_panels": (C.c_int, [_VP, _VP, _VP, C.c_int, C.POINTER(_VP)]),
    "bsp_chain_distributed": (C.c_int, [_VP, _VP, C.POINTER(_VP), C.c_int32, C.c_int32, C.POINTER(_VP), C.POINTER(Stats),
                                        C.c_char_p, C.c_size_t]),
    "bsp_panel_bounds": (C.c_int, [_VP, _VP, _VP, C.c_int32, C.POINTER(C.c_int32)]),
    "bsp_mat_write_bin": (C.c_int, [_VP, _VP, C.c_char_p]),
    "bsp_mat_read_bin": (C.c_int, [_VP, C.c_char_p, C.POINTER(_VP)]),
}

_lib = None


def load_library(path=LIB_PATH):
    """Load libbspgemm.so and bind every declared symbol. Raises if the library is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(path):
            raise FileNotFoundError(
                path + " not built: run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc); "
                "there is no CPU fallback")
        L = C.CDLL(path)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64).reshape(-1)


class Matrix:
    """Device-resident block-sparse matrix (the reference's one_matrix, cu:26-32)."""

    def __init__(self, ctx, handle):
        self.ctx = ctx
        self.handle = handle

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def free(self):
        if self.handle and self.ctx and self.ctx.handle:
            self.ctx.lib.bsp_mat_free(self.ctx.handle, self.handle)
        self.handle = None

    def info(self):
        r, c, k, n = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64()
        self.ctx._ck(self.ctx.lib.bsp_mat_info(self.handle, C.byref(r), C.byref(c), C.byref(k), C.byref(n)))
        return r.value, c.value, k.value, n.value

    rows = property(lambda s: s.info()[0])
    cols = property(lambda s: s.info()[1])
    k = property(lambda s: s.info()[2])
    blocks = property(lambda s: s.info()[3])  # one_matrix::blocks
    nnzb = blocks

    @property
    def is_f32(self):
        return bool(self.ctx.lib.bsp_mat_is_f32(self.handle))

    def to_coo(self, out=None):
        """(brow, bcol, vals[nnzb,k,k]) sorted by (brow, bcol). `out` = preallocated triple.
        vals are uint64, or float32 for a float matrix."""
        _, _, k, n = self.info()
        f32 = self.is_f32
        if out is None:
            out = (np.empty(n, np.int32), np.empty(n, np.int32), np.empty((n, k, k), np.float32 if f32 else np.uint64))
        br, bc, v = out
        fn = self.ctx.lib.bsp_mat_to_coo_f32 if f32 else self.ctx.lib.bsp_mat_to_coo
        self.ctx._ck(fn(self.ctx.handle, self.handle, br.ctypes.data, bc.ctypes.data, v.ctypes.data))
        return br, bc, v

    def device_view(self):
        """(nbrows, nbcols, rowptr_ptr, colidx_ptr, vals_ptr) raw device addresses."""
        nr, nc = C.c_int32(), C.c_int32()
        p1, p2, p3 = _VP(), _VP(), _VP()
        self.ctx._ck(self.ctx.lib.bsp_mat_device_view(self.handle, C.byref(nr), C.byref(nc), C.byref(p1), C.byref(p2),
                                                      C.byref(p3)))
        return nr.value, nc.value, p1.value or 0, p2.value or 0, p3.value or 0

    def export_device(self, d_rowptr, d_colidx, d_vals):
        """Copy rowptr / colidx / vals into caller-owned device buffers (raw addresses)."""
        self.ctx._ck(self.ctx.lib.bsp_mat_export_device(self.ctx.handle, self.handle, d_rowptr, d_colidx, d_vals))

    def prune_zero(self):
        self.ctx._ck(self.ctx.lib.bsp_prune_zero(self.ctx.handle, self.handle))
        return self

    def write_text(self, path):
        self.ctx._ck(self.ctx.lib.bsp_mat_write_text(self.ctx.handle, self.handle, os.fsencode(path)))

    def write_bin(self, path):
        self.ctx._ck(self.ctx.lib.bsp_mat_write_bin(self.ctx.handle, self.handle, os.fsencode(path)))

    def checksum(self):
        """64-bit content checksum computed on the device (csrc/bsp_checksum.cu)."""
        v = C.c_uint64()
        self.ctx._ck(self.ctx.lib.bsp_mat_checksum(self.ctx.handle, self.handle, C.byref(v)))
        return v.value

    def checksum_body(self):
        """Order-independent tile sum of this matrix / row panel (add the panels' sums mod 2^64, then
        checksum_finish with the header of the whole matrix)."""
        v = C.c_uint64()
        self.ctx._ck(self.ctx.lib.bsp_mat_checksum_body(self.ctx.handle, self.handle, C.byref(v)))
        return v.value


def checksum_finish(rows, cols, k, esize, nnzb, body):
    return load_library().bsp_checksum_finish(rows, cols, k, esize, nnzb, body & 0xFFFFFFFFFFFFFFFF)


COMM_ID_BYTES = 128


def comm_unique_id():
    """Rendezvous token for Comm (made by one rank, handed to the others by any means)."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    if load_library().bsp_comm_unique_id(buf, COMM_ID_BYTES) != BSP_OK:
        raise BspError(2, "bsp_comm_unique_id failed (NCCL not available?)")
    return buf.raw


class Comm:
    """One NCCL communicator of the C ABI (replaces MPI_COMM_WORLD of cu:404-408): collective constructor."""

    def __init__(self, ctx, unique_id, nranks, rank):
        self.ctx = ctx
        h = _VP()
        ctx._ck(ctx.lib.bsp_comm_create(ctx.handle, unique_id, nranks, rank, C.byref(h)))
        self.handle = h.value
        self.nranks, self.rank = nranks, rank

    def close(self):
        if self.handle and self.ctx.handle:
            self.ctx.lib.bsp_comm_destroy(self.ctx.handle, self.handle)
        self.handle = None

    @property
    def bytes_sent(self):
        v = C.c_int64()
        self.ctx._ck(self.ctx.lib.bsp_comm_info(self.handle, None, None, C.byref(v)))
        return v.value

    def send(self, mat, peer):
        self.ctx._ck(self.ctx.lib.bsp_mat_send(self.ctx.handle, self.handle, mat.handle, peer))

    def recv(self, peer):
        h = _VP()
        self.ctx._ck(self.ctx.lib.bsp_mat_recv(self.ctx.handle, self.handle, peer, C.byref(h)))
        return Matrix(self.ctx, h.value)

    def bcast(self, mat, root):
        """Returns the root's matrix on every rank (the root gets its own object back)."""
        h = _VP()
        self.ctx._ck(self.ctx.lib.bsp_mat_bcast(self.ctx.handle, self.handle, root, mat.handle if mat is not None else None,
                                                C.byref(h)))
        return mat if self.rank == root else Matrix(self.ctx, h.value)

    def gather_panels(self, panel, root=0):
        h = _VP()
        self.ctx._ck(self.ctx.lib.bsp_gather_panels(self.ctx.handle, self.handle, panel.handle, root, C.byref(h)))
        return Matrix(self.ctx, h.value) if self.rank == root else None

    def chain(self, arr, stats=None, log_base=0, want_log=False):
        """The multi-rank chain of main (cu:438-571): `arr` is THIS rank's sub-chain. Result on rank 0."""
        n = len(arr)
        hs = (_VP * max(n, 1))(*[m.handle for m in arr])
        h = _VP()
        st = stats if stats is not None else Stats()
        buf = C.create_string_buffer(64 * max(n, 1) + 64) if want_log else None
        self.ctx._ck(self.ctx.lib.bsp_chain_distributed(self.ctx.handle, self.handle, hs, n, log_base, C.byref(h), C.byref(st),
                                                        buf, len(buf) if buf else 0))
        out = Matrix(self.ctx, h.value) if h.value else None
        return (out, buf.value.decode()) if want_log else out


class Context:
    """One device + one stream (replaces the staging buffers main lends to helper, cu:421-436)."""

    def __init__(self, device=0, stream=None):
        self.lib = load_library()
        h = _VP()
        rc = self.lib.bsp_ctx_create(device, _VP(stream) if stream else None, C.byref(h))
        self.handle = h.value
        if rc != BSP_OK:
            msg = self.lib.bsp_last_error(self.handle).decode() if self.handle else "context creation failed"
            if self.handle:
                self.lib.bsp_ctx_destroy(self.handle)
                self.handle = None
            raise BspError(rc, msg)

    def close(self):
        if self.handle:
            self.lib.bsp_ctx_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != BSP_OK:
            raise BspError(rc, self.lib.bsp_last_error(self.handle).decode())

    def sync(self):
        self._ck(self.lib.bsp_ctx_sync(self.handle))

    def launch_count(self):
        return self.lib.bsp_ctx_launch_count(self.handle)

    # ---- matrices -------------------------------------------------------------------------
    def matrix_from_coo(self, rows, cols, k, brow, bcol, vals):
        br, bc, v = _i32(brow), _i32(bcol), _u64(vals)
        if v.size != br.size * k * k or bc.size != br.size:
            raise ValueError("coordinate list shape mismatch")
        h = _VP()
        self._ck(self.lib.bsp_mat_from_coo(self.handle, rows, cols, k, br.size, br.ctypes.data, bc.ctypes.data,
                                           v.ctypes.data, C.byref(h)))
        return Matrix(self, h.value)

    def matrix_from_coo_f32(self, rows, cols, k, brow, bcol, vals):
        """float32 tiles (extension, BASELINE.json configs[4])."""
        br, bc = _i32(brow), _i32(bcol)
        v = np.ascontiguousarray(vals, dtype=np.float32).reshape(-1)
        if v.size != br.size * k * k or bc.size != br.size:
            raise ValueError("coordinate list shape mismatch")
        h = _VP()
        self._ck(self.lib.bsp_mat_from_coo_f32(self.handle, rows, cols, k, br.size, br.ctypes.data, bc.ctypes.data,
                                               v.ctypes.data, C.byref(h)))
        return Matrix(self, h.value)

    def set_f32_path(self, path):
        """'auto' | 'simt' | 'tensor' for float32 products."""
        self._ck(self.lib.bsp_ctx_set_f32_path(self.handle, {"auto": 0, "simt": 1, "tensor": 2}[path]))

    def matrix_from_device_coo(self, rows, cols, k, nnzb, d_brow, d_bcol, d_vals):
        h = _VP()
        self._ck(self.lib.bsp_mat_from_device_coo(self.handle, rows, cols, k, nnzb, d_brow, d_bcol, d_vals, C.byref(h)))
        return Matrix(self, h.value)

    def matrix_from_device_bcsr(self, rows, cols, k, nbrows, nbcols, nnzb, d_rowptr, d_colidx, d_vals):
        h = _VP()
        self._ck(self.lib.bsp_mat_from_device_bcsr(self.handle, rows, cols, k, nbrows, nbcols, nnzb, d_rowptr,
                                                   d_colidx, d_vals, C.byref(h)))
        return Matrix(self, h.value)

    def read_text(self, path, k):
        h = _VP()
        self._ck(self.lib.bsp_mat_read_text(self.handle, os.fsencode(path), k, C.byref(h)))
        return Matrix(self, h.value)

    def read_text_many(self, paths, k):
        """Parse several files with upload/ingestion overlapped with parsing; list of Matrix."""
        n = len(paths)
        arr = (C.c_char_p * n)(*[os.fsencode(p) for p in paths])
        hs = (_VP * n)()
        self._ck(self.lib.bsp_mat_read_text_many(self.handle, arr, n, k, hs))
        return [Matrix(self, hs[i]) for i in range(n)]

    def read_bin(self, path):
        h = _VP()
        self._ck(self.lib.bsp_mat_read_bin(self.handle, os.fsencode(path), C.byref(h)))
        return Matrix(self, h.value)

    # ---- the hot path -----------------------------------------------------------------------
    def helper(self, a, b, stats=None, panel=None):
        """C = A*B (cu:97-286). panel=(row_lo,row_hi) restricts to a block-row panel."""
        h = _VP()
        st = stats if stats is not None else Stats()
        if panel is None:
            rc = self.lib.bsp_spgemm(self.handle, a.handle, b.handle, C.byref(h), C.byref(st))
        else:
            rc = self.lib.bsp_spgemm_panel(self.handle, a.handle, b.handle, panel[0], panel[1], C.byref(h), C.byref(st))
        if rc != BSP_OK and h.value:
            self.lib.bsp_mat_free(self.handle, h.value)
        self._ck(rc)
        return Matrix(self, h.value)

    spgemm = helper

    def helper2(self, arr, stats=None, want_log=False, log_base=0):
        """Pairwise-tree product of the list `arr` (cu:287-327). Inputs are not modified."""
        n = len(arr)
        hs = (_VP * n)(*[m.handle for m in arr])
        h = _VP()
        st = stats if stats is not None else Stats()
        buf = C.create_string_buffer(64 * max(n, 1) + 64) if want_log else None
        self._ck(self.lib.bsp_chain(self.handle, hs, n, C.byref(h), C.byref(st), buf, len(buf) if buf else 0, log_base))
        out = Matrix(self, h.value)
        return (out, buf.value.decode()) if want_log else out

    chain = helper2

    def concat_panels(self, panels):
        """Concatenate block-row panels (ascending, disjoint row ranges of one product)."""
        n = len(panels)
        hs = (_VP * n)(*[m.handle for m in panels])
        h = _VP()
        self._ck(self.lib.bsp_mat_concat_panels(self.handle, hs, n, C.byref(h)))
        return Matrix(self, h.value)

    def panel_bounds(self, a, b, nparts):
        """Row-panel bounds of a*b for nparts GPUs, balanced by tile products per block row (device-side)."""
        out = (C.c_int32 * (nparts + 1))()
        self._ck(self.lib.bsp_panel_bounds(self.handle, a.handle, b.handle, nparts, out))
        return list(out)

    def prune_zero_noalloc(self, k, nbrows, nnzb, d_rowptr, d_colidx, d_vals, d_rowptr_out, d_colidx_out, d_vals_out):
        """Zero-tile compaction that allocates nothing (raw device addresses in and out; the
        caller owns every buffer). Returns the number of tiles kept. See include/bspgemm.h for
        the size limit and the no-overlap rule."""
        n = C.c_int64(0)
        self._ck(self.lib.bsp_prune_zero_noalloc(self.handle, k, nbrows, nnzb, d_rowptr, d_colidx, d_vals,
                                                 d_rowptr_out, d_colidx_out, d_vals_out, C.byref(n)))
        return n.value

    def multiply_host(self, k, a, b, out=None, stats=None):
        """Host buffers in, host buffers out in one C-ABI call. a, b = (rows, cols, brow, bcol, vals).
        out = preallocated (brow, bcol, vals) or None (two calls: size, then fetch)."""
        ra, ca, bra, bca, va = a
        rb, cb, brb, bcb, vb = b
        bra, bca, va, brb, bcb, vb = _i32(bra), _i32(bca), _u64(va), _i32(brb), _i32(bcb), _u64(vb)
        n = C.c_int64(0)
        st = stats if stats is not None else Stats()

        def call(cap, o):
            return self.lib.bsp_multiply_host(
                self.handle, k, ra, ca, bra.size, bra.ctypes.data, bca.ctypes.data, va.ctypes.data,
                rb, cb, brb.size, brb.ctypes.data, bcb.ctypes.data, vb.ctypes.data,
                cap, C.byref(n), o[0], o[1], o[2], C.byref(st))
        if out is None:
            self._ck(call(0, (None, None, None)))
            out = (np.empty(n.value, np.int32), np.empty(n.value, np.int32), np.empty((n.value, k, k), np.uint64))
        self._ck(call(out[0].size, (out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data)))
        m = n.value
        return out[0][:m], out[1][:m], out[2][:m]
