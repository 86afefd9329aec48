"""ctypes binding of the CHOLMOD/B200 C ABI (include/cholmod_b200.h).

This is the host-side mirror of the reference's numeric seam: `Factor(sym)` plays the role of the
GPU-side state the reference hangs on cholmod_common (CHOLMOD/GPU/cholmod_gpu.c), and
`Factor.super_numeric(Ap, Ai, Ax, beta)` has the contract of `cholmod_super_numeric`
(CHOLMOD/Supernodal/cholmod_super_numeric.c:100) for a symmetric lower-stored matrix.

There is no CPU fallback: if lib/libcholmod_b200.so is missing, or no sm_100 device is visible,
everything that computes raises `B200Error`.
"""
import ctypes as C
import os
import numpy as np

OK, NOT_POSDEF, OUT_OF_MEMORY, TOO_LARGE, INVALID, GPU_PROBLEM = 0, 1, -2, -3, -4, -5
DOUBLE, SINGLE, COMPLEX_DOUBLE, COMPLEX_SINGLE = 0, 4, 2, 6
_DTYPE_CODE = {np.dtype(np.float64): DOUBLE, np.dtype(np.float32): SINGLE,
               np.dtype(np.complex128): COMPLEX_DOUBLE, np.dtype(np.complex64): COMPLEX_SINGLE}

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcholmod_b200.so")

# every symbol include/cholmod_b200.h declares
SYMBOLS = ["b200_probe", "b200_device_count", "b200_memorysize", "b200_create", "b200_destroy",
           "b200_get_int", "b200_get_double", "b200_super_numeric_d", "b200_super_numeric_s",
           "b200_upload_matrix", "b200_factor_device", "b200_download_factor", "b200_device_Lx",
           "b200_upload_factor", "b200_solve_d", "b200_version", "b200_set_int", "b200_plan_export",
           "b200_dmma_peak_tflops", "b200_panel_factor_noalloc_d", "b200_create_dist", "b200_comm_unique_id",
           "b200_comm_init", "b200_upload_values", "b200_forget_host_buffer", "b200_super_numeric_z",
           "b200_super_numeric_c", "b200_solve_z", "b200_super_numeric_aat", "b200_upload_matrix_aat",
           "b200_solve_sweep", "b200_set_memory_limit", "b200_peer_arena", "b200_peer_handle", "b200_peer_attach",
           "b200_multi_create", "b200_multi_destroy", "b200_multi_ngpus", "b200_multi_context",
           "b200_multi_super_numeric_d", "b200_multi_download_factor", "b200_multi_forget_host_buffer"]
NOALLOC_MAX_COLS, NOALLOC_MAX_ROWS = 64, 65536


class B200Error(RuntimeError):
    def __init__(self, status, what):
        super().__init__("%s failed with status %d" % (what, status))
        self.status = status


_lib = None


def lib():
    """Load the shared library (raises B200Error if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(GPU_PROBLEM, "loading %s (run `python -c 'import __graft_entry__ as g; g.build()'`)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i64, dbl, cint = C.c_void_p, C.c_int64, C.c_double, C.c_int
    L.b200_probe.argtypes = [cint]; L.b200_probe.restype = cint
    L.b200_device_count.argtypes = []; L.b200_device_count.restype = cint
    L.b200_memorysize.argtypes = [cint, vp, vp]; L.b200_memorysize.restype = cint
    L.b200_create.argtypes = [i64, i64, vp, vp, vp, vp, cint, cint, vp]; L.b200_create.restype = vp
    L.b200_destroy.argtypes = [vp]; L.b200_destroy.restype = None
    L.b200_get_int.argtypes = [vp, C.c_char_p]; L.b200_get_int.restype = i64
    L.b200_get_double.argtypes = [vp, C.c_char_p]; L.b200_get_double.restype = dbl
    for fn in (L.b200_super_numeric_d, L.b200_super_numeric_s):
        fn.argtypes = [vp, vp, vp, vp, vp, dbl, cint, vp, vp]; fn.restype = cint
    for fn in (L.b200_super_numeric_z, L.b200_super_numeric_c):
        fn.argtypes = [vp, vp, vp, vp, vp, vp, dbl, cint, vp, vp]; fn.restype = cint
    L.b200_solve_z.argtypes = [vp, vp, vp, i64]; L.b200_solve_z.restype = cint
    L.b200_super_numeric_aat.argtypes = [vp, i64] + [vp] * 10 + [dbl, cint, vp, vp]
    L.b200_super_numeric_aat.restype = cint
    L.b200_upload_matrix_aat.argtypes = [vp, i64] + [vp] * 8; L.b200_upload_matrix_aat.restype = cint
    L.b200_upload_matrix.argtypes = [vp, vp, vp, vp, vp]; L.b200_upload_matrix.restype = cint
    L.b200_factor_device.argtypes = [vp, dbl, cint, vp]; L.b200_factor_device.restype = cint
    L.b200_download_factor.argtypes = [vp, vp]; L.b200_download_factor.restype = cint
    L.b200_upload_factor.argtypes = [vp, vp]; L.b200_upload_factor.restype = cint
    L.b200_device_Lx.argtypes = [vp]; L.b200_device_Lx.restype = vp
    L.b200_solve_d.argtypes = [vp, vp, vp, i64]; L.b200_solve_d.restype = cint
    L.b200_version.argtypes = []; L.b200_version.restype = C.c_char_p
    L.b200_set_int.argtypes = [vp, C.c_char_p, i64]; L.b200_set_int.restype = cint
    L.b200_dmma_peak_tflops.argtypes = [cint, cint]; L.b200_dmma_peak_tflops.restype = dbl
    L.b200_plan_export.argtypes = [vp, C.c_char_p, vp, i64]; L.b200_plan_export.restype = i64
    L.b200_panel_factor_noalloc_d.argtypes = [vp, i64, i64, i64, vp, vp]
    L.b200_panel_factor_noalloc_d.restype = cint
    L.b200_create_dist.argtypes = [i64, i64, vp, vp, vp, vp, cint, cint, cint, cint, vp]
    L.b200_create_dist.restype = vp
    L.b200_comm_unique_id.argtypes = [vp]; L.b200_comm_unique_id.restype = cint
    L.b200_comm_init.argtypes = [vp, vp]; L.b200_comm_init.restype = cint
    L.b200_upload_values.argtypes = [vp, vp]; L.b200_upload_values.restype = cint
    L.b200_forget_host_buffer.argtypes = [vp, vp]; L.b200_forget_host_buffer.restype = cint
    L.b200_solve_sweep.argtypes = [vp, cint, vp, vp, i64]; L.b200_solve_sweep.restype = cint
    L.b200_set_memory_limit.argtypes = [cint, C.c_size_t]; L.b200_set_memory_limit.restype = cint
    L.b200_peer_arena.argtypes = [vp]; L.b200_peer_arena.restype = vp
    L.b200_peer_handle.argtypes = [vp, vp]; L.b200_peer_handle.restype = cint
    L.b200_peer_attach.argtypes = [vp, cint, vp, vp]; L.b200_peer_attach.restype = cint
    L.b200_multi_create.argtypes = [i64, i64, vp, vp, vp, vp, cint, cint, cint, vp]; L.b200_multi_create.restype = vp
    L.b200_multi_destroy.argtypes = [vp]; L.b200_multi_destroy.restype = None
    L.b200_multi_ngpus.argtypes = [vp]; L.b200_multi_ngpus.restype = cint
    L.b200_multi_context.argtypes = [vp, cint]; L.b200_multi_context.restype = vp
    L.b200_multi_super_numeric_d.argtypes = [vp, vp, vp, vp, vp, dbl, cint, vp, vp]; L.b200_multi_super_numeric_d.restype = cint
    L.b200_multi_download_factor.argtypes = [vp, vp]; L.b200_multi_download_factor.restype = cint
    L.b200_multi_forget_host_buffer.argtypes = [vp, vp]; L.b200_multi_forget_host_buffer.restype = cint
    _lib = L
    return L


def device_count():
    return int(lib().b200_device_count())


def probe(device=0):
    return bool(lib().b200_probe(int(device)))


def set_memory_limit(device, nbytes):
    """Cap on the device memory this library's contexts may hold on `device` (0 = none)."""
    st = lib().b200_set_memory_limit(int(device), int(nbytes))
    if st < 0:
        raise B200Error(st, "b200_set_memory_limit")


def panel_factor_noalloc(d_panel, lda, nrows, jb, stream=0):
    """b200_panel_factor_noalloc_d on a device pointer the caller owns (an int address).  The library
    allocates nothing: for callers that have banned cudaMalloc after start-up or that keep all of HBM in
    their own pool.  Returns (status, info); raises B200Error on a negative status."""
    info = C.c_int64(0)
    st = lib().b200_panel_factor_noalloc_d(C.c_void_p(int(d_panel)), int(lda), int(nrows), int(jb), C.byref(info),
                                           C.c_void_p(int(stream)) if stream else None)
    if st < 0:
        raise B200Error(st, "b200_panel_factor_noalloc_d")
    return int(st), int(info.value)


def comm_unique_id():
    """A fresh NCCL unique id (128 bytes) for b200_comm_init; made by rank 0, carried to the others."""
    buf = (C.c_char * 128)()
    st = lib().b200_comm_unique_id(buf)
    if st < 0:
        raise B200Error(st, "b200_comm_unique_id")
    return bytes(buf)


def dmma_peak_tflops(device=0, iters=20000):
    """Measured FP64 tensor (DMMA) peak of the device in TFLOP/s."""
    return float(lib().b200_dmma_peak_tflops(int(device), int(iters)))


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


class Factor:
    """Device-resident supernodal factorization context for one symbolic factor.

    `sym` is a dict with n, nsuper, super, pi, px, s (the arrays of cholmod_factor after
    cholmod_analyze with a supernodal strategy).  device=-1 builds the host-side plan only.
    """

    def __init__(self, sym, dtype=np.float64, device=0, nranks=1, rank=0):
        self.L = lib()
        self.n, self.nsuper = int(sym["n"]), int(sym["nsuper"])
        self.dtype = np.dtype(dtype)
        if self.dtype not in _DTYPE_CODE:
            raise ValueError("dtype must be float64, float32, complex128 or complex64")
        self.is_complex = self.dtype.kind == "c"
        sup, pi, px, s = _i64(sym["super"]), _i64(sym["pi"]), _i64(sym["px"]), _i64(sym["s"])
        st = C.c_int(0)
        self.nranks, self.rank = int(nranks), int(rank)
        self.h = self.L.b200_create_dist(self.n, self.nsuper, sup.ctypes.data, pi.ctypes.data, px.ctypes.data,
                                         s.ctypes.data, _DTYPE_CODE[self.dtype],
                                         int(device), self.nranks, self.rank, C.byref(st))
        if not self.h:
            raise B200Error(st.value, "b200_create")
        self.xsize = self.geti("xsize")
        self.minor = self.n

    def close(self):
        if getattr(self, "h", None):
            self.L.b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def geti(self, name):
        return int(self.L.b200_get_int(self.h, name.encode()))

    def getd(self, name):
        return float(self.L.b200_get_double(self.h, name.encode()))

    def comm_init(self, unique_id):
        """Join the NCCL communicator of a distributed plan (every rank, same 128-byte id)."""
        buf = (C.c_char * 128).from_buffer_copy(bytes(unique_id))
        st = self.L.b200_comm_init(self.h, buf)
        if st < 0:
            raise B200Error(st, "b200_comm_init")

    def comm_init_torch(self, group=None):
        """Bootstrap a distributed context through torch.distributed: rank 0 makes the NCCL id, everyone gets
        it; every rank publishes the inter-process handle of its arena and attaches to all the others."""
        import torch.distributed as dist
        ids = [comm_unique_id() if self.rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, group=group)
        self.comm_init(ids[0])
        handles = [None] * self.nranks
        dist.all_gather_object(handles, self.peer_handle(), group=group)
        for r, h in enumerate(handles):
            if r != self.rank:
                self.peer_attach(r, handle=h)

    def peer_handle(self):
        buf = (C.c_char * 64)()
        st = self.L.b200_peer_handle(self.h, buf)
        if st < 0:
            raise B200Error(st, "b200_peer_handle")
        return bytes(buf)

    def peer_attach(self, rank, arena=None, handle=None):
        hb = (C.c_char * 64).from_buffer_copy(bytes(handle)) if handle is not None else None
        st = self.L.b200_peer_attach(self.h, int(rank), C.c_void_p(arena) if arena else None, hb)
        if st < 0:
            raise B200Error(st, "b200_peer_attach")

    def seti(self, name, value):
        st = self.L.b200_set_int(self.h, name.encode(), int(value))
        if st < 0:
            raise B200Error(st, "b200_set_int(%s)" % name)

    # -- the drop-in call: host A in, host L->x out ------------------------------------------------
    def super_numeric(self, Ap, Ai, Ax, beta=0.0, quick_return=False, Lx=None, Anz=None, Az=None):
        """Returns (status, Lx); status is OK or NOT_POSDEF (self.minor then holds L->minor).
        Complex contexts: Ax is complex (CHOLMOD_COMPLEX), or Ax and Az are the real and imaginary parts
        (CHOLMOD_ZOMPLEX); Lx is complex either way."""
        Ap, Ai = _i64(Ap), _i64(Ai)
        if Lx is None:
            Lx = np.empty(self.xsize, dtype=self.dtype)
        assert Lx.dtype == self.dtype and Lx.size == self.xsize and Lx.flags.c_contiguous
        nzp = None if Anz is None else _i64(Anz)
        minor = C.c_int64(-1)
        nzptr = None if nzp is None else nzp.ctypes.data
        if self.is_complex:
            real_dt = np.float64 if self.dtype == np.complex128 else np.float32
            fn = self.L.b200_super_numeric_z if self.dtype == np.complex128 else self.L.b200_super_numeric_c
            if Az is None:
                Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
                azptr = None
            else:
                Ax = np.ascontiguousarray(Ax, dtype=real_dt)
                Az = np.ascontiguousarray(Az, dtype=real_dt)
                azptr = Az.ctypes.data
            st = fn(self.h, Ap.ctypes.data, Ai.ctypes.data, nzptr, Ax.ctypes.data, azptr, float(beta),
                    int(quick_return), Lx.ctypes.data, C.byref(minor))
        else:
            Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
            fn = self.L.b200_super_numeric_d if self.dtype == np.float64 else self.L.b200_super_numeric_s
            st = fn(self.h, Ap.ctypes.data, Ai.ctypes.data, nzptr, Ax.ctypes.data, float(beta), int(quick_return),
                    Lx.ctypes.data, C.byref(minor))
        if st < 0:
            raise B200Error(st, "b200_super_numeric")
        self.minor = int(minor.value)
        self._pinned_Lx = Lx       # the library page-locked this buffer: keep it alive until close()
        return st, Lx

    def super_numeric_aat(self, ncol, Ap, Ai, Ax, Fp, Fi, Fx, beta=0.0, quick_return=False, Lx=None, Anz=None,
                          Fnz=None, zomplex=False):
        """The unsymmetric case of cholmod_super_numeric: factorize A*F + beta*I, A n-by-ncol (row-permuted),
        F = A^H ncol-by-n.  zomplex=True passes the values of a complex context as split arrays."""
        Ap, Ai, Fp, Fi = _i64(Ap), _i64(Ai), _i64(Fp), _i64(Fi)
        Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
        Fx = np.ascontiguousarray(Fx, dtype=self.dtype)
        if Lx is None:
            Lx = np.empty(self.xsize, dtype=self.dtype)
        assert Lx.dtype == self.dtype and Lx.size == self.xsize and Lx.flags.c_contiguous
        keep = [None if a is None else _i64(a) for a in (Anz, Fnz)]
        ptr = [None if a is None else a.ctypes.data for a in keep]
        if zomplex:
            assert self.is_complex
            parts = [np.ascontiguousarray(v) for v in (Ax.real, Ax.imag, Fx.real, Fx.imag)]
            ax, az, fx, fz = (v.ctypes.data for v in parts)
        else:
            ax, az, fx, fz = Ax.ctypes.data, None, Fx.ctypes.data, None
        minor = C.c_int64(-1)
        st = self.L.b200_super_numeric_aat(self.h, int(ncol), Ap.ctypes.data, Ai.ctypes.data, ptr[0], ax, az,
                                           Fp.ctypes.data, Fi.ctypes.data, ptr[1], fx, fz, float(beta),
                                           int(quick_return), Lx.ctypes.data, C.byref(minor))
        if st < 0:
            raise B200Error(st, "b200_super_numeric_aat")
        self.minor = int(minor.value)
        self._pinned_Lx = Lx
        return st, Lx

    def upload_matrix_aat(self, ncol, Ap, Ai, Ax, Fp, Fi, Fx):
        Ap, Ai, Fp, Fi = _i64(Ap), _i64(Ai), _i64(Fp), _i64(Fi)
        Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
        Fx = np.ascontiguousarray(Fx, dtype=self.dtype)
        st = self.L.b200_upload_matrix_aat(self.h, int(ncol), Ap.ctypes.data, Ai.ctypes.data, None, Ax.ctypes.data,
                                           Fp.ctypes.data, Fi.ctypes.data, None, Fx.ctypes.data)
        if st < 0:
            raise B200Error(st, "b200_upload_matrix_aat")

    # -- the same in three steps ---------------------------------------------------------------------
    def upload_matrix(self, Ap, Ai, Ax, Anz=None):
        Ap, Ai = _i64(Ap), _i64(Ai)
        Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
        nzp = None if Anz is None else _i64(Anz)
        st = self.L.b200_upload_matrix(self.h, Ap.ctypes.data, Ai.ctypes.data,
                                       None if nzp is None else nzp.ctypes.data, Ax.ctypes.data)
        if st < 0:
            raise B200Error(st, "b200_upload_matrix")

    def upload_values(self, Ax):
        """New values for the pattern of the last upload_matrix (re-factorization fast path)."""
        Ax = np.ascontiguousarray(Ax, dtype=self.dtype)
        st = self.L.b200_upload_values(self.h, Ax.ctypes.data)
        if st < 0:
            raise B200Error(st, "b200_upload_values")

    def factor_device(self, beta=0.0, quick_return=False):
        minor = C.c_int64(-1)
        st = self.L.b200_factor_device(self.h, float(beta), int(quick_return), C.byref(minor))
        if st < 0:
            raise B200Error(st, "b200_factor_device")
        self.minor = int(minor.value)
        return st

    def download_factor(self, Lx=None):
        if Lx is None:
            Lx = np.empty(self.xsize, dtype=self.dtype)
        st = self.L.b200_download_factor(self.h, Lx.ctypes.data)
        if st < 0:
            raise B200Error(st, "b200_download_factor")
        return Lx

    def upload_factor(self, Lx):
        Lx = np.ascontiguousarray(Lx, dtype=self.dtype)
        st = self.L.b200_upload_factor(self.h, Lx.ctypes.data)
        if st < 0:
            raise B200Error(st, "b200_upload_factor")

    def device_Lx(self):
        return self.L.b200_device_Lx(self.h)

    def solve(self, B):
        """X = L'\\(L\\B) in the permuted ordering (L^H for a complex factor); B is n or n-by-nrhs."""
        B = np.asfortranarray(B, dtype=np.complex128 if self.is_complex else np.float64)
        nrhs = 1 if B.ndim == 1 else B.shape[1]
        X = np.empty_like(B, order="F")
        fn = self.L.b200_solve_z if self.is_complex else self.L.b200_solve_d
        st = fn(self.h, B.ctypes.data, X.ctypes.data, nrhs)
        if st < 0:
            raise B200Error(st, "b200_solve_d")
        return X

    def solve_sweep(self, which, B):
        """One half of the solve: which = 1 forward (L y = b), 2 backward (L^H x = y), 3 both."""
        B = np.asfortranarray(B, dtype=np.complex128 if self.is_complex else np.float64)
        nrhs = 1 if B.ndim == 1 else B.shape[1]
        X = np.empty_like(B, order="F")
        st = self.L.b200_solve_sweep(self.h, int(which), B.ctypes.data, X.ctypes.data, nrhs)
        if st < 0:
            raise B200Error(st, "b200_solve_sweep")
        return X

    # -- plan introspection (tests of the host planner) ------------------------------------------------
    def plan_table(self, which, dtype):
        nbytes = int(self.L.b200_plan_export(self.h, which.encode(), None, 0))
        if nbytes < 0:
            raise KeyError(which)
        out = np.zeros(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        if nbytes:
            self.L.b200_plan_export(self.h, which.encode(), out.ctypes.data, nbytes)
        return out


ITEM_DT = np.dtype([("a_off", "<i8"), ("c_off", "<i8"), ("map_off", "<i8"), ("lda", "<i4"), ("ldc", "<i4"),
                    ("M", "<i4"), ("N", "<i4"), ("K", "<i4"), ("shift", "<i4")])
TILE_DT = np.dtype([("item", "<i4"), ("ti", "<i4"), ("tj", "<i4"), ("k0", "<i4"), ("k1", "<i4")])
POTRF_DT = np.dtype([("off", "<i8"), ("col", "<i8"), ("lda", "<i4"), ("jb", "<i4"), ("slot", "<i4"), ("pad", "<i4")])
TRSM_DT = np.dtype([("diag_off", "<i8"), ("row_off", "<i8"), ("lda", "<i4"), ("jb", "<i4"), ("nrows", "<i4"),
                    ("slot", "<i4")])
LAUNCH_DT = np.dtype([("kind", "<i4"), ("count", "<i4"), ("first", "<i8"), ("level", "<i4"), ("stream", "<i4"),
                      ("wait_ev", "<i4"), ("rec_ev", "<i4")])
K_FACTOR_SMALL, K_FACTOR_TINY, K_UPD_TINY = 9, 10, 11
K_JOIN, K_COPYOUT, K_BCAST, K_BARRIER, K_PULL = 12, 13, 14, 15, 16
BCAST_DT = np.dtype([("off", "<i8"), ("len", "<i8"), ("root", "<i4"), ("pad", "<i4")])
COPY_DT = np.dtype([("off", "<i8"), ("len", "<i8"), ("hoff", "<i8")])
PAIR_DT = np.dtype([("d", "<i4"), ("s", "<i4"), ("lpos", "<i4"), ("ndrow1", "<i4"), ("ndrow2", "<i4")])
PULL_DT = np.dtype([("off", "<i8"), ("nrows", "<i4"), ("ncols", "<i4"), ("ld", "<i4"), ("mask", "<u4")])
K_UPD_SIMT, K_UPD_D64, K_UPD_D128, K_POTRF16, K_POTRF32, K_POTRF64, K_TRSM, K_TRSM_INV32, K_TRSM_INV64 = range(9)
K_POTRF_KINDS = (K_POTRF16, K_POTRF32, K_POTRF64)
K_TRSM_KINDS = (K_TRSM, K_TRSM_INV32, K_TRSM_INV64)
KIND_NAMES = ["upd_simt", "upd_dmma64", "upd_dmma128", "potrf16", "potrf32", "potrf64", "trsm_subst", "trsm_inv32",
              "trsm_inv64", "factor_small", "factor_tiny", "upd_tiny"]
# (a K_UPD_TINY "tile" is the whole item)
TILE_SHAPE = {K_UPD_SIMT: (32, 8), K_UPD_D64: (64, 64), K_UPD_D128: (128, 64), K_UPD_TINY: (1 << 30, 1 << 30)}


class Multi:
    """One factorization over `ngpus` devices driven from this process (b200_multi_*): host A in, one complete
    host L->x out, every device moving its share.  Real double symmetric matrices."""

    def __init__(self, sym, first_device=0, ngpus=2):
        self.L = lib()
        self.n, self.xsize = int(sym["n"]), int(sym["xsize"])
        sup, pi, px, s = _i64(sym["super"]), _i64(sym["pi"]), _i64(sym["px"]), _i64(sym["s"])
        st = C.c_int(0)
        self.h = self.L.b200_multi_create(self.n, int(sym["nsuper"]), sup.ctypes.data, pi.ctypes.data, px.ctypes.data,
                                          s.ctypes.data, DOUBLE, int(first_device), int(ngpus), C.byref(st))
        if not self.h:
            raise B200Error(st.value, "b200_multi_create")
        self.ngpus, self.minor = int(ngpus), self.n

    def context_geti(self, rank, name):
        return int(self.L.b200_get_int(self.L.b200_multi_context(self.h, int(rank)), name.encode()))

    def context_getd(self, rank, name):
        return float(self.L.b200_get_double(self.L.b200_multi_context(self.h, int(rank)), name.encode()))

    def super_numeric(self, Ap, Ai, Ax, beta=0.0, quick_return=False, Lx=None):
        Ap, Ai, Ax = _i64(Ap), _i64(Ai), np.ascontiguousarray(Ax, dtype=np.float64)
        if Lx is None:
            Lx = np.empty(self.xsize)
        minor = C.c_int64(-1)
        st = self.L.b200_multi_super_numeric_d(self.h, Ap.ctypes.data, Ai.ctypes.data, None, Ax.ctypes.data, float(beta),
                                               int(quick_return), Lx.ctypes.data, C.byref(minor))
        if st < 0:
            raise B200Error(st, "b200_multi_super_numeric_d")
        self.minor = int(minor.value)
        self._pinned_Lx = Lx
        return st, Lx

    def close(self):
        if getattr(self, "h", None):
            self.L.b200_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
