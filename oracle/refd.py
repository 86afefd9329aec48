"""oracle/refd.py -- TEST INFRASTRUCTURE: ctypes bindings for the oracle.

`Ref` wraps oracle/_ref/libref_driver.so (oracle/ref_driver.c, a plain-pointer driver over the
unmodified reference CHOLMOD in oracle/_ref/libcholmod_ref.so); `port_factor` / `port_solve`
wrap the plain-C restatement oracle/super_numeric_port.c (oracle/_ref/libsnport.so).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200_REFD_FLAVOUR=cuda: the reference built with CHOLMOD_HAS_CUDA (its own cuBLAS-offload GPU branch,
# oracle/Makefile target `cuda`), used by `bench.py --impl reference-gpu` only
_REF = os.path.join(_HERE, "_ref_cuda" if os.environ.get("B200_REFD_FLAVOUR") == "cuda" else "_ref")

CHOLMOD_NATURAL, CHOLMOD_GIVEN, CHOLMOD_AMD, CHOLMOD_METIS, CHOLMOD_NESDIS = 0, 1, 2, 3, 4
CHOLMOD_SIMPLICIAL, CHOLMOD_AUTO, CHOLMOD_SUPERNODAL = 0, 1, 2
CHOLMOD_OK, CHOLMOD_NOT_POSDEF = 0, 1


def have_ref():
    return os.path.exists(os.path.join(_REF, "libref_driver.so"))


def have_port():
    return os.path.exists(os.path.join(_REF, "libsnport.so"))


_lib = None


def _driver():
    global _lib
    if _lib is None:
        # the reference goes into the global symbol scope, like a library an application links against,
        # so that a preloaded libcholmod_b200_shim.so can call back into it (tests/test_gpu_dropin.py)
        C.CDLL(os.path.join(_REF, "libcholmod_ref.so"), mode=C.RTLD_GLOBAL)
        lib = C.CDLL(os.path.join(_REF, "libref_driver.so"))
        lib.refd_open.restype = C.c_void_p
        lib.refd_open.argtypes = [C.c_int]
        lib.refd_close.argtypes = [C.c_void_p]
        lib.refd_set_matrix.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        lib.refd_set_matrix_x.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int, C.c_int, C.c_int]
        lib.refd_set_matrix_rect.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int, C.c_int]
        lib.refd_read_matrix.argtypes = [C.c_void_p, C.c_char_p]
        lib.refd_analyze.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.refd_seam.argtypes = [C.c_void_p]
        lib.refd_super_numeric.argtypes = [C.c_void_p, C.c_double]
        lib.refd_factorize.argtypes = [C.c_void_p, C.c_double]
        lib.refd_numeric_Lx.restype = C.c_void_p
        lib.refd_numeric_Lx.argtypes = [C.c_void_p, C.c_int]
        lib.refd_set_minor.argtypes = [C.c_void_p, C.c_int64]
        lib.refd_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        lib.refd_resid.restype = C.c_double
        lib.refd_resid.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.refd_set_int.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
        lib.refd_get_int.restype = C.c_int64
        lib.refd_get_int.argtypes = [C.c_void_p, C.c_char_p]
        lib.refd_get_double.restype = C.c_double
        lib.refd_get_double.argtypes = [C.c_void_p, C.c_char_p]
        lib.refd_rcond.restype = C.c_double
        lib.refd_rcond.argtypes = [C.c_void_p]
        lib.refd_get_ptr.restype = C.c_void_p
        lib.refd_get_ptr.argtypes = [C.c_void_p, C.c_char_p]
        _lib = lib
    return _lib


def _view(ptr, count, dtype):
    if count == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


class Ref:
    """One reference CHOLMOD session (cholmod_common + A + symbolic/numeric L)."""

    def __init__(self, nthreads=0):
        self.lib = _driver()
        self.h = self.lib.refd_open(int(nthreads))
        self.single = False
        self.xtype = 1          # CHOLMOD_REAL 1, CHOLMOD_COMPLEX 2, CHOLMOD_ZOMPLEX 3

    def _xd(self):
        """numpy dtype of the entries of a complex-or-real array of this session (S, L)."""
        if self.xtype == 1:
            return np.float32 if self.single else np.float64
        return np.complex64 if self.single else np.complex128

    def close(self):
        if self.h:
            self.lib.refd_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def geti(self, name):
        return int(self.lib.refd_get_int(self.h, name.encode()))

    def seti(self, name, v):
        if self.lib.refd_set_int(self.h, name.encode(), int(v)):
            raise KeyError(name)

    def getd(self, name):
        return float(self.lib.refd_get_double(self.h, name.encode()))

    def _ptr(self, name):
        return self.lib.refd_get_ptr(self.h, name.encode())

    def set_matrix(self, n, Ap, Ai, Ax, stype=-1, single=False, zomplex=False):
        """Real Ax: a symmetric matrix.  Complex Ax: a Hermitian one, stored as CHOLMOD_COMPLEX, or as
        CHOLMOD_ZOMPLEX (split real / imaginary arrays) with zomplex=True."""
        Ap = np.ascontiguousarray(Ap, dtype=np.int64)
        Ai = np.ascontiguousarray(Ai, dtype=np.int64)
        self.single = bool(single)
        if np.iscomplexobj(Ax):
            Ax = np.ascontiguousarray(Ax, dtype=np.complex128)
            if zomplex:
                self.xtype = 3
                re, im = np.ascontiguousarray(Ax.real), np.ascontiguousarray(Ax.imag)
                rc = self.lib.refd_set_matrix_x(self.h, n, Ap.ctypes.data, Ai.ctypes.data, re.ctypes.data,
                                                im.ctypes.data, stype, int(single), 3)
            else:
                self.xtype = 2
                rc = self.lib.refd_set_matrix_x(self.h, n, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data, None,
                                                stype, int(single), 2)
        else:
            self.xtype = 1
            Ax = np.ascontiguousarray(Ax, dtype=np.float64)
            rc = self.lib.refd_set_matrix(self.h, n, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data, stype,
                                          int(single))
        if rc:
            raise RuntimeError("refd_set_matrix failed")

    def set_matrix_rect(self, nrow, ncol, Ap, Ai, Ax, single=False, zomplex=False):
        """An unsymmetric nrow-by-ncol A (stype 0): analysis and factorization are those of A*A^H."""
        Ap = np.ascontiguousarray(Ap, dtype=np.int64)
        Ai = np.ascontiguousarray(Ai, dtype=np.int64)
        self.single = bool(single)
        if np.iscomplexobj(Ax):
            Ax = np.ascontiguousarray(Ax, dtype=np.complex128)
            self.xtype = 3 if zomplex else 2
            if zomplex:
                re, im = np.ascontiguousarray(Ax.real), np.ascontiguousarray(Ax.imag)
                rc = self.lib.refd_set_matrix_rect(self.h, nrow, ncol, Ap.ctypes.data, Ai.ctypes.data, re.ctypes.data,
                                                   im.ctypes.data, int(single), 3)
            else:
                rc = self.lib.refd_set_matrix_rect(self.h, nrow, ncol, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data,
                                                   None, int(single), 2)
        else:
            self.xtype = 1
            Ax = np.ascontiguousarray(Ax, dtype=np.float64)
            rc = self.lib.refd_set_matrix_rect(self.h, nrow, ncol, Ap.ctypes.data, Ai.ctypes.data, Ax.ctypes.data,
                                               None, int(single), 1)
        if rc:
            raise RuntimeError("refd_set_matrix_rect failed")

    def read_matrix(self, path):
        rc = self.lib.refd_read_matrix(self.h, path.encode())
        if rc:
            raise RuntimeError("refd_read_matrix failed: %d" % rc)
        self.xtype = self.geti("xtypeA")            # 1 real, 2 complex, 3 zomplex
        self.single = self.geti("dtypeA") == 4

    def matrix(self):
        n = self.geti("n")
        Ap, Ai, Ax = self._csc("A", self.geti("ncol"), self.geti("nnzA"))
        return (n, Ap, Ai, Ax, self.geti("stypeA"))

    def analyze(self, ordering=CHOLMOD_AMD, supernodal=CHOLMOD_SUPERNODAL):
        rc = self.lib.refd_analyze(self.h, ordering, supernodal)
        if rc < 0:
            raise RuntimeError("refd_analyze failed: %d" % rc)
        return rc

    def symbolic(self):
        """The supernodal symbolic factor as numpy copies (what the numeric seam reads)."""
        n, ns, ss = self.geti("n"), self.geti("nsuper"), self.geti("ssize")
        return dict(
            n=n, nsuper=ns, ssize=ss, xsize=self.geti("xsize"), maxcsize=self.geti("maxcsize"),
            maxesize=self.geti("maxesize"), fl=self.getd("fl"), lnz=self.getd("lnz"),
            Perm=_view(self._ptr("Perm"), n, np.int64).copy(),
            super=_view(self._ptr("super"), ns + 1, np.int64).copy(),
            pi=_view(self._ptr("pi"), ns + 1, np.int64).copy(),
            px=_view(self._ptr("px"), ns + 1, np.int64).copy(),
            s=_view(self._ptr("s"), ss, np.int64).copy())

    def seam(self):
        """S = tril(P A P') in CSC, as cholmod_factorize_p hands it to cholmod_super_numeric."""
        rc = self.lib.refd_seam(self.h)
        if rc:
            raise RuntimeError("refd_seam failed")
        return self._csc("S", self.geti("ncol"), self.geti("nnzS"))

    def _csc(self, which, ncols, nz):
        p, i = _view(self._ptr(which + "p"), ncols + 1, np.int64).copy(), _view(self._ptr(which + "i"), nz, np.int64).copy()
        if self.xtype == 3:     # zomplex: returned as one complex array
            rd = np.float32 if self.single else np.float64
            x = _view(self._ptr(which + "x"), nz, rd) + 1j * _view(self._ptr(which + "z"), nz, rd)
            return p, i, x.astype(self._xd())
        return p, i, _view(self._ptr(which + "x"), nz, self._xd()).copy()

    def seam_F(self):
        """Unsymmetric case, after seam(): F = A(p,:)^H (ncol-by-n), the second operand of cholmod_super_numeric."""
        return self._csc("F", self.geti("n"), self.geti("nnzF"))

    def super_numeric(self, beta=0.0):
        """Reference CPU numeric phase at the seam; returns Common->status."""
        return int(self.lib.refd_super_numeric(self.h, float(beta)))

    def factorize(self, beta=0.0):
        """Reference cholmod_l_factorize_p (permute + numeric); returns Common->status."""
        return int(self.lib.refd_factorize(self.h, float(beta)))

    def Lx(self):
        single = self.geti("dtype") == 4
        if self.geti("xtype") == 2:
            xd = np.complex64 if single else np.complex128
        else:
            xd = np.float32 if single else np.float64
        return _view(self._ptr("Lx"), self.geti("xsize"), xd)

    def numeric_Lx(self):
        """Host L->x (allocated like cholmod_super_numeric.c:219-231 if L is still symbolic)."""
        p = self.lib.refd_numeric_Lx(self.h, int(self.single))
        if not p:
            raise RuntimeError("refd_numeric_Lx failed")
        return _view(p, self.geti("xsize"), self._xd())

    def set_minor(self, minor):
        self.lib.refd_set_minor(self.h, int(minor))

    def rcond(self):
        return float(_driver().refd_rcond(self.h))

    def solve(self, b):
        b = np.asfortranarray(b, dtype=np.float64 if self.xtype == 1 else np.complex128)
        n = self.geti("n")
        nrhs = 1 if b.ndim == 1 else b.shape[1]
        x = np.zeros_like(b, order="F")
        rc = self.lib.refd_solve(self.h, b.ctypes.data, x.ctypes.data, nrhs)
        if rc:
            raise RuntimeError("refd_solve failed: %d" % rc)
        assert x.shape[0] == n
        return x

    def resid(self, x, b):
        if self.xtype != 1:
            return self._resid_complex(x, b)
        x = np.ascontiguousarray(x, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        return float(self.lib.refd_resid(self.h, x.ctypes.data, b.ctypes.data))

    def _resid_complex(self, x, b):
        """The same scaled residual for a Hermitian A (one triangle stored), in numpy."""
        n, nz = self.geti("n"), self.geti("nnzA")
        Ap, Ai = _view(self._ptr("Ap"), n + 1, np.int64), _view(self._ptr("Ai"), nz, np.int64)
        rd = np.float32 if self.single else np.float64
        if self.xtype == 3:
            Ax = _view(self._ptr("Ax"), nz, rd).astype(np.float64) + 1j * _view(self._ptr("Az"), nz, rd)
        else:
            Ax = _view(self._ptr("Ax"), nz, np.complex64 if self.single else np.complex128).astype(np.complex128)
        col = np.repeat(np.arange(n), np.diff(Ap))
        stype = self.geti("stypeA")
        keep = (Ai >= col) if stype < 0 else (Ai <= col)
        i, j, a = Ai[keep], col[keep], Ax[keep]
        x = np.asarray(x, dtype=np.complex128).ravel()
        b = np.asarray(b, dtype=np.complex128).ravel()
        r = -b.copy()
        rowsum = np.zeros(n)
        np.add.at(r, i, a * x[j])
        np.add.at(rowsum, i, np.abs(a))
        off = i != j
        np.add.at(r, j[off], np.conj(a[off]) * x[i[off]])
        np.add.at(rowsum, j[off], np.abs(a[off]))
        return float(np.abs(r).max() / (rowsum.max() * np.abs(x).max() + np.abs(b).max()))


_port = None


def _portlib():
    global _port
    if _port is None:
        lib = C.CDLL(os.path.join(_REF, "libsnport.so"))
        for f in (lib.sn_port_d, lib.sn_port_s, lib.sn_port_z, lib.sn_port_c):
            f.restype = C.c_int
            f.argtypes = [C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                          C.c_void_p, C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        lib.sn_port_solve_d.argtypes = [C.c_int64] + [C.c_void_p] * 6
        lib.sn_port_solve_z.argtypes = [C.c_int64] + [C.c_void_p] * 6
        _port = lib
    return _port


def port_factor(sym, Sp, Si, Sx, beta=0.0, quick_return=False, Snz=None, F=None):
    """Run the plain-C restatement.  Returns (status, minor, Lx).  F = (Fp, Fi, Fx) selects the unsymmetric
    case: S is then A(p,:) and S*F is factorized."""
    lib = _portlib()
    xd = np.asarray(Sx).dtype
    if xd not in (np.float32, np.float64, np.complex64, np.complex128):
        xd = np.dtype(np.float64)
    Sp = np.ascontiguousarray(Sp, dtype=np.int64)
    Si = np.ascontiguousarray(Si, dtype=np.int64)
    Sx = np.ascontiguousarray(Sx, dtype=xd)
    Lx = np.full(sym["xsize"], np.nan, dtype=xd)
    minor = C.c_int64(-1)
    sup, pi, px, s = (np.ascontiguousarray(sym[k], dtype=np.int64) for k in ("super", "pi", "px", "s"))
    fn = {np.dtype(np.float64): lib.sn_port_d, np.dtype(np.float32): lib.sn_port_s,
          np.dtype(np.complex128): lib.sn_port_z, np.dtype(np.complex64): lib.sn_port_c}[np.dtype(xd)]
    nzp = None if Snz is None else np.ascontiguousarray(Snz, dtype=np.int64).ctypes.data
    fp = fi = fx = None
    if F is not None:
        Fp, Fi = (np.ascontiguousarray(a, dtype=np.int64) for a in F[:2])
        Fx = np.ascontiguousarray(F[2], dtype=xd)
        fp, fi, fx = Fp.ctypes.data, Fi.ctypes.data, Fx.ctypes.data
    st = fn(sym["n"], sym["nsuper"], sup.ctypes.data, pi.ctypes.data, px.ctypes.data, s.ctypes.data,
            sym["xsize"], Sp.ctypes.data, Si.ctypes.data, nzp, Sx.ctypes.data, fp, fi, None, fx, float(beta),
            int(quick_return), Lx.ctypes.data, C.byref(minor))
    return int(st), int(minor.value), Lx


def port_solve(sym, Lx, b_perm):
    """Solve L L' x = b in the permuted ordering with the plain-C restatement."""
    lib = _portlib()
    sup, pi, px, s = (np.ascontiguousarray(sym[k], dtype=np.int64) for k in ("super", "pi", "px", "s"))
    cplx = np.iscomplexobj(Lx)
    Lx = np.ascontiguousarray(Lx, dtype=np.complex128 if cplx else np.float64)
    x = np.array(b_perm, dtype=np.complex128 if cplx else np.float64, copy=True)
    (lib.sn_port_solve_z if cplx else lib.sn_port_solve_d)(sym["nsuper"], sup.ctypes.data, pi.ctypes.data, px.ctypes.data, s.ctypes.data,
                        Lx.ctypes.data, x.ctypes.data)
    return x
