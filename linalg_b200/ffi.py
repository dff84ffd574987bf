"""ctypes binding of the CBLAS + Fortran-LAPACK C ABI the reference links against.

This is the Python counterpart of the reference's two `lib` blocks
(`lib LibCBLAS`, /root/reference/src/linalg/libCBLAS.cr:8-222, and `lib LibLAPACK`,
src/linalg/libLAPACK_linux.cr:5-2225) restricted to the hot path: it binds the same
symbol names with the same argument conventions (CBLAS: scalars by value, complex by
pointer; LAPACK: every argument by pointer, `char*` flags without hidden lengths,
`info` last).  A `BlasLapackABI` can be pointed at ANY shared library exporting that
ABI, optionally with a symbol prefix; the product binds it to libla_b200.so
(`linalg_b200.lib()`), the tests also bind it to the CPU oracle to run the same host
code against both.
"""
from __future__ import annotations

import ctypes as C
from ctypes import POINTER, byref, c_char, c_char_p, c_double, c_float, c_int, c_int64, c_void_p

import numpy as np

# CBLAS enum values, libCBLAS.cr:19-48
ROW_MAJOR, COL_MAJOR = 101, 102
NO_TRANS, TRANS, CONJ_TRANS = 111, 112, 113
UPPER, LOWER = 121, 122
NON_UNIT, UNIT = 131, 132
LEFT, RIGHT = 141, 142

LETTER = {np.dtype(np.float32): "s", np.dtype(np.float64): "d", np.dtype(np.complex128): "z"}
_REAL = {"s": c_float, "d": c_double}

_P = c_void_p
_IP = POINTER(c_int)


def _scalar_by_value(letter):
    return _REAL[letter] if letter in _REAL else _P


class BlasLapackABI:
    """Typed access to `<prefix>cblas_?gemm`, `<prefix>?getrf_`, ... of one library."""

    def __init__(self, cdll: C.CDLL, prefix: str = "", name: str = "?"):
        self.cdll = cdll
        self.prefix = prefix
        self.name = name
        self._cache = {}

    # -- symbol lookup --------------------------------------------------------------
    def has(self, sym: str) -> bool:
        try:
            getattr(self.cdll, self.prefix + sym)
            return True
        except AttributeError:
            return False

    def _fn(self, sym, argtypes, restype=None):
        f = self._cache.get(sym)
        if f is None:
            f = getattr(self.cdll, self.prefix + sym)
            f.argtypes = argtypes
            f.restype = restype
            self._cache[sym] = f
        return f

    # -- CBLAS ----------------------------------------------------------------------
    def gemm(self, letter):
        sc = _scalar_by_value(letter)
        return self._fn(f"cblas_{letter}gemm",
                        [c_int, c_int, c_int, c_int, c_int, c_int, sc, _P, c_int, _P, c_int, sc,
                         _P, c_int])

    def symm(self, letter, herm=False):
        sc = _scalar_by_value(letter)
        nm = "hemm" if herm else "symm"
        return self._fn(f"cblas_{letter}{nm}",
                        [c_int, c_int, c_int, c_int, c_int, sc, _P, c_int, _P, c_int, sc, _P,
                         c_int])

    def trmm(self, letter):
        sc = _scalar_by_value(letter)
        return self._fn(f"cblas_{letter}trmm",
                        [c_int, c_int, c_int, c_int, c_int, c_int, c_int, sc, _P, c_int, _P,
                         c_int])

    def trsm(self, letter):
        sc = _scalar_by_value(letter)
        return self._fn(f"cblas_{letter}trsm",
                        [c_int, c_int, c_int, c_int, c_int, c_int, c_int, sc, _P, c_int, _P,
                         c_int])

    def syrk(self, letter):
        # real: ?syrk; complex: zherk (real alpha/beta by value in both)
        nm = "zherk" if letter == "z" else f"{letter}syrk"
        sc = c_double if letter == "z" else _REAL[letter]
        return self._fn(f"cblas_{nm}",
                        [c_int, c_int, c_int, c_int, c_int, sc, _P, c_int, sc, _P, c_int])

    # -- LAPACK (all-pointer) --------------------------------------------------------
    def lapack(self, routine: str, letter: str):
        """Return the ctypes function `<letter><routine>_` (with or/un renaming for complex,
        as src/linalg/lapack_helper.cr:201-203 does)."""
        if letter == "z" and routine.startswith("or"):
            routine = "un" + routine[2:]
        sigs = {
            "getrf": [_IP, _IP, _P, _IP, _P, _IP],
            "getrs": [c_char_p, _IP, _IP, _P, _IP, _P, _P, _IP, _IP],
            "gesv": [_IP, _IP, _P, _IP, _P, _P, _IP, _IP],
            "getri": [_IP, _P, _IP, _P, _P, _IP, _IP],
            "potrf": [c_char_p, _IP, _P, _IP, _IP],
            "potrs": [c_char_p, _IP, _IP, _P, _IP, _P, _IP, _IP],
            "posv": [c_char_p, _IP, _IP, _P, _IP, _P, _IP, _IP],
            "potri": [c_char_p, _IP, _P, _IP, _IP],
            "trtri": [c_char_p, c_char_p, _IP, _P, _IP, _IP],
            "trtrs": [c_char_p, c_char_p, c_char_p, _IP, _IP, _P, _IP, _P, _IP, _IP],
            "geqrf": [_IP, _IP, _P, _IP, _P, _P, _IP, _IP],
            "ormqr": [c_char_p, c_char_p, _IP, _IP, _IP, _P, _IP, _P, _P, _IP, _P, _IP, _IP],
            "unmqr": [c_char_p, c_char_p, _IP, _IP, _IP, _P, _IP, _P, _P, _IP, _P, _IP, _IP],
            "orgqr": [_IP, _IP, _IP, _P, _IP, _P, _P, _IP, _IP],
            "ungqr": [_IP, _IP, _IP, _P, _IP, _P, _P, _IP, _IP],
            "gels": [c_char_p, _IP, _IP, _IP, _P, _IP, _P, _IP, _P, _IP, _IP],
            "geqp3": [_IP, _IP, _P, _IP, _P, _P, _P, _IP] + ([_P] if letter == "z" else []) + [_IP],
            "lacpy": [c_char_p, _IP, _IP, _P, _IP, _P, _IP],
            "laset": [c_char_p, _IP, _IP, _P, _P, _P, _IP],
        }
        for fam in ("gerqf", "gelqf", "geqlf"):
            sigs[fam] = sigs["geqrf"]
        for fam in ("orgrq", "orglq", "orgql", "ungrq", "unglq", "ungql"):
            sigs[fam] = sigs["orgqr"]
        return self._fn(f"{letter}{routine}_", sigs[routine])

    def norm(self, routine: str, letter: str):
        """?lange / ?lansy / zlanhe / ?lantr: functions returning the norm (float for the REAL
        routines, FloatReturn = Float, libLAPACK_linux.cr:17)."""
        sigs = {"lange": [c_char_p, _IP, _IP, _P, _IP, _P],
                "lansy": [c_char_p, c_char_p, _IP, _P, _IP, _P],
                "lanhe": [c_char_p, c_char_p, _IP, _P, _IP, _P],
                "lantr": [c_char_p, c_char_p, c_char_p, _IP, _IP, _P, _IP, _P]}
        return self._fn(f"{letter}{routine}_", sigs[routine], c_float if letter == "s" else c_double)

    def ilaver(self):
        f = self._fn("ilaver_", [_IP, _IP, _IP])
        a, b, c = c_int(0), c_int(0), c_int(0)
        f(byref(a), byref(b), byref(c))
        return a.value, b.value, c.value


def ptr(a: np.ndarray):
    return a.ctypes.data_as(c_void_p)


def scalar_arg(letter, value):
    """CBLAS scalar: by value for s/d, pointer to {re, im} for z (blas_const, mult.cr:17-23).
    Returns (argument, keepalive)."""
    if letter == "z":
        buf = np.array([complex(value)], dtype=np.complex128)
        return ptr(buf), buf
    return (_REAL[letter](float(value)), None)


# every symbol include/la_b200.h declares; tests check the library exports all of them
def declared_symbols():
    syms = []
    for l in "sdz":
        syms += [f"cblas_{l}gemm", f"cblas_{l}symm", f"cblas_{l}trmm", f"cblas_{l}trsm"]
        for r in ("getrf", "getrs", "gesv", "potrf", "potrs", "posv", "geqrf", "gels", "trtrs",
                  "getri", "potri", "trtri", "gerqf", "gelqf", "geqlf", "geqp3", "lange", "lansy",
                  "lantr", "lacpy", "laset"):
            syms.append(f"{l}{r}_")
        for r in ("trsm", "herk", "getrf", "getrs", "potrf", "potrs", "geqrf", "ormqr", "gels",
                  "trtrs", "laswp", "gemm_stair", "lange", "lacpy", "laset", "geadd", "geabs"):
            syms.append(f"la_b200_{l}{r}_dev")
        syms.append(f"la_b200_{l}gemm_dev")
    syms += ["cblas_zhemm", "cblas_ssyrk", "cblas_dsyrk", "cblas_zherk", "sormqr_", "dormqr_",
             "zunmqr_", "sorgqr_", "dorgqr_", "zungqr_", "ilaver_", "openblas_set_num_threads",
             "openblas_get_num_threads", "openblas_get_num_procs", "openblas_get_config",
             "openblas_get_corename", "openblas_get_parallel", "la_b200_version",
             "la_b200_workspace_bytes", "la_b200_last_gemm_path", "la_b200_release_workspace",
             "la_b200_set_sm_limit", "la_b200_dgetrf_dev_noalloc", "zlanhe_", "sorgrq_", "dorgrq_",
             "zungrq_", "sorglq_", "dorglq_", "zunglq_", "sorgql_", "dorgql_", "zungql_",
             "la_b200_mat_create", "la_b200_mat_free", "la_b200_mat_ptr", "la_b200_mat_ld",
             "la_b200_mat_rows", "la_b200_mat_cols", "la_b200_mat_upload", "la_b200_mat_download"]
    return syms
