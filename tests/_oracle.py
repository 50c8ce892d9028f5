"""ctypes view of oracle/libkatib_oracle.so -- the CPU checker.  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "libkatib_oracle.so")


def build():
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("tpe.c", "sobol.c", "cmaes.c", "katib_oracle.h")]
    if not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return _SO


_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_up = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


class ParzenOpts(C.Structure):
    _fields_ = [("consider_prior", C.c_int), ("prior_weight", C.c_double), ("consider_magic_clip", C.c_int),
                ("consider_endpoints", C.c_int)]


def load():
    lib = C.CDLL(build())
    lib.kbo_parzen_defaults.restype = ParzenOpts
    lib.kbo_tpe_gamma.argtypes = [C.c_int]
    lib.kbo_tpe_weights.argtypes = [C.c_int, _dp]
    lib.kbo_tpe_split.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
    lib.kbo_parzen_build.argtypes = [_dp, C.c_int, C.c_double, C.c_double, ParzenOpts, _dp, _dp, _dp]
    lib.kbo_normal_cdf.argtypes = [C.c_double] * 3
    lib.kbo_normal_cdf.restype = C.c_double
    lib.kbo_gmm_log_pdf.argtypes = [_dp, C.c_int64, C.c_int, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, _dp]
    lib.kbo_tpe_score_numeric.argtypes = [_dp, C.c_int64, C.c_int, _dp, _dp, C.c_double, C.c_double, C.c_double,
                                          C.c_void_p]
    lib.kbo_tpe_score_numeric.restype = C.c_int64
    lib.kbo_tpe_categorical_tables.argtypes = [C.c_int, _dp, _dp, C.c_int, C.c_double, _dp, _dp]
    lib.kbo_tpe_score_categorical.argtypes = [_dp, C.c_int64, C.c_int, _dp, _dp, C.c_int, C.c_void_p]
    lib.kbo_tpe_score_categorical.restype = C.c_int64
    lib.kbo_gmm_sample.argtypes = [C.c_uint64, C.c_int64, C.c_int, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double,
                                   _dp]
    lib.kbo_sobol_fill.argtypes = [C.c_int, C.c_int, _ip, _ip, C.c_uint64, C.c_int64, _dp]
    lib.kbo_sobol_fill_u32.argtypes = [C.c_int, _ip, _ip, C.c_uint64, C.c_int64, _up]
    lib.kbo_cma_create.argtypes = [C.c_int, _dp, C.c_double, C.c_int, C.c_void_p, C.c_void_p]
    lib.kbo_cma_create.restype = C.c_void_p
    lib.kbo_cma_destroy.argtypes = [C.c_void_p]
    for f in ("kbo_cma_popsize", "kbo_cma_mu", "kbo_cma_generation"):
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.kbo_cma_constants.argtypes = [C.c_void_p, _dp, C.c_void_p]
    lib.kbo_cma_eig.argtypes = [C.c_void_p, _dp, _dp]
    lib.kbo_cma_set_eig.argtypes = [C.c_void_p, _dp, _dp]
    lib.kbo_cma_ask_batch.argtypes = [C.c_void_p, C.c_int, _dp, _dp]
    lib.kbo_cma_tell.argtypes = [C.c_void_p, _dp, _dp]
    lib.kbo_cma_state.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_double), _dp, _dp, _dp]
    lib.kbo_jacobi_eigh.argtypes = [C.c_int, _dp, _dp, _dp]
    return lib


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _LIB = load()
    return _LIB


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


# ---- convenience wrappers -------------------------------------------------------------------
def parzen_build(obs, low, high):
    obs = f64(obs)
    n = len(obs)
    w, mu, sg = np.zeros(n + 1), np.zeros(n + 1), np.zeros(n + 1)
    k = lib().kbo_parzen_build(obs if n else np.zeros(1), n, low, high, lib().kbo_parzen_defaults(), w, mu, sg)
    return w[:k], mu[:k], sg[:k]


def gmm_log_pdf(x, w, mu, sg, low, high, q=0.0):
    x = f64(x)
    out = np.zeros(len(x))
    lib().kbo_gmm_log_pdf(x, len(x), len(w), f64(w), f64(mu), f64(sg), low, high, q, out)
    return out


def tpe_split(values, loss):
    values, loss = f64(values), f64(loss)
    n = len(values)
    b, a = np.zeros(max(n, 1)), np.zeros(max(n, 1))
    nb = lib().kbo_tpe_split(n, values, loss, b, a)
    return b[:nb], a[:n - nb]


def tpe_score_numeric(x, values, loss, low, high, q=0.0):
    x, values, loss = f64(x), f64(values), f64(loss)
    sc = np.zeros(len(x))
    best = lib().kbo_tpe_score_numeric(x, len(x), len(values), values, loss, low, high, q, sc.ctypes.data)
    return sc, int(best)


def tpe_score_categorical(x, values, loss, n_choices):
    x, values, loss = f64(x), f64(values), f64(loss)
    sc = np.zeros(len(x))
    best = lib().kbo_tpe_score_categorical(x, len(x), len(values), values, loss, n_choices, sc.ctypes.data)
    return sc, int(best)


def categorical_tables(values, loss, n_choices):
    pl, pg = np.zeros(n_choices), np.zeros(n_choices)
    lib().kbo_tpe_categorical_tables(len(values), f64(values), f64(loss), n_choices, 1.0, pl, pg)
    return pl, pg


def joe_kuo(golden_dir):
    d = np.load(os.path.join(golden_dir, "joe_kuo_d1024.npz"))
    return np.ascontiguousarray(d["poly"]), np.ascontiguousarray(d["vinit"])


def sobol(golden_dir, dim, first, n, bits=32):
    poly, vinit = joe_kuo(golden_dir)
    out = np.zeros((n, dim))
    rc = lib().kbo_sobol_fill(dim, bits, poly, vinit, first, n, out)
    assert rc == 0
    return out


def sobol_u32(golden_dir, dim, first, n):
    poly, vinit = joe_kuo(golden_dir)
    out = np.zeros((n, dim), dtype=np.uint32)
    assert lib().kbo_sobol_fill_u32(dim, poly, vinit, first, n, out) == 0
    return out


class Cma:
    def __init__(self, mean, sigma, popsize=0):
        self.n = len(mean)
        self.h = lib().kbo_cma_create(self.n, f64(mean), sigma, popsize, None, None)
        self.popsize = lib().kbo_cma_popsize(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().kbo_cma_destroy(self.h)
            self.h = None

    def constants(self):
        out, w = np.zeros(9), np.zeros(self.popsize)
        lib().kbo_cma_constants(self.h, out, w.ctypes.data)
        return out, w

    def eig(self):
        B, D = np.zeros((self.n, self.n)), np.zeros(self.n)
        lib().kbo_cma_eig(self.h, B, D)
        return B, D

    def set_eig(self, B, D):
        lib().kbo_cma_set_eig(self.h, f64(B), f64(D))

    def ask_batch(self, z):
        z = f64(z)
        x = np.zeros_like(z)
        lib().kbo_cma_ask_batch(self.h, z.shape[0], z, x)
        return x

    def tell(self, x, f):
        lib().kbo_cma_tell(self.h, f64(x), f64(f))

    def state(self):
        n = self.n
        m, Cm, ps, pc = np.zeros(n), np.zeros((n, n)), np.zeros(n), np.zeros(n)
        s = C.c_double()
        lib().kbo_cma_state(self.h, m, C.byref(s), Cm, ps, pc)
        return m, s.value, Cm, ps, pc

    def state_vec(self):
        m, s, Cm, ps, pc = self.state()
        return np.concatenate([m, [s], Cm.ravel(), ps, pc])


def jacobi_eigh(A):
    A = f64(A)
    n = A.shape[0]
    w, V = np.zeros(n), np.zeros((n, n))
    lib().kbo_jacobi_eigh(n, A, w, V)
    return w, V
