"""ctypes bindings for libkatib_b200.so (include/katib_b200.h).  No compute happens in Python."""
import ctypes as C
import os
import re
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
_SO = os.path.join(_HERE, "libkatib_b200.so")
_HEADER = os.path.join(ROOT, "include", "katib_b200.h")


class KbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"katib_b200 error {code}: {msg}")
        self.code = code


def lib_path():
    return _SO


def build(force=False):
    """Compile the CUDA library in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-C", src, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", src, "-j8"], stdout=subprocess.DEVNULL)
    return _SO


def header_symbols():
    """Every function the public header declares."""
    text = open(_HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kb_[a-z0-9_]+)\s*\(", text)))


_dp = C.POINTER(C.c_double)
_LIB = None


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(_SO):
        raise ImportError(f"{_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(_SO)
    vp, i32, i64, u64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double
    sig = {
        "kb_version": ([], i32),
        "kb_last_error": ([vp], C.c_char_p),
        "kb_create": ([i32, C.POINTER(vp)], i32),
        "kb_destroy": ([vp], i32),
        "kb_device_info": ([vp, C.POINTER(i64)], i32),
        "kb_stream": ([vp], vp),
        "kb_comm_unique_id": ([vp], i32),
        "kb_comm_init": ([vp, i32, i32, vp], i32),
        "kb_comm_world": ([vp, C.POINTER(i32), C.POINTER(i32)], i32),
        "kb_space_set": ([vp, i32, vp, vp, vp, vp, vp], i32),
        "kb_tpe_set_history": ([vp, i64, vp, vp], i32),
        "kb_tpe_append_history": ([vp, i64, vp, vp], i32),
        "kb_tpe_set_option": ([vp, i32, i32], i32),
        "kb_tpe_get_parzen": ([vp, i32, i32, C.POINTER(i32), vp, vp, vp, i32, C.POINTER(dbl)], i32),
        "kb_tpe_scores": ([vp, i64, vp, vp], i32),
        "kb_tpe_score_select": ([vp, i64, vp, i64, vp, vp, vp], i32),
        "kb_tpe_score_select_device": ([vp, i64, vp, i64, vp, vp, vp, vp], i32),
        "kb_tpe_suggest": ([vp, i64, u64, vp], i32),
        "kb_tpe_score_error_device": ([vp, i64, vp, vp], i32),
        "kb_tpe_local_stats": ([vp, vp], i32),
        "kb_tpe_sample_candidates": ([vp, i64, u64, vp, vp, vp], i32),
        "kb_sobol_fill": ([vp, i32, u64, i64, vp, vp, vp], i32),
        "kb_sobol_fill_device": ([vp, i32, u64, i64, vp, vp, vp, vp], i32),
        "kb_cmaes_create": ([vp, i32, vp, dbl, i32, vp, vp], i32),
        "kb_cmaes_ask": ([vp, i32, vp, u64, vp], i32),
        "kb_cmaes_tell": ([vp, vp, vp], i32),
        "kb_cmaes_get_state": ([vp, vp, C.POINTER(dbl), vp, vp, vp, vp, vp, C.POINTER(i32)], i32),
        "kb_cmaes_set_eig": ([vp, vp, vp], i32),
        "kb_eigh": ([vp, i32, vp, vp, vp], i32),
        "kb_random_fill": ([vp, u64, u64, i64, vp], i32),
        "kb_eigh_stats": ([vp, C.POINTER(i32)], i32),
        "kb_peak_measure": ([vp, i32, C.POINTER(dbl)], i32),
        "kb_launch_count": ([vp], i64),
        "kb_tpe_scores_noalloc": ([i32, i64, vp, i32, vp, vp, vp, i32, vp, vp, vp, dbl, dbl, vp, vp], i32),
        "kb_service_create": ([i32, C.POINTER(vp)], i32),
        "kb_service_create_multi": ([i32, C.POINTER(i32), C.POINTER(vp)], i32),
        "kb_service_destroy": ([vp], i32),
        "kb_service_get_suggestions": ([vp, vp, C.c_size_t, C.POINTER(vp), C.POINTER(C.c_size_t)], i32),
        "kb_service_validate_algorithm_settings": ([vp, vp, C.c_size_t], i32),
        "kb_service_last_message": ([vp], C.c_char_p),
        "kb_service_stat": ([vp, i32], C.c_longlong),
        "kb_service_set_option": ([vp, i32, i32], i32),
        "kb_profile": ([vp, i32], i32),
        "kb_profile_read": ([vp, i32, C.POINTER(dbl), C.POINTER(i64)], i32),
    }
    for name, (args, res) in sig.items():
        f = getattr(L, name)
        f.argtypes, f.restype = args, res
    _LIB = L
    return L


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a):
    return None if a is None else a.ctypes.data


def tpe_scores_noalloc(device, n_cand, cand_ptr, below, above, low, high, out_ptr, stream=None):
    """kb_tpe_scores_noalloc on raw device pointers; below/above = (K, w_ptr, mu_ptr, sigma_ptr)."""
    rc = lib().kb_tpe_scores_noalloc(device, n_cand, cand_ptr, below[0], below[1], below[2], below[3], above[0], above[1],
                                     above[2], above[3], low, high, out_ptr, stream)
    if rc:
        raise KbError(rc, lib().kb_last_error(None).decode())


class Service:
    """kb_service: the suggestion service above the kernels.  Requests / replies are serialized protobuf bytes."""
    OK, INVALID_ARGUMENT, INTERNAL = 0, 3, 13

    def __init__(self, device=0, devices=None):
        """devices: a list of CUDA device indices -> kb_service_create_multi (one service over several GPUs)"""
        self._h = C.c_void_p()
        if devices is not None and len(devices) > 1:
            arr = (C.c_int * len(devices))(*devices)
            rc = lib().kb_service_create_multi(len(devices), arr, C.byref(self._h))
        else:
            rc = lib().kb_service_create(devices[0] if devices else device, C.byref(self._h))
        if rc:
            raise KbError(rc, "kb_service_create failed: " + lib().kb_last_error(None).decode())

    def close(self):
        if getattr(self, "_h", None):
            lib().kb_service_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def message(self):
        return lib().kb_service_last_message(self._h).decode()

    def get_suggestions(self, request_bytes):
        """-> (grpc status code, reply bytes, status message)"""
        reply, n = C.c_void_p(), C.c_size_t()
        code = lib().kb_service_get_suggestions(self._h, request_bytes, len(request_bytes), C.byref(reply), C.byref(n))
        data = C.string_at(reply, n.value) if n.value else b""
        return code, data, self.message()

    def stat(self, which):
        return int(lib().kb_service_stat(self._h, which))

    def set_option(self, option, value):
        rc = lib().kb_service_set_option(self._h, option, value)
        if rc:
            raise KbError(rc, "kb_service_set_option: unknown option")

    def validate(self, request_bytes):
        """-> (grpc status code, status message); request_bytes None = nil request"""
        code = lib().kb_service_validate_algorithm_settings(self._h, request_bytes, len(request_bytes) if request_bytes else 0)
        return code, self.message()


class Context:
    """One kb_ctx (one GPU).  Thin: every method is one C-ABI call on host numpy buffers or raw device pointers."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().kb_create(device, C.byref(self._h))
        if rc:
            raise KbError(rc, lib().kb_last_error(None).decode())
        self.D = 0
        self._cma_n = 0

    def close(self):
        if getattr(self, "_h", None):
            lib().kb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise KbError(rc, lib().kb_last_error(self._h).decode())

    # -- info
    def device_info(self):
        out = (C.c_int64 * 5)()
        self._ck(lib().kb_device_info(self._h, out))
        return dict(sm_count=out[0], cc=(out[1], out[2]), sm_clock_khz=out[3], l2_bytes=out[4])

    @property
    def stream(self):
        return lib().kb_stream(self._h)

    def launch_count(self):
        return int(lib().kb_launch_count(self._h))

    def peak(self, which):
        v = C.c_double()
        self._ck(lib().kb_peak_measure(self._h, which, C.byref(v)))
        return v.value

    def profile(self, enable):
        self._ck(lib().kb_profile(self._h, 1 if enable else 0))

    def profile_read(self, kernel_class):
        ms, n = C.c_double(), C.c_int64()
        self._ck(lib().kb_profile_read(self._h, kernel_class, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # -- comm
    @staticmethod
    def unique_id():
        buf = (C.c_char * 128)()
        rc = lib().kb_comm_unique_id(buf)
        if rc:
            raise KbError(rc, lib().kb_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, world, rank, uid):
        buf = (C.c_char * 128).from_buffer_copy(uid)
        self._ck(lib().kb_comm_init(self._h, world, rank, buf))

    # -- space / history
    def space_set(self, kind, low, high, step=None, n_choices=None):
        D = len(kind)
        kind = np.ascontiguousarray(kind, dtype=np.int32)
        low, high = _f64(low), _f64(high)
        step = _f64(step if step is not None else np.zeros(D))
        nch = np.ascontiguousarray(n_choices if n_choices is not None else np.zeros(D), dtype=np.int32)
        self._ck(lib().kb_space_set(self._h, D, _ptr(kind), _ptr(low), _ptr(high), _ptr(step), _ptr(nch)))
        self.D = D

    def tpe_set_option(self, option, value):
        self._ck(lib().kb_tpe_set_option(self._h, option, value))

    def tpe_set_history(self, values, loss):
        values, loss = _f64(values), _f64(loss)
        n = len(loss)
        self._ck(lib().kb_tpe_set_history(self._h, n, _ptr(values), _ptr(loss)))

    def tpe_append_history(self, values, loss):
        values, loss = _f64(values), _f64(loss)
        self._ck(lib().kb_tpe_append_history(self._h, len(loss), _ptr(values), _ptr(loss)))

    def tpe_get_parzen(self, param, side, cap=None):
        K = C.c_int()
        pa = C.c_double()
        lib().kb_tpe_get_parzen(self._h, param, side, C.byref(K), None, None, None, 0, None)
        n = K.value if cap is None else cap
        w, mu, sg = np.zeros(n), np.zeros(n), np.zeros(n)
        self._ck(lib().kb_tpe_get_parzen(self._h, param, side, C.byref(K), _ptr(w), _ptr(mu), _ptr(sg), n, C.byref(pa)))
        return w[:K.value], mu[:K.value], sg[:K.value], pa.value

    def tpe_scores(self, cand):
        cand = _f64(cand)
        n = cand.shape[0]
        out = np.zeros((n, self.D))
        self._ck(lib().kb_tpe_scores(self._h, n, _ptr(cand), _ptr(out)))
        return out

    def tpe_score_select(self, cand, index_offset=0):
        cand = _f64(cand)
        n = cand.shape[0]
        idx, val, sc = np.zeros(self.D, dtype=np.int64), np.zeros(self.D), np.zeros(self.D)
        self._ck(lib().kb_tpe_score_select(self._h, n, _ptr(cand) if n else None, index_offset, _ptr(idx), _ptr(val),
                                           _ptr(sc)))
        return idx, val, sc

    def tpe_score_select_ptr(self, n, host_ptr, index_offset=0):
        """Same, from a raw (pinned) host pointer -- used by bench.py's end-to-end leg."""
        idx, val, sc = np.zeros(self.D, dtype=np.int64), np.zeros(self.D), np.zeros(self.D)
        self._ck(lib().kb_tpe_score_select(self._h, n, host_ptr, index_offset, _ptr(idx), _ptr(val), _ptr(sc)))
        return idx, val, sc

    def tpe_score_select_device(self, n, dev_ptr, index_offset=0, stream=None):
        idx, val, sc = np.zeros(self.D, dtype=np.int64), np.zeros(self.D), np.zeros(self.D)
        self._ck(lib().kb_tpe_score_select_device(self._h, n, dev_ptr, index_offset, _ptr(idx), _ptr(val), _ptr(sc),
                                                  stream))
        return idx, val, sc

    def tpe_score_error_device(self, n, dev_ptr):
        """per-parameter worst |FP32-pipeline score - FP64 score| / max(1, |FP64 score|) over n candidates in HBM"""
        out = np.zeros(self.D)
        self._ck(lib().kb_tpe_score_error_device(self._h, n, dev_ptr, _ptr(out)))
        return out

    def tpe_local_stats(self):
        """(blocks that used their own origin, blocks that kept the (hi, lo) form) of the KB_OPT_LOCAL_ORIGIN kernel"""
        out = np.zeros(2, dtype=np.uint64)
        self._ck(lib().kb_tpe_local_stats(self._h, _ptr(out)))
        return int(out[0]), int(out[1])

    def tpe_suggest(self, n_ei_candidates, seed):
        out = np.zeros(self.D)
        self._ck(lib().kb_tpe_suggest(self._h, n_ei_candidates, seed, _ptr(out)))
        return out

    def tpe_sample_candidates(self, n, seed):
        out = np.zeros((n, self.D))
        self._ck(lib().kb_tpe_sample_candidates(self._h, n, seed, _ptr(out), None, None))
        return out

    # -- sobol
    def sobol_fill(self, dim, first, n, low=None, high=None):
        out = np.zeros((n, dim))
        lo = _f64(low) if low is not None else None
        hi = _f64(high) if high is not None else None
        self._ck(lib().kb_sobol_fill(self._h, dim, first, n, _ptr(lo), _ptr(hi), _ptr(out) if n else None))
        return out

    def sobol_fill_device(self, dim, first, n, dev_ptr, low=None, high=None, stream=None):
        lo = _f64(low) if low is not None else None
        hi = _f64(high) if high is not None else None
        self._ck(lib().kb_sobol_fill_device(self._h, dim, first, n, _ptr(lo), _ptr(hi), dev_ptr, stream))

    # -- cma-es
    def cmaes_create(self, mean, sigma, popsize=0, lo=None, hi=None):
        mean = _f64(mean)
        lo_a = _f64(lo) if lo is not None else None
        hi_a = _f64(hi) if hi is not None else None
        self._cma_n = len(mean)
        self._ck(lib().kb_cmaes_create(self._h, len(mean), _ptr(mean), sigma, popsize, _ptr(lo_a), _ptr(hi_a)))

    def cmaes_ask(self, lam, z=None, seed=0, out=None):
        """out: optional C-contiguous float64 [lam, n] buffer to receive x (e.g. a view of pinned memory)"""
        x = out if out is not None else np.zeros((lam, self._cma_n))
        assert x.dtype == np.float64 and x.flags.c_contiguous and x.shape == (lam, self._cma_n)
        zz = _f64(z) if z is not None else None
        self._ck(lib().kb_cmaes_ask(self._h, lam, _ptr(zz), seed, _ptr(x)))
        return x

    def cmaes_tell(self, x, f):
        x, f = _f64(x), _f64(f)
        self._ck(lib().kb_cmaes_tell(self._h, _ptr(x), _ptr(f)))

    def cmaes_state(self, want_eig=False):
        n = self._cma_n
        m, Cm, ps, pc = np.zeros(n), np.zeros((n, n)), np.zeros(n), np.zeros(n)
        B, D = (np.zeros((n, n)), np.zeros(n)) if want_eig else (None, None)
        s, g = C.c_double(), C.c_int()
        self._ck(lib().kb_cmaes_get_state(self._h, _ptr(m), C.byref(s), _ptr(Cm), _ptr(ps), _ptr(pc), _ptr(B), _ptr(D),
                                          C.byref(g)))
        return dict(mean=m, sigma=s.value, C=Cm, p_sigma=ps, p_c=pc, B=B, D=D, generation=g.value)

    def cmaes_set_eig(self, B, D):
        B, D = _f64(B), _f64(D)
        self._ck(lib().kb_cmaes_set_eig(self._h, _ptr(B), _ptr(D)))

    def eigh(self, A):
        A = _f64(A)
        n = A.shape[0]
        w, V = np.zeros(n), np.zeros((n, n))
        self._ck(lib().kb_eigh(self._h, n, _ptr(A), _ptr(w), _ptr(V)))
        return w, V

    def eigh_sweeps(self):
        v = C.c_int()
        self._ck(lib().kb_eigh_stats(self._h, C.byref(v)))
        return v.value

    def random_fill(self, seed, first, n):
        out = np.zeros((n, self.D))
        self._ck(lib().kb_random_fill(self._h, seed, first, n, _ptr(out)))
        return out
