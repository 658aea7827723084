"""ctypes front end of oracle/vq_oracle.c.  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libvq_oracle.so")
_lib = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_up = C.POINTER(C.c_uint16)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "vq_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_distortion.restype = C.c_double
        L.orc_distortion.argtypes = [_dp, _dp, C.c_int]
        L.orc_assign.argtypes = [_dp, C.c_long, C.c_int, _dp, C.c_int, _ip, _dp, _dp]
        L.orc_quantize.restype = C.c_double
        L.orc_quantize.argtypes = [_dp, C.c_int, C.c_int, _dp, C.c_long, _up]
        L.orc_lpca_r.restype = C.c_int
        L.orc_lpca_r.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
        L.orc_lpca_rc.argtypes = [C.c_int, _dp, _dp]
        L.orc_refl_to_raas.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.orc_initial.restype = C.c_int
        L.orc_initial.argtypes = [C.c_int, _dp, _dp]
        L.orc_grow.restype = C.c_int
        L.orc_grow.argtypes = [C.c_int, C.c_int, _dp, _dp]
        L.orc_accumulate.argtypes = [_dp, C.c_long, C.c_int, _ip, _dp, C.c_int, _dp, _ip, _dp, _dp]
        L.orc_centroids.argtypes = [C.c_int, C.c_int, _dp, _ip, _dp, _dp]
        L.orc_sigma.restype = C.c_double
        L.orc_sigma.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_double]
        L.orc_inertia.restype = C.c_double
        L.orc_inertia.argtypes = [_dp, C.c_long, _dp, C.c_int, C.c_int]
        L.orc_learn.restype = C.c_int
        L.orc_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def distortion(rx, ra) -> float:
    rx, ra = _f64(rx), _f64(ra)
    return lib().orc_distortion(_d(rx), _d(ra), len(rx) - 1)


def assign(x, raas):
    """-> (idx int32[N], ddmin f64[N], second f64[N])"""
    x, raas = _f64(x), _f64(raas)
    N, D = x.shape
    K = raas.shape[0]
    idx = np.empty(N, np.int32)
    dd = np.empty(N, np.float64)
    d2 = np.empty(N, np.float64)
    lib().orc_assign(_d(x), N, D - 1, _d(raas), K, _i(idx), _d(dd), _d(d2))
    return idx, dd, d2


def quantize(raas, x):
    """-> (seq uint16[T], average distortion)"""
    x, raas = _f64(x), _f64(raas)
    T, D = x.shape
    seq = np.empty(T, np.uint16)
    avg = lib().orc_quantize(_d(raas), raas.shape[0], D - 1, _d(x), T, seq.ctypes.data_as(_up))
    return seq, avg


def lpca_r(r):
    r = _f64(r)
    P = len(r) - 1
    rc = np.zeros(P + 1)
    a = np.zeros(P + 1)
    pe = C.c_double(0.0)
    code = lib().orc_lpca_r(P, _d(r), _d(rc), _d(a), C.byref(pe))
    return code, rc, a, pe.value


def refl_to_raas(refl):
    refl = _f64(refl)
    K, D = refl.shape
    raas = np.empty_like(refl)
    lib().orc_refl_to_raas(_d(refl), _d(raas), K, D - 1)
    return raas


def initial(P):
    refl = np.zeros((2, P + 1))
    raas = np.zeros((2, P + 1))
    lib().orc_initial(P, _d(refl), _d(raas))
    return refl, raas


def grow(refl):
    refl = _f64(refl)
    K, D = refl.shape
    r2 = np.zeros((2 * K, D))
    r2[:K] = refl
    raas = np.zeros((2 * K, D))
    lib().orc_grow(D - 1, K, _d(r2), _d(raas))
    return r2, raas


def accumulate(x, idx, ddmin, K):
    x = _f64(x)
    idx = np.ascontiguousarray(idx, np.int32)
    ddmin = _f64(ddmin)
    N, D = x.shape
    cells = np.zeros((K, D))
    cardd = np.zeros(K, np.int32)
    discel = np.zeros(K)
    DD = C.c_double(0.0)
    lib().orc_accumulate(_d(x), N, D - 1, _i(idx), _d(ddmin), K, _d(cells), _i(cardd), _d(discel), C.byref(DD))
    return cells, cardd, discel, DD.value


def centroids(cells, cardd, refl, raas):
    """-> (normalised cells, new refl, new raas); inputs are not modified."""
    cells, refl, raas = _f64(cells).copy(), _f64(refl).copy(), _f64(raas).copy()
    cardd = np.ascontiguousarray(cardd, np.int32)
    K, D = refl.shape
    lib().orc_centroids(D - 1, K, _d(cells), _i(cardd), _d(refl), _d(raas))
    return cells, refl, raas


def kmeans_pass(x, refl, raas):
    """One teacher-forced k-means pass from (refl, raas). Returns a dict."""
    idx, dd, d2 = assign(x, raas)
    K = raas.shape[0]
    cells, cardd, discel, DD = accumulate(x, idx, dd, K)
    ncells, nrefl, nraas = centroids(cells, cardd, refl, raas)
    return dict(idx=idx, ddmin=dd, second=d2, cells=cells, cardd=cardd, discel=discel, DD=DD,
                cells_norm=ncells, refl=nrefl, raas=nraas)


def sigma(raas, cells_norm, avg_distortion) -> float:
    raas, cells_norm = _f64(raas), _f64(cells_norm)
    K, D = raas.shape
    return lib().orc_sigma(_d(raas), _d(cells_norm), K, D - 1, avg_distortion)


def inertia(x, raas) -> float:
    x, raas = _f64(x), _f64(raas)
    N, D = x.shape
    return lib().orc_inertia(_d(x), N, _d(raas), raas.shape[0], D - 1)


_STAGE_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                        _dp, _dp, _ip, _dp, _ip, _dp)
_PASS_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_double)


def learn(x, eps: float = 0.05, max_M: int = 4096, base_refl=None, keep_assignments: bool = False):
    """Full LBG run.  Returns (stages, passes): a list of per-stage dicts and the
    per-pass trace [(M, pass, DD), ...]."""
    x = _f64(x)
    N, D = x.shape
    stages, trace = [], []

    def on_stage(_ctx, M, passes, avg, sig, inert, refl, raas, cardd, discel, idx, mind):
        st = dict(M=M, passes=passes, avg_distortion=avg, sigma=sig, inertia=inert,
                  refl=np.ctypeslib.as_array(refl, (M, D)).copy(),
                  raas=np.ctypeslib.as_array(raas, (M, D)).copy(),
                  cardd=np.ctypeslib.as_array(cardd, (M,)).copy(),
                  discel=np.ctypeslib.as_array(discel, (M,)).copy())
        if keep_assignments:
            st["idx"] = np.ctypeslib.as_array(idx, (N,)).copy()
            st["mindist"] = np.ctypeslib.as_array(mind, (N,)).copy()
        stages.append(st)

    def on_pass(_ctx, M, p, DD):
        trace.append((M, p, DD))

    scb, pcb = _STAGE_CB(on_stage), _PASS_CB(on_pass)
    L = lib()
    L.orc_learn.argtypes = [_dp, C.c_long, C.c_int, C.c_double, C.c_int, _dp, C.c_int,
                            _STAGE_CB, _PASS_CB, C.c_void_p]
    if base_refl is not None:
        base_refl = _f64(base_refl)
        rc = L.orc_learn(_d(x), N, D - 1, eps, max_M, _d(base_refl), base_refl.shape[0], scb, pcb, None)
    else:
        rc = L.orc_learn(_d(x), N, D - 1, eps, max_M, None, 0, scb, pcb, None)
    if rc:
        raise RuntimeError(f"orc_learn failed: {rc}")
    return stages, trace


def num_threads() -> int:
    return lib().orc_num_threads()
