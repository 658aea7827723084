"""Host-side mirror of the reference's LPC front end over the C ABI (include/ecoz2_lpc_b200.h).

`lpa_on_signal` mirrors reference ecoz2/src/lpc/lpa_on_signal.c:81-157 (signal ->
gain-normalised autocorrelation vectors, the content of a `.prd`), `lpca` mirrors
ecoz2/src/lpc/lpca.c:9-59 and `lpc_signals` the file-level ecoz2_lpc_signals
(ecoz2/src/lpc/lpc_signals.c:38-153).  All compute runs in the CUDA library; without a GPU
every call fails loudly.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import vq

# every symbol include/ecoz2_lpc_b200.h declares (tests check the library exports them all)
EXPORTED_SYMBOLS = (
    "ecoz2_lpc_signals", "ecoz2_lpca",
    "ecoz2b200_lpa_num_frames", "ecoz2b200_lpa_on_signal", "ecoz2b200_lpa_on_signal_device",
    "ecoz2b200_lpc_last_error", "ecoz2b200_dfma_peak_tflops",
)

_dp = C.POINTER(C.c_double)
_configured = False


class LpcError(RuntimeError):
    pass


def _lib():
    global _configured
    L = vq.lib()
    if not _configured:
        L.ecoz2b200_lpc_last_error.restype = C.c_char_p
        L.ecoz2b200_lpa_num_frames.restype = C.c_long
        L.ecoz2b200_lpa_num_frames.argtypes = [C.c_long, C.c_double, C.c_int, C.c_int]
        L.ecoz2b200_lpa_on_signal.restype = C.c_long
        L.ecoz2b200_lpa_on_signal.argtypes = [_dp, C.c_long, C.c_double, C.c_int, C.c_int, C.c_int, _dp,
                                              C.c_long, C.POINTER(C.c_int)]
        L.ecoz2b200_lpa_on_signal_device.restype = C.c_long
        L.ecoz2b200_lpa_on_signal_device.argtypes = [C.c_void_p, C.c_long, C.c_double, C.c_int, C.c_int,
                                                     C.c_int, C.c_void_p, C.c_long, C.POINTER(C.c_int),
                                                     C.c_void_p]
        L.ecoz2b200_dfma_peak_tflops.restype = C.c_double
        L.ecoz2b200_dfma_peak_tflops.argtypes = [C.c_int, C.c_double]
        L.ecoz2_lpca.restype = C.c_int
        L.ecoz2_lpca.argtypes = [_dp, C.c_int, C.c_int, _dp, _dp, _dp, _dp]
        L.ecoz2_lpc_signals.restype = C.c_int
        L.ecoz2_lpc_signals.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                        C.POINTER(C.c_char_p), C.c_int, C.c_float, C.c_int]
        _configured = True
    return L


def _err() -> str:
    return (_lib().ecoz2b200_lpc_last_error() or b"").decode()


def dfma_peak_tflops(device: int = 0, millis: float = 20.0) -> float:
    """Measured FP64 FMA peak of `device` (TFLOP/s): denominator of lpc_autocorr's roofline."""
    return float(_lib().ecoz2b200_dfma_peak_tflops(device, millis))


def num_frames(num_samples: int, sample_rate: float, window_ms: int, offset_ms: int) -> int:
    """Frames lpa_on_signal makes of a signal (-1: shorter than one window)."""
    return int(_lib().ecoz2b200_lpa_num_frames(num_samples, float(sample_rate), window_ms, offset_ms))


def lpa_on_signal(signal, sample_rate: float, P: int, window_ms: int = 45, offset_ms: int = 15,
                  return_errors: bool = False):
    """signal (host array) -> vectors [T, P+1] float64, through host buffers."""
    x = np.ascontiguousarray(signal, dtype=np.float64)
    T = num_frames(len(x), sample_rate, window_ms, offset_ms)
    if T < 0:
        raise LpcError(_err())
    out = np.empty((T, P + 1), dtype=np.float64)
    ne = C.c_int(0)
    got = _lib().ecoz2b200_lpa_on_signal(x.ctypes.data_as(_dp), len(x), float(sample_rate), P, window_ms,
                                         offset_ms, out.ctypes.data_as(_dp), T, C.byref(ne))
    if got < 0:
        raise LpcError(_err())
    assert got == T
    return (out, ne.value) if return_errors else out


def lpa_on_signal_device(d_signal_ptr: int, num_samples: int, sample_rate: float, P: int, window_ms: int,
                         offset_ms: int, d_vectors_ptr: int, capacity_frames: int, stream: int = 0) -> int:
    """Device-resident variant: raw device pointers (e.g. torch tensors' data_ptr()); returns T."""
    ne = C.c_int(0)
    got = _lib().ecoz2b200_lpa_on_signal_device(d_signal_ptr, num_samples, float(sample_rate), P, window_ms,
                                                offset_ms, d_vectors_ptr, capacity_frames, C.byref(ne),
                                                stream or None)
    if got < 0:
        raise LpcError(_err())
    return int(got)


def lpca(x, P: int):
    """-> (status, r, rc, a, pe) like ecoz2_lpca."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    r = np.zeros(P + 1)
    rc = np.zeros(P + 1)
    a = np.zeros(P + 1)
    pe = C.c_double(0.)
    st = _lib().ecoz2_lpca(x.ctypes.data_as(_dp), len(x), P, r.ctypes.data_as(_dp), rc.ctypes.data_as(_dp),
                           a.ctypes.data_as(_dp), C.byref(pe))
    if st < 0:
        raise LpcError(_err() or "ecoz2_lpca failed")
    return st, r, rc, a, pe.value


def lpc_signals(P: int, window_ms: int, offset_ms: int, sgn_filenames, minpc: int = 0, split: float = 0.,
                mintrpt: float = 0., verbose: bool = False) -> int:
    """ecoz2_lpc_signals: wav files -> data/predictors[/TRAIN|/TEST]/<class>/<name>.prd (cwd-relative)."""
    arr = (C.c_char_p * len(sgn_filenames))(*[f.encode() for f in sgn_filenames])
    return int(_lib().ecoz2_lpc_signals(P, window_ms, offset_ms, minpc, split, arr, len(sgn_filenames),
                                        mintrpt, 1 if verbose else 0))
