"""Host-side mirror of the reference's `vq` call surface over the C-ABI library.

The product is `csrc/libecoz2_vq_b200.so` (hand-written sm_100a CUDA kernels + a
C++ host engine, no torch inside).  This module is the thin ctypes binding the
parity tests, `bench.py` and Python users call; it mirrors

* the three public entry points of the reference's C API
  (`ecoz2_vq_learn`, `ecoz2_vq_learn_using_base_codebook`, `ecoz2_vq_quantize`,
  reference ecoz2/src/include/ecoz2.h:56-87, bound by the Rust crate at
  src/ecoz2_lib/mod.rs:96-122 and wrapped at :252-342), and
* the in-memory steps of the algorithm (`ecoz2b200_*`, include/ecoz2_vq_b200.h).

There is no CPU fallback: without the built library or without a CUDA device the
calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ECOZ2_VQ_B200_LIB") or os.path.join(_HERE, "csrc", "libecoz2_vq_b200.so")

MODE_FP32, MODE_FP64, MODE_TENSOR = 0, 1, 2
OK, ERR_ARG, ERR_IO, ERR_CUDA, ERR_NCCL = 0, 1, 2, 3, 4     # ECOZ2B200_OK / ECOZ2B200_ERR_*
MODE_TF32X3 = MODE_TENSOR   # first name of the mode, kept as an alias

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_up = C.POINTER(C.c_uint16)

VQ_LEARN_CALLBACK = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double)


class StageInfo(C.Structure):
    _fields_ = [("M", C.c_int), ("passes", C.c_int), ("avgDistortion", C.c_double),
                ("sigma", C.c_double), ("inertia", C.c_double), ("reflections", _dp),
                ("raas", _dp), ("cardd", _ip), ("discel", _dp), ("stage_seconds", C.c_double),
                ("assign_ms", C.c_double), ("evals", C.c_double)]


class Counters(C.Structure):
    _fields_ = [("assign_ms", C.c_double), ("accumulate_ms", C.c_double),
                ("update_ms", C.c_double), ("allreduce_ms", C.c_double),
                ("kernel_launches", C.c_long), ("evals", C.c_double), ("rechecked", C.c_long),
                ("search_main_ms", C.c_double), ("search_main_launches", C.c_long),
                ("audited", C.c_long), ("audit_idx_diff", C.c_long), ("audit_mismatches", C.c_long)]


_STAGE_CB = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(StageInfo))
_PASS_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int)

# every symbol include/ecoz2_vq_b200.h declares (tests check the library exports them all)
EXPORTED_SYMBOLS = (
    "ecoz2_vq_learn", "ecoz2_vq_learn_using_base_codebook", "ecoz2_vq_quantize", "ecoz2_vq_classify",
    "vq_learn", "vq_learn_using_base_codebook", "vq_quantize", "vq_classify", "quantize",
    "ecoz2b200_version", "ecoz2b200_last_error", "ecoz2b200_device_count",
    "ecoz2b200_create", "ecoz2b200_destroy", "ecoz2b200_set_mode", "ecoz2b200_set_recheck_kappa",
    "ecoz2b200_set_recheck_kappa_tc", "ecoz2b200_set_audit", "ecoz2b200_debug_tc_distortions",
    "ecoz2b200_tc_issued_flops_per_eval", "ecoz2b200_write_min_dists_csv", "ecoz2b200_min_dists_text", "ecoz2b200_debug_format_lines", "ecoz2b200_format_lines_noalloc",
    "ecoz2b200_nccl_unique_id", "ecoz2b200_comm_init", "ecoz2b200_shard_range", "ecoz2b200_packed_layout",
    "ecoz2b200_frames_clear", "ecoz2b200_frames_reserve", "ecoz2b200_frames_append_host",
    "ecoz2b200_frames_append_device", "ecoz2b200_frames_commit", "ecoz2b200_frames_count",
    "ecoz2b200_codebook_initial", "ecoz2b200_codebook_set_reflections",
    "ecoz2b200_codebook_set_raas", "ecoz2b200_codebook_grow", "ecoz2b200_codebook_size",
    "ecoz2b200_codebook_get",
    "ecoz2b200_assign", "ecoz2b200_kmeans_pass", "ecoz2b200_kmeans_pass_host", "ecoz2b200_sigma",
    "ecoz2b200_inertia",
    "ecoz2b200_last_assignment", "ecoz2b200_learn",
    "ecoz2b200_quantize_host", "ecoz2b200_quantize_resident", "ecoz2b200_quantize_models",
    "ecoz2b200_counters_reset", "ecoz2b200_counters_get", "ecoz2b200_ffma_peak_tflops",
    "ecoz2b200_prd_info", "ecoz2b200_prd_read", "ecoz2b200_cbook_save", "ecoz2b200_cbook_info",
    "ecoz2b200_cbook_read", "ecoz2b200_seq_save", "ecoz2b200_class_name_of", "ecoz2b200_output_filename",
)

_lib = None


def lib():
    """Load the C-ABI library (fails loudly if it was not built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (make -C ecoz2rs_b200/csrc). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        vp, i, d, l = C.c_void_p, C.c_int, C.c_double, C.c_long
        cpp = C.POINTER(C.c_char_p)
        sig = {
            "ecoz2_vq_learn": (i, [i, d, C.c_char_p, cpp, i, vp, VQ_LEARN_CALLBACK]),
            "ecoz2_vq_learn_using_base_codebook": (i, [C.c_char_p, d, cpp, i, vp, VQ_LEARN_CALLBACK]),
            "ecoz2_vq_quantize": (i, [C.c_char_p, cpp, i, i]),
            "ecoz2_vq_classify": (i, [cpp, i, cpp, i, i]),
            "ecoz2b200_version": (C.c_char_p, []),
            "ecoz2b200_last_error": (C.c_char_p, []),
            "ecoz2b200_device_count": (i, []),
            "ecoz2b200_create": (vp, [_ip, i, i, i]),
            "ecoz2b200_destroy": (None, [vp]),
            "ecoz2b200_set_mode": (i, [vp, i]),
            "ecoz2b200_set_recheck_kappa": (i, [vp, d]),
            "ecoz2b200_set_recheck_kappa_tc": (i, [vp, d]),
            "ecoz2b200_set_audit": (i, [vp, i]),
            "ecoz2b200_debug_tc_distortions": (i, [vp, vp, l]),
            "ecoz2b200_tc_issued_flops_per_eval": (i, [i, _ip]),
            "ecoz2b200_write_min_dists_csv": (i, [C.c_char_p, _ip, vp, l]),
            "ecoz2b200_min_dists_text": (i, [vp, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
            "ecoz2b200_debug_format_lines": (i, [_ip, vp, l, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
            "ecoz2b200_format_lines_noalloc": (i, [vp, vp, l, vp, C.c_size_t, vp, vp]),
            "ecoz2b200_nccl_unique_id": (i, [C.c_char_p]),
            "ecoz2b200_comm_init": (i, [vp, C.c_char_p, i, i]),
            "ecoz2b200_shard_range": (i, [l, i, i, C.POINTER(l), C.POINTER(l)]),
            "ecoz2b200_packed_layout": (i, [i, i, C.POINTER(l)]),
            "ecoz2b200_frames_clear": (i, [vp]),
            "ecoz2b200_frames_reserve": (i, [vp, l]),
            "ecoz2b200_frames_append_host": (i, [vp, vp, l]),
            "ecoz2b200_frames_append_device": (i, [vp, vp, l]),
            "ecoz2b200_frames_commit": (i, [vp]),
            "ecoz2b200_frames_count": (l, [vp]),
            "ecoz2b200_codebook_initial": (i, [vp]),
            "ecoz2b200_codebook_set_reflections": (i, [vp, _dp, i]),
            "ecoz2b200_codebook_set_raas": (i, [vp, _dp, i]),
            "ecoz2b200_codebook_grow": (i, [vp]),
            "ecoz2b200_codebook_size": (i, [vp]),
            "ecoz2b200_codebook_get": (i, [vp, _dp, _dp]),
            "ecoz2b200_assign": (i, [vp, _ip, _dp, C.POINTER(l)]),
            "ecoz2b200_kmeans_pass": (i, [vp, _dp, _ip, _dp, _dp]),
            "ecoz2b200_kmeans_pass_host": (i, [vp, vp, l, _dp, _ip, _dp, _dp]),
            "ecoz2b200_sigma": (i, [vp, d, _dp]),
            "ecoz2b200_inertia": (i, [vp, _dp]),
            "ecoz2b200_last_assignment": (i, [vp, _ip, _dp]),
            "ecoz2b200_learn": (i, [vp, d, i, _STAGE_CB, _PASS_CB, vp]),
            "ecoz2b200_quantize_host": (i, [vp, vp, l, vp, _dp]),
            "ecoz2b200_quantize_resident": (i, [vp, vp, _dp]),
            "ecoz2b200_quantize_models": (i, [vp, vp, i, i, vp, i, vp, vp, C.POINTER(C.c_int)]),
            "ecoz2b200_counters_reset": (i, [vp]),
            "ecoz2b200_counters_get": (i, [vp, C.POINTER(Counters)]),
            "ecoz2b200_ffma_peak_tflops": (d, [i, d]),
            "ecoz2b200_prd_info": (i, [C.c_char_p, C.POINTER(i), C.POINTER(i), C.c_char_p]),
            "ecoz2b200_prd_read": (i, [C.c_char_p, _dp, l]),
            "ecoz2b200_cbook_save": (i, [C.c_char_p, C.c_char_p, _dp, i, i]),
            "ecoz2b200_cbook_info": (i, [C.c_char_p, C.POINTER(i), C.POINTER(i), C.c_char_p]),
            "ecoz2b200_cbook_read": (i, [C.c_char_p, _dp, l]),
            "ecoz2b200_seq_save": (i, [C.c_char_p, C.c_char_p, _up, i, i]),
            "ecoz2b200_class_name_of": (i, [C.c_char_p, C.c_char_p]),
            "ecoz2b200_output_filename": (i, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, l]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def device_count() -> int:
    return lib().ecoz2b200_device_count()


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _d(a):
    return a.ctypes.data_as(_dp)


class VqError(RuntimeError):
    pass


def _check(rc: int, what: str):
    if rc != 0:
        raise VqError(f"{what} failed (code {rc}): {lib().ecoz2b200_last_error().decode()}")


_nccl_preloaded = False


def _preload_bundled_nccl():
    """The library binds NCCL with dlopen("libnccl.so.2") and the dynamic loader keeps one library
    per soname: if that were the system copy and torch is imported later in the same process, torch
    would be handed an NCCL older than the one it was built against.  Load the copy bundled with
    torch's wheels first (when there is one), so both use the same."""
    global _nccl_preloaded
    if _nccl_preloaded:
        return
    _nccl_preloaded = True
    try:
        import glob
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for d in (spec.submodule_search_locations if spec else ()):
            for f in sorted(glob.glob(os.path.join(d, "lib", "libnccl.so*"))):
                C.CDLL(f, mode=C.RTLD_GLOBAL)
                return
    except Exception:
        pass


class VqEngine:
    """Frames of this process on one or more local B200s + the replicated codebook."""

    def __init__(self, P: int, max_M: int = 4096, devices: Sequence[int] = (0,), mode: Optional[int] = None):
        L = lib()
        if len(devices) > 1:
            _preload_bundled_nccl()
        devs = (C.c_int32 * len(devices))(*devices)
        self._h = L.ecoz2b200_create(devs, len(devices), P, max_M)
        if not self._h:
            raise VqError("ecoz2b200_create failed: " + L.ecoz2b200_last_error().decode())
        self.P, self.D, self.max_M = P, P + 1, max_M
        self.devices = tuple(devices)
        if mode is not None:
            self.set_mode(mode)

    def close(self):
        if getattr(self, "_h", None):
            lib().ecoz2b200_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- configuration
    def set_mode(self, mode: int):
        _check(lib().ecoz2b200_set_mode(self._h, mode), "set_mode")

    def set_recheck_kappa(self, kappa: float):
        _check(lib().ecoz2b200_set_recheck_kappa(self._h, kappa), "set_recheck_kappa")

    def set_recheck_kappa_tc(self, kappa: float):
        _check(lib().ecoz2b200_set_recheck_kappa_tc(self._h, kappa), "set_recheck_kappa_tc")

    def set_audit(self, every: int):
        """Every `every`-th frame of every search also gets the all-FP64 scan (0 = off); see counters()."""
        _check(lib().ecoz2b200_set_audit(self._h, every), "set_audit")

    def debug_tc_distortions(self, n_rows: int) -> np.ndarray:
        """Raw tensor-core (split-TF32) distortion values of the first n_rows resident frames."""
        out = np.empty((n_rows, self.K), dtype=np.float32)
        _check(lib().ecoz2b200_debug_tc_distortions(self._h, out.ctypes.data, n_rows),
               "debug_tc_distortions")
        return out

    def comm_init(self, unique_id: bytes, rank: int, world_size: int):
        assert len(unique_id) == 128
        _preload_bundled_nccl()
        _check(lib().ecoz2b200_comm_init(self._h, unique_id, rank, world_size), "comm_init")

    # -- frames
    def set_frames(self, x: np.ndarray):
        """Upload host frames [N, P+1] float64 (replaces any resident frames)."""
        x = _f64(x)
        assert x.ndim == 2 and x.shape[1] == self.D
        L = lib()
        _check(L.ecoz2b200_frames_clear(self._h), "frames_clear")
        _check(L.ecoz2b200_frames_reserve(self._h, x.shape[0]), "frames_reserve")
        _check(L.ecoz2b200_frames_append_host(self._h, x.ctypes.data, x.shape[0]), "frames_append_host")
        _check(L.ecoz2b200_frames_commit(self._h), "frames_commit")

    def set_frames_ptr(self, host_ptr: int, n: int):
        """Same from a raw host pointer (e.g. pinned memory owned by the caller)."""
        L = lib()
        _check(L.ecoz2b200_frames_clear(self._h), "frames_clear")
        _check(L.ecoz2b200_frames_reserve(self._h, n), "frames_reserve")
        _check(L.ecoz2b200_frames_append_host(self._h, host_ptr, n), "frames_append_host")
        _check(L.ecoz2b200_frames_commit(self._h), "frames_commit")

    def set_frames_device(self, dev_ptr: int, n: int):
        """Frames already in HBM on devices[0] (float64 [n, P+1], e.g. a torch tensor's data_ptr)."""
        L = lib()
        _check(L.ecoz2b200_frames_clear(self._h), "frames_clear")
        _check(L.ecoz2b200_frames_reserve(self._h, n), "frames_reserve")
        _check(L.ecoz2b200_frames_append_device(self._h, dev_ptr, n), "frames_append_device")
        _check(L.ecoz2b200_frames_commit(self._h), "frames_commit")

    @property
    def num_frames(self) -> int:
        return lib().ecoz2b200_frames_count(self._h)

    # -- codebook
    def codebook_initial(self):
        _check(lib().ecoz2b200_codebook_initial(self._h), "codebook_initial")

    def set_reflections(self, refl: np.ndarray):
        refl = _f64(refl)
        assert refl.ndim == 2 and refl.shape[1] == self.D
        _check(lib().ecoz2b200_codebook_set_reflections(self._h, _d(refl), refl.shape[0]),
               "codebook_set_reflections")

    def set_raas(self, raas: np.ndarray):
        raas = _f64(raas)
        assert raas.ndim == 2 and raas.shape[1] == self.D
        _check(lib().ecoz2b200_codebook_set_raas(self._h, _d(raas), raas.shape[0]), "codebook_set_raas")

    def grow(self):
        _check(lib().ecoz2b200_codebook_grow(self._h), "codebook_grow")

    @property
    def K(self) -> int:
        return lib().ecoz2b200_codebook_size(self._h)

    def codebook(self):
        """-> (reflections [K, D], raas [K, D])"""
        K = self.K
        refl = np.empty((K, self.D))
        raas = np.empty((K, self.D))
        _check(lib().ecoz2b200_codebook_get(self._h, _d(refl), _d(raas)), "codebook_get")
        return refl, raas

    # -- steps
    def assign(self):
        """-> (idx int32 [N], ddmin float64 [N], frames re-evaluated in FP64)"""
        N = self.num_frames
        idx = np.empty(N, np.int32)
        dd = np.empty(N, np.float64)
        nre = C.c_long(0)
        _check(lib().ecoz2b200_assign(self._h, idx.ctypes.data_as(_ip), _d(dd), C.byref(nre)), "assign")
        return idx, dd, nre.value

    def kmeans_pass(self, want_cells: bool = True):
        K = self.K
        cells = np.empty((K, self.D)) if want_cells else None
        cardd = np.empty(K, np.int32)
        discel = np.empty(K)
        DD = C.c_double(0)
        _check(lib().ecoz2b200_kmeans_pass(self._h, _d(cells) if want_cells else None,
                                           cardd.ctypes.data_as(_ip), _d(discel), C.byref(DD)),
               "kmeans_pass")
        return dict(cells=cells, cardd=cardd, discel=discel, DD=DD.value)

    def kmeans_pass_host_ptr(self, host_ptr: int, n: int, want_cells: bool = True):
        """One pass over n frames at host address host_ptr (pinned memory for full speed): the upload
        overlaps the search chunk by chunk; the frames replace the resident set."""
        K = self.K
        cells = np.empty((K, self.D)) if want_cells else None
        cardd = np.empty(K, np.int32)
        discel = np.empty(K)
        DD = C.c_double(0)
        _check(lib().ecoz2b200_kmeans_pass_host(self._h, host_ptr, n, _d(cells) if want_cells else None,
                                                cardd.ctypes.data_as(_ip), _d(discel), C.byref(DD)),
               "kmeans_pass_host")
        return dict(cells=cells, cardd=cardd, discel=discel, DD=DD.value)

    def kmeans_pass_host(self, x: np.ndarray, want_cells: bool = True):
        x = _f64(x)
        assert x.ndim == 2 and x.shape[1] == self.D
        return self.kmeans_pass_host_ptr(x.ctypes.data, x.shape[0], want_cells)

    def sigma(self, avg_distortion: float) -> float:
        out = C.c_double(0)
        _check(lib().ecoz2b200_sigma(self._h, avg_distortion, C.byref(out)), "sigma")
        return out.value

    def inertia(self) -> float:
        out = C.c_double(0)
        _check(lib().ecoz2b200_inertia(self._h, C.byref(out)), "inertia")
        return out.value

    def last_assignment(self):
        N = self.num_frames
        idx = np.empty(N, np.int32)
        md = np.empty(N, np.float64)
        _check(lib().ecoz2b200_last_assignment(self._h, idx.ctypes.data_as(_ip), _d(md)), "last_assignment")
        return idx, md

    def min_dists_text(self) -> bytes:
        """The "codeword,minDistortion" lines of the last pass (report.c:118-121), formatted on the device."""
        ptr, n = C.c_void_p(0), C.c_size_t(0)
        _check(lib().ecoz2b200_min_dists_text(self._h, C.byref(ptr), C.byref(n)), "min_dists_text")
        return C.string_at(ptr.value, n.value)

    def learn(self, eps: float = 0.05, stop_M: int = 4096, keep_assignments: bool = False,
              on_stage: Optional[Callable[[dict], None]] = None):
        """Full LBG run from the current codebook. -> (stages, trace) like oracle.learn."""
        D = self.D
        stages, trace = [], []

        def _stage(_ctx, info_p):
            s = info_p.contents
            M = s.M
            st = dict(M=M, passes=s.passes, avg_distortion=s.avgDistortion, sigma=s.sigma,
                      inertia=s.inertia,
                      refl=np.ctypeslib.as_array(s.reflections, (M, D)).copy(),
                      raas=np.ctypeslib.as_array(s.raas, (M, D)).copy(),
                      cardd=np.ctypeslib.as_array(s.cardd, (M,)).copy(),
                      discel=np.ctypeslib.as_array(s.discel, (M,)).copy(),
                      stage_seconds=s.stage_seconds, assign_ms=s.assign_ms, evals=s.evals)
            if keep_assignments:
                st["idx"], st["mindist"] = self.last_assignment()
            stages.append(st)
            if on_stage:
                on_stage(st)

        def _pass(_ctx, M, p, DD, secs, empty):
            trace.append((M, p, DD, secs))

        scb, pcb = _STAGE_CB(_stage), _PASS_CB(_pass)
        _check(lib().ecoz2b200_learn(self._h, eps, stop_M, scb, pcb, None), "learn")
        return stages, trace

    def quantize(self, x: np.ndarray):
        """Host frames in, symbols out (the e2e path). -> (seq uint16 [T], avg distortion)"""
        x = _f64(x)
        assert x.ndim == 2 and x.shape[1] == self.D
        seq = np.empty(x.shape[0], np.uint16)
        avg = C.c_double(0)
        _check(lib().ecoz2b200_quantize_host(self._h, x.ctypes.data, x.shape[0], seq.ctypes.data,
                                             C.byref(avg)), "quantize_host")
        return seq, avg.value

    def quantize_ptr(self, host_ptr: int, T: int, seq_ptr: int) -> float:
        avg = C.c_double(0)
        _check(lib().ecoz2b200_quantize_host(self._h, host_ptr, T, seq_ptr, C.byref(avg)), "quantize_host")
        return avg.value

    def quantize_resident(self, want_seq: bool = True):
        seq = np.empty(self.num_frames, np.uint16) if want_seq else None
        avg = C.c_double(0)
        _check(lib().ecoz2b200_quantize_resident(self._h, seq.ctypes.data if want_seq else None,
                                                 C.byref(avg)), "quantize_resident")
        return seq, avg.value

    def quantize_models(self, reflections: np.ndarray, seg_offsets=None, want_seq: bool = True):
        """cbook_quantize of all resident frames against R codebooks at once (reflections [R, K, 1+P]);
        seg_offsets = frame offsets of the predictor files (default: one segment).
        -> (seq [R, n] u16 or None, dists [R, num_segments], one_sweep)."""
        r = np.ascontiguousarray(reflections, dtype=np.float64)
        assert r.ndim == 3 and r.shape[2] == self.D
        R, K = r.shape[0], r.shape[1]
        n = self.num_frames
        so = np.ascontiguousarray([0, n] if seg_offsets is None else seg_offsets, dtype=np.int64)
        nseg = len(so) - 1
        seq = np.empty((R, n), np.uint16) if want_seq else None
        dists = np.empty((R, nseg), np.float64)
        one = C.c_int(0)
        _check(lib().ecoz2b200_quantize_models(self._h, r.ctypes.data, R, K, so.ctypes.data, nseg,
                                               seq.ctypes.data if want_seq else None, dists.ctypes.data,
                                               C.byref(one)), "quantize_models")
        return seq, dists, bool(one.value)

    def quantize_resident_ptr(self, seq_ptr: int) -> float:
        """Resident frames -> u16 symbols at host address seq_ptr (pinned memory for full speed)."""
        avg = C.c_double(0)
        _check(lib().ecoz2b200_quantize_resident(self._h, seq_ptr, C.byref(avg)), "quantize_resident")
        return avg.value

    # -- counters
    def counters_reset(self):
        lib().ecoz2b200_counters_reset(self._h)

    def counters(self) -> dict:
        c = Counters()
        _check(lib().ecoz2b200_counters_get(self._h, C.byref(c)), "counters_get")
        return {k: getattr(c, k) for k, _ in Counters._fields_}


# ---------------------------------------------------------------------- file-level API
def _file_array(paths: Sequence[str]):
    arr = (C.c_char_p * len(paths))(*[os.fsencode(p) for p in paths])
    return arr


def vq_learn(prediction_order: int, epsilon: float, codebook_class_name: str,
             predictor_filenames: Sequence[str],
             callback: Optional[Callable[[int, float, float, float], None]] = None) -> int:
    """`ecoz2_vq_learn` (reference src/ecoz2_lib/mod.rs:252-323). Outputs under ./data."""
    cb = VQ_LEARN_CALLBACK(lambda _t, M, a, s, i: callback(M, a, s, i) if callback else None)
    return lib().ecoz2_vq_learn(prediction_order, epsilon, codebook_class_name.encode(),
                                _file_array(predictor_filenames), len(predictor_filenames), None, cb)


def vq_learn_using_base_codebook(base_codebook: str, epsilon: float,
                                 predictor_filenames: Sequence[str],
                                 callback: Optional[Callable[[int, float, float, float], None]] = None) -> int:
    cb = VQ_LEARN_CALLBACK(lambda _t, M, a, s, i: callback(M, a, s, i) if callback else None)
    return lib().ecoz2_vq_learn_using_base_codebook(os.fsencode(base_codebook), epsilon,
                                                    _file_array(predictor_filenames),
                                                    len(predictor_filenames), None, cb)


def vq_quantize(nom_raas: str, predictor_filenames: Sequence[str], show_filenames: bool = False) -> int:
    """`ecoz2_vq_quantize` (reference src/ecoz2_lib/mod.rs:325-342)."""
    return lib().ecoz2_vq_quantize(os.fsencode(nom_raas), _file_array(predictor_filenames),
                                   len(predictor_filenames), int(show_filenames))


def vq_classify(cb_filenames: Sequence[str], predictor_filenames: Sequence[str],
                show_ranked: bool = False) -> int:
    """`ecoz2_vq_classify` (reference src/ecoz2_lib/mod.rs:344-360, vq/vq_classify.c:185-327)."""
    return lib().ecoz2_vq_classify(_file_array(cb_filenames), len(cb_filenames),
                                   _file_array(predictor_filenames), len(predictor_filenames),
                                   int(show_ranked))


def ffma_peak_tflops(device: int = 0, millis: float = 20.0) -> float:
    """Measured FP32 FFMA peak of `device` (TFLOP/s): denominator of the FP32 roofline."""
    return lib().ecoz2b200_ffma_peak_tflops(device, millis)


def tc_issued_flops_per_eval(P: int):
    """(tensor-core FLOPs the TF32X3 search issues per frame x codeword pair, operand format)."""
    f16 = C.c_int32(0)
    n = lib().ecoz2b200_tc_issued_flops_per_eval(P, C.byref(f16))
    return n, ("f16" if f16.value else "tf32")


def write_min_dists_csv(path: str, codeword: np.ndarray, min_dist: np.ndarray) -> None:
    """The learn report's `<cbook>.min_dists.csv` writer (host threads, exact "%f")."""
    cw = np.ascontiguousarray(codeword, dtype=np.int32)
    md = np.ascontiguousarray(min_dist, dtype=np.float64)
    assert cw.shape == md.shape and cw.ndim == 1
    _check(lib().ecoz2b200_write_min_dists_csv(os.fsencode(path), cw.ctypes.data_as(_ip), md.ctypes.data,
                                               cw.shape[0]), "write_min_dists_csv")


def format_lines_on_device(codeword: np.ndarray, min_dist: np.ndarray):
    """Test aid: the device report formatter on host arrays -> (return code, text bytes)."""
    cw = np.ascontiguousarray(codeword, dtype=np.int32)
    md = np.ascontiguousarray(min_dist, dtype=np.float64)
    buf = C.create_string_buffer(40 * max(1, cw.shape[0]))
    n = C.c_size_t(0)
    rc = lib().ecoz2b200_debug_format_lines(cw.ctypes.data_as(_ip), md.ctypes.data, cw.shape[0], buf, len(buf),
                                            C.byref(n))
    return rc, buf.raw[:n.value]


def last_error() -> str:
    """Message of the last failed call on this thread (ecoz2b200_last_error)."""
    return lib().ecoz2b200_last_error().decode()


NOALLOC_MAX_LINES = 4194304        # ECOZ2B200_NOALLOC_MAX_LINES


def format_lines_noalloc(d_codeword: int, d_min_dist: int, n: int, d_text: int, text_capacity: int,
                         d_bytes: int, stream: int = 0) -> int:
    """The device report formatter on caller-owned DEVICE buffers (raw addresses), queued on `stream`;
    allocates nothing (its scratch is a fixed device array of the library).  Returns the C return code;
    `last_error()` has the message of a failure."""
    return lib().ecoz2b200_format_lines_noalloc(d_codeword, d_min_dist, n, d_text, text_capacity, d_bytes,
                                                stream or None)


def nccl_unique_id() -> bytes:
    _preload_bundled_nccl()
    buf = C.create_string_buffer(128)
    _check(lib().ecoz2b200_nccl_unique_id(buf), "nccl_unique_id")
    return buf.raw
