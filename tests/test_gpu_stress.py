"""Race hunting without compute-sanitizer (closed on this GPU pool): (1) many repetitions of the LBG run,
the fused small-codebook pass, the LPC front end and the tensor-core search in one process, every
repetition compared with the first / with the all-FP64 search; (2) the parity tests of the search once
more with every SM's shared memory filled with NaN patterns before each search
(ECOZ2_VQ_DEBUG_POISON=1, read at start-up: hence a subprocess), so that a tile read before its
mbarrier fired cannot pass by luck of stale contents."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(cmd, env=None, timeout=500):
    e = dict(os.environ)
    e["PYTHONPATH"] = ROOT + os.pathsep + e.get("PYTHONPATH", "")
    if env:
        e.update(env)
    r = subprocess.run(cmd, cwd=ROOT, env=e, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, (r.stdout[-3000:] + "\n" + r.stderr[-3000:])
    return r.stdout


def test_repeated_learn_runs_are_identical(built_lib):
    out = _run([sys.executable, "tests/tools/stress_learn_repeat.py", "--reps", "12"])
    assert "STRESS OK" in out


def test_repeated_small_pass_and_lpc_are_identical(built_lib):
    out = _run([sys.executable, "tests/tools/stress_small_lpc.py", "--reps", "12"])
    assert "STRESS OK" in out


def test_repeated_tensor_core_search_equals_fp64(built_lib):
    _run([sys.executable, "tests/tools/stress_tc.py", "8"])


def test_repeated_passes_with_prestaged_operands_and_host_chunks(built_lib):
    """The search fed by bulk copies of pre-staged operands (no loader warps pacing the pipeline) and the
    chunked host pass, alternating, 3 x 24 repetitions: identical counts every time."""
    out = _run([sys.executable, "tests/tools/stress_pre.py", "24"])
    assert "STRESS OK" in out


def test_codebook_uploaded_right_after_create_survives_a_busy_default_stream(built_lib):
    """Nothing the constructor queues may land after the caller's first upload (it used cudaMemset on the
    legacy default stream once; see tests/tools/create_upload_order.py)."""
    out = _run([sys.executable, "tests/tools/create_upload_order.py", "20"])
    assert "STRESS OK" in out


def test_search_parity_with_poisoned_shared_memory(built_lib):
    out = _run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-x", "-p", "no:cacheprovider",
                "tests/test_gpu_tc.py", "tests/test_gpu_parity.py", "-k",
                "shapes or k4096 or assign or teacher_forced or fused or learn_end_to_end"],
               env={"ECOZ2_VQ_DEBUG_POISON": "1"})
    assert " passed" in out and "failed" not in out
