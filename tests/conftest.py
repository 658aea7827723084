"""pytest configuration: the `gpu` marker and shared fixtures.

`-m "not gpu"`: oracle vs golden vectors, formats, host logic, C-ABI symbol export
                (runs without a GPU in a few minutes).
`-m gpu`      : parity of the CUDA path against the oracle, through the C ABI.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def built_lib():
    """Build (if needed) and load the C-ABI library."""
    import __graft_entry__ as g
    g.build()
    from ecoz2rs_b200 import vq
    return vq.lib()


@pytest.fixture(scope="session")
def frames36():
    from ecoz2rs_b200.synth import ar_mix
    return ar_mix(20000, 36, seed=20260101)


@pytest.fixture(scope="session")
def frames12():
    from ecoz2rs_b200.synth import ar_mix
    return ar_mix(30000, 12, seed=20260105)


@pytest.fixture(scope="session")
def oracle_run36(frames36):
    """Oracle LBG 2 -> 256 on frames36: (stages, trace)."""
    from oracle import oracle as O
    return O.learn(frames36, eps=0.05, max_M=256, keep_assignments=True)


@pytest.fixture(scope="session")
def oracle_run12(frames12):
    from oracle import oracle as O
    return O.learn(frames12, eps=0.05, max_M=128, keep_assignments=True)


def allowed_index_mismatches(idx_gpu, idx_ref, ddmin_ref, second_ref, rel=1e-6):
    """north_star: indices identical except where the reference's best-two gap is
    below 1e-6 relative.  Returns the mismatching positions that are NOT exempt."""
    bad = np.nonzero(idx_gpu != idx_ref)[0]
    if bad.size == 0:
        return bad
    gap = (second_ref[bad] - ddmin_ref[bad]) / np.abs(ddmin_ref[bad])
    return bad[gap >= rel]
