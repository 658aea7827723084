"""CPU tests of the LPC front end: the numpy oracle (oracle/lpc_oracle.py)
against golden vectors produced by the reference itself (tests/golden/make_golden_lpc.py:
the reference's own fixture `signal_frame.inputs` through its C `lpca`, and seeded signals
through its `lpa_on_signal`), and the C-ABI symbol export of include/ecoz2_lpc_b200.h."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "lpc_front.npz")
RTOL = 1e-9   # the tolerance the reference's own test uses between its lpca variants (lpca_rs.rs:262)


def _rel(u, v):
    return float(np.max(np.abs(u - v) / np.maximum(np.abs(v), 1e-300)))


def test_oracle_lpca_matches_reference_on_its_fixture():
    from oracle import lpc_oracle as lo
    g = np.load(GOLD)
    x = g["frame_x"]
    assert len(x) == 1440 and int(g["frame_p"]) == 36
    r, rc, a, pe, st = lo.lpca(x[None, :], 36)
    assert st[0] == 0
    assert _rel(r[0], g["frame_r"]) < RTOL
    assert _rel(rc[0][1:], g["frame_rc"][1:]) < RTOL
    assert _rel(a[0], g["frame_a"]) < RTOL
    assert abs(pe[0] - float(g["frame_pe"])) < RTOL * float(g["frame_pe"])


def test_oracle_lpa_on_signal_matches_reference():
    from oracle import lpc_oracle as lo
    g = np.load(GOLD)
    for i, (P, W, O) in enumerate(g["sig_cfgs"]):
        want = g[f"sig_v{i}"]
        got = lo.lpa_on_signal(g["sig_x"], float(g["sig_rate"]), int(P), int(W), int(O))
        assert got.shape == want.shape
        assert _rel(got, want) < RTOL, (P, W, O)


def test_oracle_frame_geometry_edges():
    from oracle import lpc_oracle as lo
    # shorter than one window
    assert lo.frame_geometry(100, 32000., 45, 15)[2] == -1
    # exactly one window; tail shorter than an offset is dropped
    win, off, T = lo.frame_geometry(1440, 32000., 45, 15)
    assert (win, off, T) == (1440, 480, 1)
    assert lo.frame_geometry(1440 + 479, 32000., 45, 15)[2] == 1
    assert lo.frame_geometry(1440 + 480, 32000., 45, 15)[2] == 2
    # fractional sample rates are truncated (lpa_on_signal.c:84)
    assert lo.frame_geometry(48000, 22050.7, 45, 15)[:2] == (992, 330)


def test_lpc_header_symbols_exported(built_lib):
    from ecoz2rs_b200 import lpc
    with open(os.path.join(ROOT, "include", "ecoz2_lpc_b200.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(ecoz2b200_[a-z0-9_]+|ecoz2_lp[a-z_]+)\s*\(", header))
    assert declared == set(lpc.EXPORTED_SYMBOLS), declared ^ set(lpc.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(built_lib, name), name


def test_lpc_no_cpu_fallback(built_lib):
    from ecoz2rs_b200 import lpc, vq
    if vq.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(lpc.LpcError, match="no CUDA device"):
        lpc.lpa_on_signal(np.zeros(4000), 32000., 12)
    with pytest.raises(lpc.LpcError):
        lpc.lpca(np.ones(100), 12)
