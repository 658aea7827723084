"""GPU parity of the LPC front end, through the C ABI
(include/ecoz2_lpc_b200.h): against the golden vectors the reference itself produced, and
against the numpy oracle on seeded signals (edge cases: one frame, ragged tail, silence,
DC, orders that need several lag passes / the any-order recursion).
Tolerance 1e-9 relative: FP64 sums in a different order (lpca_rs.rs:262 uses the same)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "lpc_front.npz")
RTOL = 1e-9


def _rel(u, v):
    return float(np.max(np.abs(u - v) / np.maximum(np.abs(v), 1e-300)))


def _signal(n, seed):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 32000.0
    x = 0.3 * np.sin(2 * np.pi * 440.0 * t * (1 + 0.1 * np.sin(2 * np.pi * 0.5 * t)))
    w = rng.standard_normal(n)
    w[1:] += 0.8 * w[:-1]
    return (x + 0.05 * w + 0.01).astype(np.float32).astype(np.float64)


def test_lpca_reference_fixture(built_lib):
    from ecoz2rs_b200 import lpc
    g = np.load(GOLD)
    st, r, rc, a, pe = lpc.lpca(g["frame_x"], 36)
    assert st == 0
    assert _rel(r, g["frame_r"]) < RTOL
    assert _rel(rc[1:], g["frame_rc"][1:]) < RTOL
    assert _rel(a, g["frame_a"]) < RTOL
    assert abs(pe - float(g["frame_pe"])) < RTOL * float(g["frame_pe"])


def test_lpa_on_signal_golden(built_lib):
    from ecoz2rs_b200 import lpc
    g = np.load(GOLD)
    for i, (P, W, O) in enumerate(g["sig_cfgs"]):
        want = g[f"sig_v{i}"]
        got, ne = lpc.lpa_on_signal(g["sig_x"], float(g["sig_rate"]), int(P), int(W), int(O), return_errors=True)
        assert got.shape == want.shape and ne == 0
        assert _rel(got, want) < RTOL, (P, W, O)


@pytest.mark.parametrize("P,W,O,n", [
    (36, 45, 15, 1440),            # exactly one frame
    (36, 45, 15, 1440 + 479),      # ragged tail dropped
    (12, 45, 15, 320000),          # 10 s
    (20, 20, 5, 100001),
    (36, 45, 45, 64000),           # no overlap
    (60, 45, 15, 32000),           # two passes of 40 lags, any-order recursion
    (100, 45, 15, 20000),
    (5, 2, 1, 3000),               # tiny windows (64 samples)
    (36, 500, 100, 64000),         # window too long to stage 32 frames
    (36, 3000, 1000, 400000),      # window too long to stage at all
])
def test_lpa_on_signal_matches_oracle(built_lib, P, W, O, n):
    from ecoz2rs_b200 import lpc
    from oracle import lpc_oracle as lo
    x = _signal(n, 100 + P + n % 97)
    want = lo.lpa_on_signal(x, 32000., P, W, O)
    got = lpc.lpa_on_signal(x, 32000., P, W, O)
    assert got.shape == want.shape
    # high orders on short windows are ill-conditioned: the recursion amplifies rounding
    tol = RTOL if P <= 44 else 1e-6
    assert _rel(got, want) < tol


def test_silence_dc_and_errors(built_lib):
    from ecoz2rs_b200 import lpc
    from oracle import lpc_oracle as lo
    # digital silence in the middle: lpca status 1 on all-zero frames, vectors stay zero
    x = _signal(32000, 7)
    x[8000:16000] = 0.0
    got, ne = lpc.lpa_on_signal(x, 32000., 12, 45, 15, return_errors=True)
    want = lo.lpa_on_signal(x, 32000., 12, 45, 15)
    zero = np.all(want == 0.0, axis=1)
    assert zero.any() and ne >= int(zero.sum())
    assert np.all(got[zero] == 0.0)
    ok = ~zero & np.all(np.isfinite(want), axis=1)
    assert _rel(got[ok], want[ok]) < 1e-7
    # too short
    assert lpc.num_frames(100, 32000., 45, 15) == -1
    with pytest.raises(lpc.LpcError, match="too short"):
        lpc.lpa_on_signal(np.zeros(100), 32000., 12, 45, 15)
    with pytest.raises(lpc.LpcError, match="order"):
        lpc.lpa_on_signal(x, 32000., 300, 45, 15)
    # lpca on an all-zero frame: status 1 like lpca.c:23-24
    st, r, rc, a, pe = lpc.lpca(np.zeros(256), 12)
    assert st == 1 and np.all(r == 0.0)


def test_lpca_many_lengths(built_lib):
    from ecoz2rs_b200 import lpc
    from oracle import lpc_oracle as lo
    for N, P in ((37, 36), (64, 12), (1000, 20), (1441, 36), (5000, 44), (300, 80)):
        x = _signal(N, N + P)
        st, r, rc, a, pe = lpc.lpca(x, P)
        r0, rc0, a0, pe0, st0 = lo.lpca(x[None, :], P)
        assert st == int(st0[0])
        assert _rel(r, r0[0]) < RTOL
        if st == 0:
            tol = RTOL if P <= 44 else 1e-6
            assert _rel(rc[1:], rc0[0][1:]) < tol and abs(pe - pe0[0]) < tol * abs(pe0[0])


def test_device_resident_matches_host_path(built_lib):
    import torch
    from ecoz2rs_b200 import lpc
    x = _signal(640000, 99)
    want = lpc.lpa_on_signal(x, 32000., 36, 45, 15)
    dx = torch.from_numpy(x).cuda()
    T = lpc.num_frames(len(x), 32000., 45, 15)
    dv = torch.empty((T, 37), dtype=torch.float64, device="cuda")
    got_T = lpc.lpa_on_signal_device(dx.data_ptr(), len(x), 32000., 36, 45, 15, dv.data_ptr(), T)
    assert got_T == T
    assert np.array_equal(dv.cpu().numpy(), want)


def test_lpc_signals_writes_prd_files(built_lib, tmp_path):
    """wav -> .prd through ecoz2_lpc_signals; the .prd feeds the VQ path's own reader."""
    import struct
    from ecoz2rs_b200 import lpc, formats
    from oracle import lpc_oracle as lo
    x = _signal(48000, 3)
    pcm = np.clip(np.round(x * 32768.0), -32768, 32767).astype("<i2")
    d = tmp_path / "signals" / "whale"
    d.mkdir(parents=True)
    wav = d / "a.wav"
    hdr = b"RIFF" + struct.pack("<I", 36 + pcm.nbytes) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 1, 32000, 64000, 2, 16)
    wav.write_bytes(hdr + b"data" + struct.pack("<I", pcm.nbytes) + pcm.tobytes())
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        assert lpc.lpc_signals(36, 45, 15, [str(wav)]) == 0
        out = tmp_path / "data" / "predictors" / "whale" / "a.prd"
        assert out.exists()
        cls, vec = formats.read_prd(str(out))
    finally:
        os.chdir(cwd)
    assert cls == "whale"
    want = lo.lpa_on_signal((pcm.astype(np.float32) / np.float32(32768.0)).astype(np.float64), 32000., 36, 45, 15)
    assert vec.shape == want.shape and _rel(vec, want) < RTOL


def test_device_signal_not_16_byte_aligned(built_lib):
    """A signal pointer that is only 8-byte aligned takes the cp.async staging instead of the bulk
    copy; odd window lengths exercise the hand-copied last element.  Same bits either way."""
    import torch
    from ecoz2rs_b200 import lpc
    for rate, P, W, O in ((32000., 36, 45, 15), (22050., 12, 45, 15), (11025., 20, 45, 15)):
        x = _signal(200001, 41)
        want = lpc.lpa_on_signal(x, rate, P, W, O)
        buf = torch.empty(len(x) + 1, dtype=torch.float64, device="cuda")
        buf[1:].copy_(torch.from_numpy(x))
        T = lpc.num_frames(len(x), rate, W, O)
        dv = torch.empty((T, P + 1), dtype=torch.float64, device="cuda")
        assert (buf.data_ptr() + 8) % 16 == 8
        assert lpc.lpa_on_signal_device(buf.data_ptr() + 8, len(x), rate, P, W, O, dv.data_ptr(), T) == T
        assert np.array_equal(dv.cpu().numpy(), want), (rate, P)


def test_cli_wav_to_symbols(built_lib, tmp_path):
    """The whole chain with the reference's command lines: `ecoz2 lpc` (wav -> .prd), `ecoz2 vq learn`
    on those predictors, `ecoz2 vq quantize` -> .seq; the .prd holds what the oracle computes."""
    import struct
    import subprocess
    from ecoz2rs_b200 import formats
    from oracle import lpc_oracle as lo
    exe = os.path.join(ROOT, "ecoz2rs_b200", "csrc", "ecoz2")
    wavs = []
    for k, cls in enumerate(("A", "A", "Bm")):
        x = _signal(64000 + 1000 * k, 50 + k)
        pcm = np.clip(np.round(x * 32768.0), -32768, 32767).astype("<i2")
        d = tmp_path / "signals" / cls
        d.mkdir(parents=True, exist_ok=True)
        w = d / f"s{k}.wav"
        hdr = b"RIFF" + struct.pack("<I", 36 + pcm.nbytes) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 1, 32000, 64000, 2, 16)
        w.write_bytes(hdr + b"data" + struct.pack("<I", pcm.nbytes) + pcm.tobytes())
        wavs.append((w, cls, pcm))
    env = dict(os.environ, ECOZ2_VQ_MAX_M="16", ECOZ2_VQ_MIN_DISTS="0")
    r = subprocess.run([exe, "lpc", "-P", "20", "-W", "45", "-O", "15", "--signals", str(tmp_path / "signals")],
                       cwd=tmp_path, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    total = 0
    for w, cls, pcm in wavs:
        got_cls, vec = formats.read_prd(str(tmp_path / "data" / "predictors" / cls / (w.stem + ".prd")))
        want = lo.lpa_on_signal((pcm.astype(np.float32) / np.float32(32768.0)).astype(np.float64), 32000., 20, 45, 15)
        assert got_cls == cls and vec.shape == want.shape and _rel(vec, want) < RTOL
        total += len(vec)
    r = subprocess.run([exe, "vq", "learn", "-P", "20", "--predictors", str(tmp_path / "data" / "predictors")],
                       cwd=tmp_path, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "Ecoz2ObserverRef.step: M=16" in r.stdout, r.stdout + r.stderr
    cb = tmp_path / "data" / "codebooks" / "_" / "eps_0.05_M_0016.cbook"
    r = subprocess.run([exe, "vq", "quantize", "--codebook", str(cb), "--predictors",
                        str(tmp_path / "data" / "predictors")], cwd=tmp_path, env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    n = 0
    for w, cls, _ in wavs:
        c, M, seq = formats.read_seq(str(tmp_path / "data" / "sequences" / "M16" / cls / (w.stem + ".seq")))
        assert (c, M) == (cls, 16) and seq.max() < 16
        n += len(seq)
    assert n == total


def test_signal_to_symbols_without_leaving_hbm(built_lib):
    """the vectors of lpa_on_signal_device are handed to the VQ engine with
    frames_append_device; codebook learning and the symbols equal the route through host arrays."""
    import torch
    from ecoz2rs_b200 import lpc, vq
    P, W, O = 20, 45, 15
    x = _signal(32000 * 120, 17)
    host_vec = lpc.lpa_on_signal(x, 32000., P, W, O)
    dx = torch.from_numpy(x).cuda()
    T = lpc.num_frames(len(x), 32000., W, O)
    dv = torch.empty((T, P + 1), dtype=torch.float64, device="cuda")
    assert lpc.lpa_on_signal_device(dx.data_ptr(), len(x), 32000., P, W, O, dv.data_ptr(), T) == T
    with vq.VqEngine(P, max_M=64) as a, vq.VqEngine(P, max_M=64) as b:
        a.set_frames_device(dv.data_ptr(), T)
        b.set_frames(host_vec)
        for e in (a, b):
            e.codebook_initial()
        sa, _ = a.learn(eps=0.05, stop_M=64)
        sb, _ = b.learn(eps=0.05, stop_M=64)
        assert [s["passes"] for s in sa] == [s["passes"] for s in sb]
        # the cell sums meet in FP64 atomics: equal to rounding, not bit for bit, between two runs
        assert np.allclose(sa[-1]["refl"], sb[-1]["refl"], rtol=1e-9, atol=1e-12)
        for e in (a, b):
            e.set_reflections(sa[-1]["refl"])
        seq_a, avg_a = a.quantize_resident()
        seq_b, avg_b = b.quantize(host_vec)
        assert np.array_equal(seq_a, seq_b) and abs(avg_a - avg_b) <= 1e-12 * abs(avg_b)
        assert len(np.unique(seq_a)) > 32
