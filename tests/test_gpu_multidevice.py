"""One process driving several GPUs (ECOZ2_VQ_GPUS=all / devices=(0, 1)) and work on a non-zero device:
the LBG run, the symbols and the LPC vectors do not depend on where they were computed.  Needs two
GPUs (skipped on a single-GPU box; run with `gpurun --gpus 2`).  torch is imported AFTER the first
multi-GPU call on purpose: the NCCL bound by the library must be one torch can live with."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_engines_on_several_devices_agree(built_lib):
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix
    if vq.device_count() < 2:
        pytest.skip("needs two GPUs")
    x = ar_mix(300_001, 36, seed=3)
    res = []
    for devs in ((0,), (0, 1), (1,)):
        with vq.VqEngine(36, max_M=512, devices=devs) as eng:
            eng.set_frames(x)
            eng.codebook_initial()
            st, _ = eng.learn(eps=0.05, stop_M=512)
            seq, avg = eng.quantize(x[:50_001])
            res.append(([(s["M"], s["passes"]) for s in st], st[-1]["avg_distortion"], st[-1]["refl"], seq, avg))
    for r in res[1:]:
        assert r[0] == res[0][0] and abs(r[1] - res[0][1]) <= 1e-9 * res[0][1]
        assert np.allclose(r[2], res[0][2], rtol=1e-7, atol=1e-10) and (r[3] != res[0][3]).mean() < 1e-4
    # LPC front end on device 1
    import torch
    from ecoz2rs_b200 import lpc
    sig = np.random.default_rng(0).standard_normal(32000 * 30) * 0.1
    want = lpc.lpa_on_signal(sig, 32000., 36, 45, 15)
    with torch.cuda.device(1):
        d = torch.from_numpy(sig).cuda()
        T = lpc.num_frames(len(sig), 32000., 45, 15)
        v = torch.empty((T, 37), dtype=torch.float64, device="cuda")
        assert lpc.lpa_on_signal_device(d.data_ptr(), len(sig), 32000., 36, 45, 15, v.data_ptr(), T) == T
        assert np.array_equal(v.cpu().numpy(), want)


def test_entry_points_leave_the_current_device_alone(built_lib, tmp_path, monkeypatch):
    """Every extern "C" entry point restores the caller's current CUDA context (DeviceGuard): a host
    that works on device 0 and runs an engine on device 1 still finds device 0 current afterwards."""
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix
    if vq.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")

    def current():
        return torch.cuda.current_device()

    x = ar_mix(50_000, 36, seed=5)
    assert current() == 0
    with vq.VqEngine(36, max_M=64, devices=(1,)) as eng:
        assert current() == 0
        eng.set_frames(x)
        assert current() == 0
        eng.codebook_initial()
        eng.learn(eps=0.05, stop_M=64)
        assert current() == 0
        eng.quantize(x[:1000])
        assert current() == 0
        eng.counters()
    assert current() == 0
    # the file-level entry point on devices (1, 0)
    f = str(tmp_path / "predictors" / "TRAIN" / "g" / "a.prd")
    formats.write_prd(f, x, "g")
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("ECOZ2_VQ_GPUS", "1,0")
    monkeypatch.setenv("ECOZ2_VQ_MAX_M", "16")
    assert vq.vq_learn(36, 0.05, "g", [f]) == 0
    assert current() == 0
    assert vq.ffma_peak_tflops(1, 5.0) > 0 and current() == 0


def test_file_level_entry_points_on_two_devices(built_lib, tmp_path, monkeypatch):
    """ecoz2_vq_learn / ecoz2_vq_quantize / quantize_models with the frames sharded over two GPUs of one
    process: the parallel loader places chunks that straddle the shard boundary on both devices, the
    report text comes in one segment per shard, the per-file distortions are summed over shards --
    all of it must reproduce the one-GPU run."""
    import shutil
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix
    if vq.device_count() < 2:
        pytest.skip("needs two GPUs")
    x = ar_mix(150_001, 36, seed=11)
    cuts = [0, 70_000, 70_001, 150_001]
    files = []
    for k in range(3):
        f = str(tmp_path / "predictors" / "TRAIN" / "m" / f"p{k}.prd")
        formats.write_prd(f, x[cuts[k]:cuts[k + 1]], "m")
        files.append(f)
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("ECOZ2_VQ_MAX_M", "64")
    outs = {}
    for gpus in ("0", "0,1"):
        monkeypatch.setenv("ECOZ2_VQ_GPUS", gpus)
        assert vq.vq_learn(36, 0.05, "m", files) == 0
        cb = tmp_path / "data" / "codebooks" / "m" / "eps_0.05_M_0064.cbook"
        assert vq.vq_quantize(str(cb), files) == 0
        res = {}
        for p_ in sorted((tmp_path / "data").rglob("*")):
            if p_.is_file():
                res[str(p_.relative_to(tmp_path))] = p_.read_bytes()
        outs[gpus] = res
        shutil.rmtree(tmp_path / "data")
    a, b = outs["0"], outs["0,1"]
    assert sorted(a) == sorted(b)
    for name in a:
        if name.endswith(".seq"):
            assert a[name] == b[name], name
        elif name.endswith(".cbook"):
            ra, rb = np.frombuffer(a[name][120:], "<f8"), np.frombuffer(b[name][120:], "<f8")
            assert np.allclose(ra, rb, rtol=0, atol=1e-10), name
        elif name.endswith(".min_dists.csv"):
            ta = np.loadtxt(__import__("io").BytesIO(a[name]), delimiter=",", skiprows=1)
            tb = np.loadtxt(__import__("io").BytesIO(b[name]), delimiter=",", skiprows=1)
            assert ta.shape == tb.shape == (150_001, 2) and np.count_nonzero(ta[:, 0] != tb[:, 0]) <= 2, name
    # several codebooks at once on both devices
    refl = np.frombuffer(a["data/codebooks/m/eps_0.05_M_0064.cbook"][120:], "<f8").reshape(64, 37)   # 120-byte header
    models = np.stack([refl, np.clip(refl * 1.01, -0.999, 0.999), np.clip(refl * 0.98, -0.999, 0.999)])
    models[:, :, 0] = 0.0
    got = []
    for devs in ((0,), (0, 1)):
        with vq.VqEngine(36, max_M=64, devices=devs) as eng:
            eng.set_frames(x)
            seq, dists, one = eng.quantize_models(models, cuts)
            assert one
            got.append((seq, dists))
    assert np.array_equal(got[0][0], got[1][0])
    assert np.allclose(got[0][1], got[1][1], rtol=1e-12, atol=1e-14)
