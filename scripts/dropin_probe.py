"""Dev probe: wall time of the drop-in `ecoz2_vq_learn` (files in, files out) with and without
the per-stage min_dists.csv reports."""
import os, sys, time, tempfile, shutil
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ecoz2rs_b200 import vq, formats
from ecoz2rs_b200.synth import ar_mix

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
maxm = sys.argv[2] if len(sys.argv) > 2 else "2048"
tmp = tempfile.mkdtemp(prefix="dropin_")
try:
    t0 = time.time()
    x = ar_mix(n, 36, seed=5)
    files = formats.write_prd_shards(os.path.join(tmp, "predictors", "TRAIN", "c"), x, "c", shard_frames=1_000_000)
    print(f"wrote {len(files)} prd files in {time.time()-t0:.1f}s", flush=True)
    os.chdir(tmp)
    os.environ["ECOZ2_VQ_MAX_M"] = maxm
    for md in ("1", "0"):
        os.environ["ECOZ2_VQ_MIN_DISTS"] = md
        shutil.rmtree(os.path.join(tmp, "data"), ignore_errors=True)
        fd = os.dup(1); dn = os.open(os.devnull, os.O_WRONLY); os.dup2(dn, 1)
        t0 = time.time()
        rc = vq.vq_learn(36, 0.05, "c", files)
        dt = time.time() - t0
        os.dup2(fd, 1); os.close(fd); os.close(dn)
        print(f"ecoz2_vq_learn n={n} 2->{maxm} MIN_DISTS={md}: rc={rc} wall {dt:.2f}s", flush=True)
finally:
    os.chdir("/")
    shutil.rmtree(tmp, ignore_errors=True)
