import numpy as np, time, os, sys
sys.path.insert(0,'.')
from ecoz2rs_b200 import vq
n=10_000_000
rng=np.random.default_rng(1)
md=rng.random(n); cw=rng.integers(0,2048,n).astype(np.int32)
print("cpus", os.cpu_count(), len(os.sched_getaffinity(0)))
for path in ('/dev/null','/dev/shm/x.csv','/tmp/x.csv'):
    for rep in range(2):
        t0=time.time(); vq.write_min_dists_csv(path, cw, md); print(path, f"{(time.time()-t0)*1e3:.1f} ms")
