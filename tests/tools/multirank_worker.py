"""One rank of the multi-process deployment (one process per GPU, `ecoz2b200_comm_init` with a shared
ncclUniqueId -- what bench.py does under torchrun), without torch: the id travels through a file.

    python multirank_worker.py <rank> <world> <frames.npy> <workdir> <stop_M>

Takes the frames [lo, hi) the library's sharding rule gives this rank, runs the LBG learn 2 -> stop_M,
one more k-means pass from HOST frames, inertia and quantize, and writes everything to
<workdir>/rank<r>.npz.  Used by tests/test_gpu_multirank.py (needs >= `world` GPUs)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(sys.argv[1]), int(sys.argv[2])
    frames_path, workdir, stop_M = sys.argv[3], sys.argv[4], int(sys.argv[5])
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.sharding import shard_range
    x = np.load(frames_path, mmap_mode="r")
    N, D = x.shape
    lo, hi = shard_range(N, rank, world)
    mine = np.ascontiguousarray(x[lo:hi])
    id_path = os.path.join(workdir, "nccl_id.bin")
    if rank == 0:
        uid = vq.nccl_unique_id()
        with open(id_path + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(id_path + ".tmp", id_path)
    else:
        t0 = time.time()
        while not os.path.exists(id_path):
            if time.time() - t0 > 120:
                raise SystemExit("no NCCL id from rank 0")
            time.sleep(0.05)
        with open(id_path, "rb") as f:
            uid = f.read()
    dev = rank % vq.device_count()
    with vq.VqEngine(D - 1, max_M=stop_M, devices=(dev,)) as eng:
        eng.comm_init(uid, rank, world)
        eng.set_frames(mine)
        eng.codebook_initial()
        stages, trace = eng.learn(eps=0.05, stop_M=stop_M)
        refl = stages[-1]["refl"]
        # one more pass from the converged codebook: resident frames, then the same from HOST frames
        eng.set_reflections(refl)
        res = eng.kmeans_pass(want_cells=True)
        r_res, _ = eng.codebook()
        eng.set_reflections(refl)
        host = eng.kmeans_pass_host(mine, want_cells=True)
        r_host, _ = eng.codebook()
        eng.set_reflections(refl)
        inert = eng.inertia()
        seq, avg = eng.quantize_resident()
        np.savez(os.path.join(workdir, f"rank{rank}.npz"),
                 lo=lo, hi=hi, M=[s["M"] for s in stages], passes=[s["passes"] for s in stages],
                 avg=[s["avg_distortion"] for s in stages], sigma=[s["sigma"] for s in stages],
                 inertia=[s["inertia"] for s in stages],
                 trace_M=[t[0] for t in trace], trace_pass=[t[1] for t in trace], trace_DD=[t[2] for t in trace],
                 refl=np.concatenate([s["refl"].ravel() for s in stages]),
                 cardd=np.concatenate([s["cardd"] for s in stages]),
                 res_cells=res["cells"], res_cardd=res["cardd"], res_DD=res["DD"], res_refl=r_res,
                 host_cells=host["cells"], host_cardd=host["cardd"], host_DD=host["DD"], host_refl=r_host,
                 inert=inert, seq=seq, q_avg=avg)


if __name__ == "__main__":
    main()
