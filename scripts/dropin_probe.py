"""Dev probe: wall time of ecoz2_vq_learn on .prd shards from tmpfs, with / without min_dists.csv."""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402

bench.capture_stdout()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
stop = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
out = bench._dropin_cfg2(torch.device("cuda", 0), n, stop)
bench.emit_line(out)
