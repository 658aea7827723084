#!/bin/bash
# Round-1f GPU call: GPU parity tests (VQ + classify + LPC front end), smoke, both bench arms,
# LPC front-end timing + ncu launch list and --set full capture of its kernels.
TAG=${1:-r01f}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi_$TAG.txt 2>&1
nproc >> $O/smi_$TAG.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1
echo "pytest exit $?" >> $O/pytest_gpu_$TAG.log
timeout 300 python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1
echo "smoke exit $?" >> $O/smoke_$TAG.log
timeout 900 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err
echo "bench exit $?" >> $O/bench_$TAG.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err
echo "bench ref exit $?" >> $O/bench_ref_$TAG.err
timeout 300 python scripts/lpc_probe.py --seconds 3600 > $O/lpc_probe_$TAG.log 2>&1
timeout 300 python scripts/lpc_probe.py --seconds 3600 --p 12 >> $O/lpc_probe_$TAG.log 2>&1
timeout 300 python scripts/lpc_probe.py --seconds 3600 --p 20 --win 20 --off 5 >> $O/lpc_probe_$TAG.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"lpc_" -c 40 --csv --log-file $O/lpc_launches_$TAG.csv \
    python scripts/lpc_probe.py --seconds 3600 --reps 2 > $O/ncu_lpc_launches_$TAG.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"lpc_" --launch-skip 6 --launch-count 2 \
    -o $O/lpc_$TAG -f python scripts/lpc_probe.py --seconds 3600 --reps 2 > $O/ncu_lpc_full_$TAG.log 2>&1
ls -la $O
