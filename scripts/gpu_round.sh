#!/bin/bash
# One gpurun call: GPU parity tests, smoke, both bench arms, ncu launch list, ncu --set full capture
# of the search and accumulate kernels at K=2048. Everything lands in gpurun_out/.
# usage: scripts/gpu_round.sh <tag>
TAG=${1:-r01}
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
timeout 600 python scripts/perf_probe.py --n 10000000 --ks 2,4,8,16,32,64,128,256,512,1024,2048,4096 > $O/probe_$TAG.log 2>&1
timeout 600 python scripts/perf_probe.py --n 10000000 --ks 64,256,1024,2048,4096 --mode 0 > $O/probe_fp32_$TAG.log 2>&1
timeout 600 python scripts/perf_probe.py --n 20000000 --p 12 --ks 2,16,64,256,1024 > $O/probe_p12_$TAG.log 2>&1
# launch list of the library's own kernels (the torch synth kernels are filtered out by name)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"vq_|ffma" -c 1500 --csv \
    --log-file $O/launches_$TAG.csv python bench.py --frames 2000000 --steps 3 --warmup 3 \
    --no-cpu-baseline --e2e-steps 1 > $O/ncu_launches_$TAG.log 2>&1
# perf_probe --ks 2048: one pass per stage K = 2..1024 (K <= 8 take the fused small-codebook kernel), then K = 2048; per pass at K >= 16 the regex matches
# search, filter, cell sums -> skip 21 lands on the first K = 2048 pass
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:"vq_assign_tc2|vq_cell_sums" --launch-skip 21 --launch-count 3 \
    -o $O/search_$TAG -f python scripts/perf_probe.py --n 2000000 --ks 2048 --reps 2 \
    > $O/ncu_full_$TAG.log 2>&1
# DRAM traffic of the dominant kernel at the bench size (10M frames, K=2048): one launch
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"vq_assign_tc2" --launch-skip 16 --launch-count 1 --csv --log-file $O/traffic_$TAG.csv \
    python scripts/perf_probe.py --n 10000000 --ks 2048 --reps 1 > $O/ncu_traffic_$TAG.log 2>&1
# SM clock seen inside the search kernel (clock64 vs globaltimer), 20 back-to-back passes
ECOZ2_VQ_TC_EXP=256 timeout 300 python scripts/perf_probe.py --n 10000000 --ks 2048 --reps 20 2>&1 \
    | grep -E "tc prof|K=" | tail -4 > $O/clock_$TAG.log
ls -la $O
# fused small-codebook pass (K = 2) and the LPC front end: ncu --set full of one launch each
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"vq_pass_small2" --launch-skip 1 --launch-count 1 \
    -o $O/small2_$TAG -f python scripts/perf_probe.py --n 10000000 --ks 2 --reps 2 > $O/ncu_small2_$TAG.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"lpc_autocorr" --launch-skip 2 --launch-count 1 \
    -o $O/lpc_$TAG -f python scripts/lpc_probe.py --seconds 3600 --reps 2 > $O/ncu_lpc_full_$TAG.log 2>&1
timeout 300 python scripts/lpc_probe.py --seconds 3600 > $O/lpc_probe_$TAG.log 2>&1
ls -la $O
