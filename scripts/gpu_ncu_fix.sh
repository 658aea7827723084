#!/bin/bash
# the three ncu captures of gpu_round.sh whose launch-skip counts changed with the fused small-codebook pass
TAG=${1:-r01j}
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:"vq_assign_tc2|vq_cell_sums" --launch-skip 21 --launch-count 3 \
    -o $O/search_$TAG -f python scripts/perf_probe.py --n 2000000 --ks 2048 --reps 2 > $O/ncu_full_$TAG.log 2>&1
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"vq_assign_tc2" --launch-skip 16 --launch-count 1 --csv --log-file $O/traffic_$TAG.csv \
    python scripts/perf_probe.py --n 10000000 --ks 2048 --reps 1 > $O/ncu_traffic_$TAG.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"vq_pass_small2" --launch-skip 1 --launch-count 1 \
    -o $O/small2_$TAG -f python scripts/perf_probe.py --n 10000000 --ks 2 --reps 2 > $O/ncu_small2_$TAG.log 2>&1
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"vq_pass_small2|vq_small_finish" -c 12 --csv --log-file $O/small2_launches_$TAG.csv \
    python scripts/perf_probe.py --n 10000000 --ks 2,4,8 --reps 1 > $O/ncu_small2_launches_$TAG.log 2>&1
tail -3 $O/ncu_full_$TAG.log $O/ncu_small2_$TAG.log; tail -4 $O/traffic_$TAG.csv | cut -c1-250
