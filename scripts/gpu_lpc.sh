#!/bin/bash
# LPC front end: parity tests, timing (fast and generic autocorrelation), ncu launch list + full capture
TAG=${1:-r01g}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_lpc.py -x -q > $O/pytest_lpc_$TAG.log 2>&1
echo "pytest exit $?" >> $O/pytest_lpc_$TAG.log
timeout 300 python scripts/lpc_probe.py --seconds 3600 > $O/lpc_probe_$TAG.log 2>&1
timeout 300 python scripts/lpc_probe.py --seconds 3600 --p 12 >> $O/lpc_probe_$TAG.log 2>&1
timeout 300 python scripts/lpc_probe.py --seconds 3600 --p 20 --win 20 --off 5 >> $O/lpc_probe_$TAG.log 2>&1
ECOZ2_LPC_GENERIC=1 timeout 300 python scripts/lpc_probe.py --seconds 3600 >> $O/lpc_probe_$TAG.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"lpc_" -c 40 --csv --log-file $O/lpc_launches_$TAG.csv \
    python scripts/lpc_probe.py --seconds 3600 --reps 2 > $O/ncu_lpc_launches_$TAG.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"lpc_autocorr" --launch-skip 2 --launch-count 1 \
    -o $O/lpc_$TAG -f python scripts/lpc_probe.py --seconds 3600 --reps 2 > $O/ncu_lpc_full_$TAG.log 2>&1
cat $O/pytest_lpc_$TAG.log $O/lpc_probe_$TAG.log
