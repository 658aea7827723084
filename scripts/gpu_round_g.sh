#!/bin/bash
# Round-1g GPU call: full-size configs 4/5, bench with the lpc_front key, LPC timing + ncu of the fast kernel
TAG=${1:-r01g}
O=gpurun_out
mkdir -p $O
timeout 900 python scripts/configs_probe.py --cfg 5 > $O/cfg5_$TAG.json 2> $O/cfg5_$TAG.err
echo "cfg5 exit $?" >> $O/cfg5_$TAG.err
timeout 900 python scripts/configs_probe.py --cfg 4 > $O/cfg4_$TAG.json 2> $O/cfg4_$TAG.err
echo "cfg4 exit $?" >> $O/cfg4_$TAG.err
timeout 900 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err
echo "bench exit $?" >> $O/bench_$TAG.err
bash scripts/gpu_lpc.sh $TAG > /dev/null 2>&1
tail -3 $O/cfg5_$TAG.err $O/cfg4_$TAG.err $O/bench_$TAG.err; cat $O/cfg5_$TAG.json $O/cfg4_$TAG.json; cat $O/lpc_probe_$TAG.log
