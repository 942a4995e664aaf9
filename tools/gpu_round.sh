#!/bin/bash
# One gpurun call: GPU tests, both bench arms, then (only after those exited) the ncu launch list and a full capture
# of the two render kernels.  Usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh TAG'
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -3 $OUT/${TAG}_pytest.log
python bench.py --steps 10 --warmup 3 > $OUT/${TAG}_bench_gsb.json 2> $OUT/${TAG}_bench_gsb.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref exit $?"
BENCH_SMALL="python bench.py --cameras 2 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $OUT/${TAG}_launches.csv $BENCH_SMALL > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'render_(bwd|fwd)_kernel' --launch-skip 8 -c 2 -f -o $OUT/${TAG}_prof_render $BENCH_SMALL > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
head -c 1500 $OUT/${TAG}_bench_gsb.json
