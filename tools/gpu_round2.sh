#!/bin/bash
# One gpurun call of round 2.  Usage: gpurun --timeout 1800 -- 'bash tools/gpu_round2.sh TAG [sections]'
# sections (default "test bench ref"): test bench ref ncu refncu full sanitize mgpu
TAG=${1:-r02}
shift
SECTIONS=${*:-test bench ref}
OUT=gpurun_out
mkdir -p $OUT
has() { [[ " $SECTIONS " == *" $1 "* ]]; }
NGPU=$(nvidia-smi -L | wc -l)
if has test; then
  python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
  tail -5 $OUT/${TAG}_pytest.log
fi
if has bench; then
  python bench.py --steps 10 --warmup 3 > $OUT/${TAG}_bench_gsb.json 2> $OUT/${TAG}_bench_gsb.err; echo "bench exit $?"
  head -c 1200 $OUT/${TAG}_bench_gsb.json; echo
fi
if has ref; then
  python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref exit $?"
  head -c 600 $OUT/${TAG}_bench_ref.json; echo
fi
BENCH_SMALL="python bench.py --cameras 2 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
if has ncu; then
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file $OUT/${TAG}_launches.csv $BENCH_SMALL > $OUT/${TAG}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
fi
if has refncu; then
  # per-kernel table of the REFERENCE build on the same box: one best-case pass over 2+2 cameras x 2 backgrounds at C4
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file $OUT/${TAG}_ref_launches.csv python bench.py --impl reference --cameras 2 --steps 1 --warmup 1 > $OUT/${TAG}_ncu_ref.log 2>&1; echo "ncu ref exit $?"
fi
if has full; then
  # full capture of the two render kernels in the BENCH configuration (all 64 cameras): roofline.traffic comes from this report
  # (the first 128 render_fwd launches of the process render the truth frames from the 250 k teacher cloud: skipped)
  CMD="ncu --set full --clock-control none --import-source on -k regex:render_(bwd|fwd)_kernel --launch-skip 170 -c 6 -f -o $OUT/${TAG}_prof_render python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
  echo "$CMD" > $OUT/${TAG}_prof_render.cmd
  timeout 900 $CMD > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
fi
if has skeleton; then
  # A/B of VERDICT r01 item 5: shipped fused render backward vs the lanes = list entries skeleton (lower bound), 6 C4 cameras
  timeout 600 python tools/gpu_bwd_skeleton.py C4 3 > $OUT/${TAG}_bwd_skeleton.json 2> $OUT/${TAG}_bwd_skeleton.err; echo "skeleton exit $?"
  head -c 400 $OUT/${TAG}_bwd_skeleton.json; echo
fi
if has staging; then
  # A/B of VERDICT r01 item 8b: record staging by cp.async against the TMA row gather (tile::gather4)
  timeout 300 python tools/gpu_staging_ab.py C4 2 > $OUT/${TAG}_staging_ab.json 2> $OUT/${TAG}_staging_ab.err; echo "staging exit $?"
  head -c 600 $OUT/${TAG}_staging_ab.json; echo; tail -3 $OUT/${TAG}_staging_ab.err
fi
if has stream; then
  # full capture of the streaming kernels (preprocess, binning chain, preprocess backward) at the bench's sizes
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'(preprocess_fwd|preprocess_bwd|emit|tile_sort|tile_scan|apply_gradients)' --launch-skip 40 -c 12 -f -o $OUT/${TAG}_prof_stream $BENCH_SMALL > $OUT/${TAG}_ncu_stream.log 2>&1; echo "ncu stream exit $?"
fi
if has sanitize; then
  timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_job.py > $OUT/${TAG}_memcheck.log 2>&1; echo "memcheck exit $?"
  tail -4 $OUT/${TAG}_memcheck.log
  timeout 1200 compute-sanitizer --tool racecheck --racecheck-report all --error-exitcode 9 python tools/sanitize_job.py > $OUT/${TAG}_racecheck.log 2>&1; echo "racecheck exit $?"
  tail -4 $OUT/${TAG}_racecheck.log
  timeout 900 compute-sanitizer --tool synccheck --error-exitcode 9 python tools/sanitize_job.py > $OUT/${TAG}_synccheck.log 2>&1; echo "synccheck exit $?"
  tail -4 $OUT/${TAG}_synccheck.log
fi
if has mgpu; then
  python -m pytest tests/test_multi_gpu_gpu.py -m gpu -x -q > $OUT/${TAG}_pytest_mgpu.log 2>&1; echo "pytest mgpu exit $?"
  tail -5 $OUT/${TAG}_pytest_mgpu.log
  for N in 2 4 8; do
    if [ $N -le $NGPU ]; then
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 10 --warmup 3 > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err; echo "bench n$N exit $?"
      head -c 400 $OUT/${TAG}_bench_n$N.json; echo
    fi
  done
fi
if has mgputest; then
  # the world-size-2 parity tests only (torchrun inside pytest) + the C++ compat demo on two devices
  python -m pytest tests/test_multi_gpu_gpu.py -m gpu -x -q > $OUT/${TAG}_pytest_mgpu.log 2>&1; echo "pytest mgpu exit $?"
  tail -5 $OUT/${TAG}_pytest_mgpu.log
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > $OUT/${TAG}_bench_n2.json 2> $OUT/${TAG}_bench_n2.err; echo "bench n2 exit $?"
  head -c 400 $OUT/${TAG}_bench_n2.json; echo
fi
if has scale8c; then
  # 8 GPUs, trimmed to what only this lease can show: the bench line at N = 8 (with the correctness job), C5 as written
  # (3 M splats, 128 cameras, densify every 2nd step), the host-link ceiling, compat Trainer on 8 devices
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus $NGPU --steps 8 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_n$NGPU.json 2> $OUT/${TAG}_bench_n$NGPU.err; echo "bench n$NGPU exit $?"
  head -c 300 $OUT/${TAG}_bench_n$NGPU.json; echo
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NGPU --config C5 --densify-every 2 --steps 6 --warmup 3 --no-mgpu-check --no-cpu-baseline > $OUT/${TAG}_bench_c5_n$NGPU.json 2> $OUT/${TAG}_bench_c5_n$NGPU.err; echo "bench c5 exit $?"
  head -c 300 $OUT/${TAG}_bench_c5_n$NGPU.json; echo
  timeout 300 python tools/h2d_probe.py --mb 133 --reps 20 > $OUT/${TAG}_h2d_probe.json 2> $OUT/${TAG}_h2d_probe.err; echo "h2d probe exit $?"
  head -c 600 $OUT/${TAG}_h2d_probe.json; echo
  (cd tests/compat && timeout 300 ./compat_demo --devices $NGPU > ../../$OUT/${TAG}_compat_devices.json 2> ../../$OUT/${TAG}_compat_devices.err; echo "compat devices exit $?")
  head -c 400 $OUT/${TAG}_compat_devices.json; echo
fi
if has scale8; then
  # the 8-GPU evidence of round 2: strong scaling at C4 with the correctness job, C5 with periodic densify, host-link ceiling
  for N in 4 8; do
    if [ $N -le $NGPU ]; then
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 10 --warmup 3 > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err; echo "bench n$N exit $?"
      head -c 300 $OUT/${TAG}_bench_n$N.json; echo
    fi
  done
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NGPU --config C5 --densify-every 2 --steps 6 --warmup 3 --no-mgpu-check > $OUT/${TAG}_bench_c5_n$NGPU.json 2> $OUT/${TAG}_bench_c5_n$NGPU.err; echo "bench c5 exit $?"
  head -c 300 $OUT/${TAG}_bench_c5_n$NGPU.json; echo
  timeout 300 python tools/h2d_probe.py --mb 133 --reps 20 > $OUT/${TAG}_h2d_probe.json 2> $OUT/${TAG}_h2d_probe.err; echo "h2d probe exit $?"
  head -c 600 $OUT/${TAG}_h2d_probe.json; echo
  (cd tests/compat && timeout 300 ./compat_demo --devices $NGPU > ../../$OUT/${TAG}_compat_devices.json 2> ../../$OUT/${TAG}_compat_devices.err; echo "compat devices exit $?")
  head -c 400 $OUT/${TAG}_compat_devices.json; echo
fi
if has scale8b; then
  # peer-window exchange at 8 GPUs: full bench line (with the correctness job + exchange microbench), the NCCL A/B, C5 with densify, compat devices
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus $NGPU --steps 10 --warmup 3 > $OUT/${TAG}_bench_n$NGPU.json 2> $OUT/${TAG}_bench_n$NGPU.err; echo "bench n$NGPU exit $?"
  head -c 300 $OUT/${TAG}_bench_n$NGPU.json; echo
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29529 bench.py --gpus $NGPU --steps 10 --warmup 3 --nccl-exchange --no-mgpu-check --no-cpu-baseline > $OUT/${TAG}_bench_n${NGPU}_nccl.json 2> $OUT/${TAG}_bench_n${NGPU}_nccl.err; echo "bench nccl exit $?"
  head -c 300 $OUT/${TAG}_bench_n${NGPU}_nccl.json; echo
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $NGPU --config C5 --densify-every 2 --steps 6 --warmup 3 --no-mgpu-check --no-cpu-baseline > $OUT/${TAG}_bench_c5_n$NGPU.json 2> $OUT/${TAG}_bench_c5_n$NGPU.err; echo "bench c5 exit $?"
  head -c 300 $OUT/${TAG}_bench_c5_n$NGPU.json; echo
  (cd tests/compat && timeout 300 ./compat_demo --devices $NGPU > ../../$OUT/${TAG}_compat_devices.json 2> ../../$OUT/${TAG}_compat_devices.err; echo "compat devices exit $?")
  head -c 400 $OUT/${TAG}_compat_devices.json; echo
fi
