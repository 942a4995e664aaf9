for v in ${AB_VARIANTS:-base}; do
  L=gaussian-splatterer_b200/libgsb200_$v.so; [ $v = base ] && L=gaussian-splatterer_b200/libgsb200.so
  GSB_LIB=$PWD/$L python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python -c "
import json,sys
d=json.load(open('gpurun_out/ab_$v.json')); print('$v', round(d['value'],1), d['stages']['render_bwd']['ms_per_launch'], d['stages']['render_fwd']['ms_per_launch'])"
done
