#!/bin/bash
# Same-box A/B of library variants (developer tool): AB_VARIANTS="base x y base" bash tools/ab_variants.sh
# Each name but "base" is gaussian-splatterer_b200/libgsb200_<name>.so (built out of tree with make EXTRA=-D...; the .so files
# travel with gpurun); one short bench.py run per variant, one line per variant:
#   name  views/s  render_bwd  render_fwd  preprocess  emit  tile_sort  preprocess_bwd (ms per launch, stage events)
# A name of the form env:VAR=VALUE runs the base library with that environment variable set (GSB_LPT=0, GSB_FWD_STREAMS=1).
for v in ${AB_VARIANTS:-base}; do
  L=gaussian-splatterer_b200/libgsb200_$v.so; [ $v = base ] && L=gaussian-splatterer_b200/libgsb200.so
  E=""; case $v in env:*) E=${v#env:}; L=gaussian-splatterer_b200/libgsb200.so; v=$(echo $E | tr '=' '_');; esac
  env $E GSB_LIB=$PWD/$L python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python -c "
import json
d=json.load(open('gpurun_out/ab_$v.json')); s=d['stages']
print('$v', round(d['value'],1), *[s[k]['ms_per_launch'] for k in ('render_bwd','render_fwd','preprocess','emit','tile_sort','preprocess_bwd')])"
done
