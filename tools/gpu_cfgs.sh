# per-config bench lines with both sweep kernels (one B200)
mkdir -p gpurun_out
for c in 2 3 4 5; do for k in 0 1; do
  PSFM_SWEEP2=$k python bench.py --config $c --steps 40 --warmup 5 --no-cpu > gpurun_out/cfg${c}_sw${k}.json 2> gpurun_out/cfg${c}_sw${k}.err
  python -c "
import json; d=json.load(open('gpurun_out/cfg${c}_sw${k}.json')); print('config $c sweep2=$k', round(d['value']), 'frames/s', round(d['ms_per_step'],4), 'ms', {a:round(b,4) for a,b in d['roofline']['all_kernels_ms_per_step'].items()}, 'frac', round(d['roofline']['frac'],3))"
done; done
