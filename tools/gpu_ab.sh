# A/B of library variants on one box: tests with the default, then a bench line + ncu key metrics per variant
# usage: bash tools/gpu_ab.sh <suffix> [<suffix> ...]   ('' = default library)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in "$@"; do
  lib=$PWD/packnet-sfm-resnet-san_b200/lib/libpsfm_b200$v.so
  PSFM_B200_LIB=$lib python bench.py --steps 100 --warmup 10 --no-cpu > gpurun_out/ab$v.json 2> gpurun_out/ab$v.err
  python -c "
import json; d=json.load(open('gpurun_out/ab$v.json')); print('variant[$v]', round(d['value']), round(d['ms_per_step'],4), d['roofline']['all_kernels_ms_per_step'])"
  PSFM_B200_LIB=$lib ncu --set full --clock-control none --import-source on -k regex:"sweep2_kernel|sweep_kernel" -s 4 -c 1 -o gpurun_out/ncu$v python bench.py --steps 3 --warmup 3 --no-cpu --no-graph > /dev/null 2> gpurun_out/ncu$v.err
  ncu -i gpurun_out/ncu$v.ncu-rep --page raw --csv > gpurun_out/ncu${v}_raw.csv
done
