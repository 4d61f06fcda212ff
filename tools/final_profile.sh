# end-of-round measurement set: bench lines (default / variants), launch list and one full ncu capture of the
# fisheye sweep.  Each ncu pass runs only after the plain command has exited 0.
set -e
mkdir -p gpurun_out/final
python bench.py --steps 200 --warmup 20 > gpurun_out/final/bench_default.json 2> gpurun_out/final/bench_default.err
python bench.py --steps 200 --warmup 20 --no-cpu --pyramid > gpurun_out/final/bench_pyramid.json 2>/dev/null
python bench.py --steps 200 --warmup 20 --no-cpu --host-images u8 > gpurun_out/final/bench_u8.json 2>/dev/null
python bench.py --steps 200 --warmup 20 --no-cpu --host-images u8 --pyramid > gpurun_out/final/bench_u8_pyramid.json 2>/dev/null
python bench.py --steps 200 --warmup 20 --no-cpu --camera fisheye > gpurun_out/final/bench_fisheye.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --camera fisheye > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final/fisheye_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --camera fisheye > gpurun_out/final/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"sweep_kernel|fisheye_ray" -s 4 -c 2 -o gpurun_out/final/fisheye_sweep \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-graph --camera fisheye > gpurun_out/final/ncu2.log 2>&1
ncu -i gpurun_out/final/fisheye_sweep.ncu-rep --page details > gpurun_out/final/fisheye_sweep_details.txt 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final/default_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/final/ncu3.log 2>&1
for f in default pyramid u8 u8_pyramid fisheye; do python -c "
import json; d=json.load(open('gpurun_out/final/bench_$f.json')); print('$f', round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['e2e']['h2d_bytes_per_step'], d['roofline']['frac'], d['clocks'])"; done
