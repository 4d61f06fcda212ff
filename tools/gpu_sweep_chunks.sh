python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/t3_tests.log
tail -3 gpurun_out/t3_tests.log
for ch in 0 32 48 64 96 192; do
  PSFM_SWEEP_CHUNK=$ch python bench.py --steps 50 --warmup 5 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunk $ch', round(d['value']), d['ms_per_step'], d['roofline']['all_kernels_ms_per_step'])"
done
