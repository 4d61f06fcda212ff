# tests + one bench line (used with gpurun during development)
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 100 --warmup 10 --no-cpu "$@" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), 'frames/s', round(d['ms_per_step'],4), 'ms/step graph', d['config']['cuda_graph'], 'e2e', round(d['e2e']['value']), d['roofline']['all_kernels_ms_per_step'], 'frac', round(d['roofline']['frac'],4))"
