# GPU tests + metrics micro-bench + one default bench line (gpurun development call)
mkdir -p gpurun_out
python -m pytest tests/test_metrics_gpu.py -x -q 2>&1 | tail -15
python tools/bench_metrics.py > gpurun_out/metrics_bench.jsonl 2> gpurun_out/metrics_bench.err; tail -3 gpurun_out/metrics_bench.err; cat gpurun_out/metrics_bench.jsonl
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; cat gpurun_out/bench_default.json
