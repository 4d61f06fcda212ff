mkdir -p gpurun_out
python -m pytest tests/test_metrics_gpu.py -x -q 2>&1 | tail -3
python tools/bench_metrics.py > gpurun_out/metrics_bench.jsonl 2> gpurun_out/metrics_bench.err; tail -3 gpurun_out/metrics_bench.err
python -c "
import json
for l in open('gpurun_out/metrics_bench.jsonl'):
    d=json.loads(l); print(d['shape'], d['pred'], d['use_gt_scale'], 'graph ms', round(d['ms_graph'],4), 'frac', round(d['frac_of_hbm_peak'],3), 'x torch', round(d['speedup_vs_torch_eager'],1))
"
python tools/profile_metrics.py > /dev/null && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/metrics_launches.csv python tools/profile_metrics.py > gpurun_out/ncu_metrics.log 2>&1
tail -12 gpurun_out/metrics_launches.csv | cut -d, -f5,10-
