# one 8-GPU box: scaling lines of every multi-GPU configuration BASELINE.json names
bash tools/scale_run.sh --config 2
# config 3: global batch 16, bf16 384x640, strong scaling over 2 / 4 / 8
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --config 3 --global-batch 16 --steps 100 --warmup 10 --no-cpu --no-gpu-eager > gpurun_out/scale/cfg3_strong_n$n.json 2> gpurun_out/scale/cfg3_strong_n$n.err
  python -c "
import json; d=json.load(open('gpurun_out/scale/cfg3_strong_n$n.json')); print('cfg3 strong n$n', round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']))"
done
python bench.py --gpus 1 --config 3 --global-batch 16 --steps 100 --warmup 10 --no-cpu --no-gpu-eager > gpurun_out/scale/cfg3_strong_n1.json 2> gpurun_out/scale/cfg3_strong_n1.err
# config 4 (DDAD, border) and config 5 (C=4, B=32): 1 and 8 GPUs (config 5: the full sweep)
python bench.py --gpus 1 --config 4 --steps 40 --warmup 5 --no-cpu > gpurun_out/scale/cfg4_n1.json 2> gpurun_out/scale/cfg4_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 8 --config 4 --steps 40 --warmup 5 --no-cpu --no-gpu-eager > gpurun_out/scale/cfg4_n8.json 2> gpurun_out/scale/cfg4_n8.err
bash tools/scale_run.sh --config 5
for f in cfg4_n1 cfg4_n8; do python -c "
import json; d=json.load(open('gpurun_out/scale/$f.json')); print('$f', d['n_gpus'], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']))"; done
