# 1/2/4/8-GPU bench lines of the final build on one box, plus the byte-frame variants at 8 GPUs
mkdir -p gpurun_out/scale
for n in 1 2 4 8; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu > gpurun_out/scale/n1.json 2> gpurun_out/scale/n1.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --warmup 10 --no-cpu > gpurun_out/scale/n$n.json 2> gpurun_out/scale/n$n.err; fi
done
for n in 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 100 --warmup 10 --no-cpu --host-images u8 > gpurun_out/scale/n${n}_u8.json 2> gpurun_out/scale/n${n}_u8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n --steps 100 --warmup 10 --no-cpu --host-images u8 --pyramid > gpurun_out/scale/n${n}_u8_pyramid.json 2> gpurun_out/scale/n${n}_u8_pyramid.err
done
nvidia-smi topo -m > gpurun_out/scale/topo.txt 2>&1
for f in n1 n2 n4 n8 n4_u8 n4_u8_pyramid n8_u8 n8_u8_pyramid; do python -c "
import json; d=json.load(open('gpurun_out/scale/$f.json')); print('$f', d['n_gpus'], round(d['value']), round(d['ms_per_step'],4), round(d['e2e']['value']), d['e2e']['h2d_bytes_per_step'], d['config'].get('numa_bound_cores'))"; done
