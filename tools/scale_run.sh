# 1/2/4/8-GPU bench lines on one box.   bash tools/scale_run.sh [--config K] [extra bench flags]
# writes gpurun_out/scale/cfg<K>_n<N>.json (+ the reference arm at N=1, the bucket line at N=8 and the bare H2D ceiling)
CFG=2
if [ "$1" = "--config" ]; then CFG=$2; shift 2; fi
STEPS=100
[ $CFG = 4 ] && STEPS=40
[ $CFG = 5 ] && STEPS=30
mkdir -p gpurun_out/scale
NG=$(nvidia-smi -L | wc -l)
for n in 1 2 4 8; do
  [ $n -gt $NG ] && continue
  out=gpurun_out/scale/cfg${CFG}_n$n
  if [ $n = 1 ]; then python bench.py --gpus 1 --config $CFG --steps $STEPS --warmup 10 --no-cpu --no-gpu-eager "$@" > $out.json 2> $out.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --config $CFG --steps $STEPS --warmup 10 --no-cpu --no-gpu-eager "$@" > $out.json 2> $out.err; fi
  python -c "
import json; d=json.load(open('$out.json')); print('cfg$CFG n$n', d['n_gpus'], round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), d['run'].get('allreduce'))"
done
if [ $CFG = 2 ]; then
  for n in 1 2 4 8; do
    [ $n -gt $NG ] && continue
    if [ $n = 1 ]; then python tools/h2d_ceiling.py > gpurun_out/scale/h2d_n$n.json 2>/dev/null
    else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n tools/h2d_ceiling.py > gpurun_out/scale/h2d_n$n.json 2>/dev/null
         python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2962$n tools/h2d_ceiling.py --bind > gpurun_out/scale/h2d_bind_n$n.json 2>/dev/null; fi
    cat gpurun_out/scale/h2d_n$n.json
  done
  n=$NG
  [ $n -gt 1 ] && python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus $n --steps 100 --warmup 10 --no-cpu --grad-bucket-mfloats 28 --allreduce async > gpurun_out/scale/cfg2_n${n}_bucket28.json 2> gpurun_out/scale/cfg2_n${n}_bucket28.err
  [ $n -gt 1 ] && python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 100 --warmup 10 --no-cpu --allreduce async > gpurun_out/scale/cfg2_n${n}_async.json 2> gpurun_out/scale/cfg2_n${n}_async.err
fi
nvidia-smi topo -m > gpurun_out/scale/topo.txt 2>&1
