#!/usr/bin/env python
"""Bare pinned-host -> device copy bandwidth with N ranks copying AT THE SAME TIME (no kernels): the ceiling of the
end-to-end bench line when every step's inputs come from host memory.

    python tools/h2d_ceiling.py                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_ceiling.py

Each rank copies a pinned buffer of --mb MiB (default 51, one config-2 step) --iters times, after a barrier; rank 0
prints one JSON line: per-rank and aggregate GB/s (max time over ranks)."""
import argparse, json, os, sys
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument('--mb', type=float, default=51.0)
ap.add_argument('--iters', type=int, default=200)
ap.add_argument('--bind', action='store_true', help='bind each rank to its slice of the cores next to its GPU (bench.py)')
args = ap.parse_args()
world, rank, local = int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0))
dev = torch.device('cuda', local)
torch.cuda.set_device(dev)
cores = None
if args.bind:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
    import bench
    cores = bench.bind_to_gpu_numa_node(local)
if world > 1:
    dist.init_process_group('nccl', device_id=dev)
n = int(args.mb * 2 ** 20)
src = torch.empty(n, dtype=torch.uint8).pin_memory()
src.fill_(rank + 1)
dst = torch.empty(n, dtype=torch.uint8, device=dev)
for _ in range(10):
    dst.copy_(src, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.iters):
    dst.copy_(src, non_blocking=True)
e1.record()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
ts = [torch.zeros_like(t) for _ in range(world)]
if world > 1:
    dist.all_gather(ts, t)
else:
    ts = [t]
if rank == 0:
    ms = [x.item() for x in ts]
    per = [n * args.iters / (m * 1e-3) / 1e9 for m in ms]
    print(json.dumps({'ranks': world, 'mb_per_copy': args.mb, 'iters': args.iters, 'bound_cores': cores,
                      'per_rank_gbs': [round(p, 2) for p in per],
                      'aggregate_gbs': round(world * n * args.iters / (max(ms) * 1e-3) / 1e9, 2)}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
