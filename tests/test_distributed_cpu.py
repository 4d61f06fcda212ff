"""world_size-2 `gloo` test (CPU) of the multi-GPU host logic: batch sharding and the single
all-reduce of the loss / gradient bucket per step.  The shard-local loss is the
CPU oracle here -- this checks the reduction algebra the N>1 bench path relies on: with equal
shards, the all-reduced mean of shard losses and the /world scaled pose gradients equal the
single-process result on the whole batch."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import photometric_oracle as O

HP = dict(num_scales=2, ssim_loss_weight=0.85, smooth_loss_weight=0.001, photometric_reduce_op='min',
          clip_loss=0.0, padding_mode='zeros', automask_loss=True)


def _loss_and_grads(image, context, invd, vecs, K):
    invd = [d.clone().requires_grad_() for d in invd]
    vecs = [v.clone().requires_grad_() for v in vecs]
    out = O.multiview_photometric_loss(image, context, invd, K, K, [O.pose_vec_to_mat(v) for v in vecs], **HP)
    out['loss'].backward()
    return out['loss'].detach(), [d.grad for d in invd], [v.grad for v in vecs]


def _inputs():
    image, context, invd, vecs, K = O.synth_S3(4, 16, 24, 2, torch.float64, kd=5, n=2)
    return image, context, [d.detach() for d in invd], [v.detach() for v in vecs], K


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    image, context, invd, vecs, K = _inputs()
    B = image.shape[0]
    sl = slice(rank * B // world, (rank + 1) * B // world)              # batch shard of this rank
    loss, d_invd, d_vecs = _loss_and_grads(image[sl], [c[sl] for c in context], [d[sl] for d in invd],
                                           [v[sl] for v in vecs], K[sl])
    # one flat bucket per step: [loss | "parameter" gradients]; here a shared pose bias stands in for the
    # pose-net parameters, its gradient is the sum over the shard's frames
    bucket = torch.cat([loss.view(1)] + [g.sum(0) for g in d_vecs])
    dist.all_reduce(bucket, op=dist.ReduceOp.SUM)
    bucket /= world
    # numpy arrays travel by value: a tensor in a Queue is a shared-memory handle the parent may fetch after this
    # process has exited (ConnectionResetError)
    q.put((rank, bucket.numpy().copy(), [g.detach().numpy().copy() for g in d_invd]))
    dist.barrier()
    dist.destroy_process_group()


def test_batch_sharded_loss_allreduce_world2():
    world = 2
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=240) for _ in range(world)], key=lambda r: r[0])
    res = [(r, torch.from_numpy(b), [torch.from_numpy(g) for g in gs]) for r, b, gs in res]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    image, context, invd, vecs, K = _inputs()
    loss, d_invd, d_vecs = _loss_and_grads(image, context, invd, vecs, K)
    want = torch.cat([loss.view(1)] + [g.sum(0) for g in d_vecs])
    for rank, bucket, shard_grads in res:
        torch.testing.assert_close(bucket, want, rtol=1e-10, atol=1e-14)   # every rank holds the global values
        B = image.shape[0]
        sl = slice(rank * B // world, (rank + 1) * B // world)
        for g, full in zip(shard_grads, d_invd):
            # activation gradients stay rank-local; the shard mean is over B/world frames
            torch.testing.assert_close(g / world, full[sl], rtol=1e-10, atol=1e-16)
