"""CPU, world_size 2 over gloo: the host-side multi-process logic — scene sharding without a collective, the
loss normaliser all-reduce (`reduce_mean`, fcaf3d_neck_with_head.py:179,187) and gradient averaging of the DDP wrap."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fcaf3d_b200.detector import reduce_mean
    from fcaf3d_b200.train import shard_scenes
    shard = shard_scenes(11, rank, world)
    n_pos = torch.tensor(float(10 + 20 * rank))
    mean = reduce_mean(n_pos)
    vec = reduce_mean(torch.tensor([1.0, 4.0 * rank, 7.0]))      # the batched loss reduces one [scenes] vector
    # DDP semantics on a tiny torch module: the averaged gradient equals the single-process gradient of the mean loss
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 1)
    ddp = torch.nn.parallel.DistributedDataParallel(lin)
    x = torch.arange(8, dtype=torch.float32).reshape(2, 4) + rank
    ddp(x).sum().backward()
    # plain python objects only: a tensor in the queue is passed by file descriptor and races with the worker's exit
    out.put((rank, shard, mean.item(), n_pos.item(), lin.weight.grad.reshape(-1).tolist(), vec.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_process_host_logic():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    shards = [r[1] for r in res]
    assert sorted(shards[0] + shards[1]) == list(range(11)) and not set(shards[0]) & set(shards[1])
    assert res[0][2] == res[1][2] == pytest.approx(20.0)       # mean of 10 and 30
    assert res[0][3] == 10.0                                    # the caller's tensor is not modified in place
    want = (torch.arange(8, dtype=torch.float32).reshape(2, 4).sum(0) + (torch.arange(8.).reshape(2, 4) + 1).sum(0)) / 2
    assert torch.allclose(torch.tensor(res[0][4]), want) and torch.allclose(torch.tensor(res[1][4]), want)
    assert res[0][5] == res[1][5] == [1.0, 2.0, 7.0]


def test_single_process_reduce_mean_is_identity():
    from fcaf3d_b200.detector import reduce_mean
    t = torch.tensor(3.0)
    assert reduce_mean(t) is t
