"""CPU, world_size 2, gloo: the N>1 path's host logic — block sharding and the single flattened
gradient all-reduce — exercised with the S3DIS tower running on the CPU oracle layers (the CUDA ops
need a GPU; the distributed plumbing does not)."""
from __future__ import annotations

import os
import socket
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from pointcloudsegmentation_b200.dist import GradAverager, broadcast_parameters, shard_blocks  # noqa: E402


def test_shard_blocks_balanced_and_complete():
    costs = [100, 90, 10, 10, 10, 80, 5]
    sh = shard_blocks(costs, 3)
    assert sorted(i for s in sh for i in s) == list(range(7))
    loads = [sum(costs[i] for i in s) for s in sh]
    assert max(loads) - min(loads) <= 25
    # fewer blocks than ranks: padded by repetition (provider.py:25-40), never an empty rank
    sh = shard_blocks([3.0, 1.0], 4)
    assert all(len(s) == 1 for s in sh) and {i for s in sh for i in s} == {0, 1}
    assert shard_blocks([], 2) == [[], []]


def _block(seed, pn=700):
    from pointcloudsegmentation_b200.synthetic import s3dis_block

    rng = np.random.default_rng(seed)
    x = s3dis_block(rng, pn)
    return (torch.from_numpy(x), torch.from_numpy(rng.uniform(0, 1, (pn, 3)).astype(np.float32)),
            torch.from_numpy(rng.integers(0, 13, pn)))


def _model():
    from oracle import cpu_layers
    from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed

    torch.manual_seed(0)
    m = Pointnet13DilatedEmbed(L=cpu_layers.layers())
    m.eval()  # no dropout: deterministic
    with torch.no_grad():
        m(*_block(99))
    return m


def _grads(m, seed):
    m.zero_grad(set_to_none=True)
    loss, _, _ = m(*_block(seed))
    loss.backward()
    return [None if p.grad is None else p.grad.clone() for p in m.parameters()]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    m = _model()
    if rank == 1:  # ranks start different; broadcast must make them equal
        with torch.no_grad():
            for p in m.parameters():
                p.add_(1.0)
    broadcast_parameters(m.parameters())
    avg = GradAverager(m.parameters())
    m.zero_grad(set_to_none=True)
    loss, _, _ = m(*_block(rank))
    loss.backward()
    avg()
    q.put((rank, [p.grad.clone().numpy() for p in m.parameters()]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_mean_matches_serial():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=600) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    m = _model()
    g0, g1 = _grads(m, 0), _grads(m, 1)
    for a, b, r0, r1 in zip(g0, g1, got[0], got[1]):
        want = (a + b).numpy() / 2
        assert np.allclose(r0, want, rtol=1e-5, atol=1e-7)
        assert np.array_equal(r0, r1)  # both ranks hold the same averaged gradient
