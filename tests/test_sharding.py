"""CPU suite: the N>1 host path (population sharding) with world_size-2 gloo processes."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnn_manip_b200.sharding import gather_population_values, partition_population  # noqa: E402


@pytest.mark.parametrize("pop,world", [(64, 1), (64, 2), (64, 8), (10, 4), (3, 8), (0, 2)])
def test_partition_covers_population_once(pop, world):
    spans = [partition_population(pop, world, r) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == pop
    for (a, b), (c, d) in zip(spans, spans[1:]):
        assert b == c and a <= b and c <= d
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_partition_rejects_bad_arguments():
    with pytest.raises(ValueError):
        partition_population(8, 2, 2)
    with pytest.raises(ValueError):
        partition_population(8, 0, 0)


def _worker(rank, world, port, pop, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = partition_population(pop, world, rank)
    local = torch.arange(lo, hi, dtype=torch.float64) * 1.5 + 0.25      # "objective" of candidate i
    full = gather_population_values(local, pop)
    q.put((rank, full.tolist()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("pop", [64, 7])
def test_gather_population_values_gloo_world2(pop):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + pop
    procs = [ctx.Process(target=_worker, args=(r, 2, port, pop, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    want = [i * 1.5 + 0.25 for i in range(pop)]
    assert got[0] == want and got[1] == want


def test_gather_single_process_is_identity():
    v = torch.arange(5, dtype=torch.float32)
    assert torch.equal(gather_population_values(v, 5), v)
