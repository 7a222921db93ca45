"""CPU suite: the data-parallel gradient exchange of the training path with
world_size-2 gloo processes -- one flat all-reduce and the per-bucket variant, averaged over the ranks -- and the bucket
layout arithmetic (decoder, processor blocks 9..0, encoders cover the 1 591 299 parameters exactly once)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def epd_buckets(node_dim=25, edge_dim=4, out_dim=3, h=128, m_steps=10):
    """The bucket layout gm_train_bucket reports (csrc/train.cu layout_model), restated on the host."""
    def mlp(in1, out3, norm):
        return h * in1 + h + h * h + h + out3 * h + out3 + (2 * out3 if norm else 0)
    enc = mlp(edge_dim, h, True) + mlp(node_dim, h, True)
    block = mlp(3 * h, h, True) + mlp(2 * h, h, True)
    dec = mlp(h, out_dim, False)
    total = enc + m_steps * block + dec
    buckets = [(total - dec, dec)]
    for s in range(m_steps - 1, -1, -1):
        buckets.append((enc + s * block, block))
    buckets.append((0, enc))
    return buckets, total


def test_bucket_layout_partitions_the_parameters():
    buckets, total = epd_buckets()
    assert total == 1591299 and len(buckets) == 12
    covered = sorted(buckets)
    assert covered[0][0] == 0 and covered[-1][0] + covered[-1][1] == total
    for (a, n), (b, _) in zip(covered, covered[1:]):
        assert a + n == b
    assert len({n for _, n in buckets[1:11]}) == 1


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gnn_manip_b200.train_dyn import average_gradients
    buckets, total = epd_buckets(m_steps=2)
    g = torch.Generator().manual_seed(rank)
    local = torch.randn((total,), generator=g)
    one = average_gradients(local.clone())
    per_bucket = average_gradients(local.clone(), buckets)
    # numpy arrays travel by value; torch tensors would be handed over as file descriptors, which needs this process alive
    q.put((rank, local.numpy(), one.numpy(), per_bucket.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_average_gradients_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        rank, local, one, per_bucket = q.get(timeout=180)
        got[rank] = (torch.from_numpy(local), torch.from_numpy(one), torch.from_numpy(per_bucket))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    want = (got[0][0] + got[1][0]) / 2
    for r in range(2):
        assert torch.equal(got[r][1], want) and torch.equal(got[r][2], want)
    assert torch.equal(got[0][1], got[1][1])          # every rank holds the same averaged gradient: Adam stays in lockstep


def test_average_gradients_single_process_is_identity():
    from gnn_manip_b200.train_dyn import average_gradients
    g = torch.arange(10, dtype=torch.float32)
    assert torch.equal(average_gradients(g.clone()), g)
