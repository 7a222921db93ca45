"""Multi-GPU sharding of the CMA-ES population: one process per GPU, the population
index range is split contiguously, every rank rolls out its own candidates as disjoint graphs in one plan, and only
the per-candidate scalars (objective values) travel back -- a control-plane all_gather of popsize floats, no
collective on the data path.  The reference evaluates candidates serially inside cma.fmin2
(gnn_manip/utils/traj_utils.py:257); this is the part that replaces that loop across GPUs."""
import torch


def partition_population(popsize, world_size, rank):
    """Contiguous, balanced partition: rank r owns [lo, hi); sizes differ by at most one, earlier ranks get the extra."""
    if popsize < 0 or world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad partition request popsize={popsize} world_size={world_size} rank={rank}")
    base, extra = divmod(popsize, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_population_values(local_values, popsize, group=None):
    """All ranks receive the [popsize] vector of per-candidate values in population order.
    local_values: 1-D tensor holding this rank's slice (partition_population order)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        if local_values.numel() != popsize:
            raise ValueError("single process must hold the whole population")
        return local_values.clone()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = partition_population(popsize, world, rank)
    if local_values.numel() != hi - lo:
        raise ValueError(f"rank {rank} holds {local_values.numel()} values, its slice has {hi - lo}")
    width = -(-popsize // world)
    pad = torch.zeros(width, dtype=local_values.dtype, device=local_values.device)
    pad[: hi - lo] = local_values
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    parts = []
    for r in range(world):
        a, b = partition_population(popsize, world, r)
        parts.append(out[r][: b - a])
    return torch.cat(parts)
