"""How the path shards over ranks (one process per GPU).  Pure index arithmetic plus
the one collective of the render path; used by bench.py (NCCL) and tests/test_multirank.py (gloo)."""


def ray_shard(n, rank, world):
    """Contiguous ray range [a, b) of rank `rank`: traversal batches shard by ray, no collective."""
    base, rem = divmod(n, world)
    a = rank * base + min(rank, rem)
    return a, a + base + (1 if rank < rem else 0)


def sample_stripe(spp, rank, world):
    """(first_sample, n_samples, stride) of rank `rank`: sample indices rank, rank + world, ... < spp.
    The union over ranks is every index in [0, spp) exactly once, so the reduced film is the
    single-GPU film up to fp64 summation order."""
    n = (spp - rank + world - 1) // world if rank < spp else 0
    return rank, n, world


def reduce_film(film, dist, dst=0):
    """Sum the per-rank film accumulators {rgbSum[3], weightSum} (fp64) onto rank `dst` - the frame's
    single collective (ncclReduce over NVLink under NCCL; plain reduce under gloo)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film
