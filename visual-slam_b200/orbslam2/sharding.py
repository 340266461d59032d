"""Frame-range sharding of the front end over ranks (one process per GPU).

The reference processes one video sequentially (src/tests/tester_map.py:60-120: extract frame i,
match it against frame i-1).  Extraction is independent per frame and matching needs only the
previous frame's descriptors, so a video shards into contiguous frame ranges with no data-path
collective; the only exchange is one descriptor block per rank boundary (the first frame of rank
r+1 goes to rank r, which owns the pair (last frame of r, first frame of r+1)).

The functions take the torch.distributed module (or any object with the same P2POp / isend / irecv
/ batch_isend_irecv / all_reduce attributes) so that the same code runs over NCCL on GPUs
(bench.py) and over gloo on CPU tensors (tests/test_sharding_gloo.py).
"""


def frame_range(rank, world, total):
    """Contiguous [begin, end) of `total` frames owned by `rank`; the remainder goes to the first ranks."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    base, rem = divmod(int(total), world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def owned_pairs(rank, world, total):
    """Match pairs (i, i+1) owned by `rank`: the pairs inside its range plus the one across its upper boundary."""
    b, e = frame_range(rank, world, total)
    last = e if (rank < world - 1 and e < total) else e - 1
    return [(i, i + 1) for i in range(b, last) if i + 1 < total]


def boundary_exchange(dist, rank, world, first_desc, first_count, nb_desc, nb_count):
    """Post the boundary transfers: this rank's first-frame descriptors/count go to rank-1, the
    neighbour's arrive in nb_desc/nb_count from rank+1.  Returns the outstanding requests."""
    if world == 1:
        return []
    ops = []
    if rank > 0:
        ops += [dist.P2POp(dist.isend, first_desc, rank - 1), dist.P2POp(dist.isend, first_count, rank - 1)]
    if rank < world - 1:
        ops += [dist.P2POp(dist.irecv, nb_desc, rank + 1), dist.P2POp(dist.irecv, nb_count, rank + 1)]
    return dist.batch_isend_irecv(ops) if ops else []


def max_over_ranks(dist, world, tensor):
    """In-place MAX all-reduce of a 1-element timing tensor (bench.py times a step as the slowest rank)."""
    if world > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.MAX)
    return tensor
