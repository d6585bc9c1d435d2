"""Row-block sharding of X across the GPUs of one box (host-side logic, no CUDA).

Mirrors the reference's distributed plan for tsmm / mmchain: X is cut into row blocks of
OptimizerUtils.DEFAULT_BLOCKSIZE = 1000 rows (J/hops/OptimizerUtils.java:75), every worker maps its
blocks locally and the m-sized (or m x m) partial results are summed (TsmmSPInstruction /
MapmmChainSPInstruction + RDDAggregateUtils.sumStable, J/runtime/instructions/spark/utils/
RDDAggregateUtils.java:58-70).  Here a worker is one GPU/rank and the sum is an all-reduce.
"""
DEFAULT_BLOCKSIZE = 1000


def row_shard(n_rows, world_size, rank, blocksize=DEFAULT_BLOCKSIZE):
    """Contiguous [start, stop) row range of `rank`: whole 1000-row blocks, spread as evenly as possible
    (the first `nblocks % world_size` ranks get one block more); the last block may be ragged."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    nblocks = (n_rows + blocksize - 1) // blocksize
    base, rest = divmod(nblocks, world_size)
    first = rank * base + min(rank, rest)
    count = base + (1 if rank < rest else 0)
    start = min(first * blocksize, n_rows)
    stop = min((first + count) * blocksize, n_rows)
    return start, stop


def all_shards(n_rows, world_size, blocksize=DEFAULT_BLOCKSIZE):
    return [row_shard(n_rows, world_size, r, blocksize) for r in range(world_size)]
