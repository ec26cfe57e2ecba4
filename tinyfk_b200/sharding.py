"""Batch partitioning across GPUs / ranks.  Configurations are independent, so the only rule is the
contiguous split: rank g of G owns rows [g*ceil(N/G), min(N, (g+1)*ceil(N/G))) of q and
the matching rows of P / J -- the same rule Engine::solve_host applies to its device list.  No collective
touches the data path; ranks only agree on the clock (max over ranks) when a rate is reported."""


def shard_bounds(n, world, rank):
    per = (n + world - 1) // world
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def aggregate_rate(n_local, ms_local, world, all_reduce_max=None, all_reduce_sum=None):
    """whole-job rate in items/s: total items over the SLOWEST rank's time."""
    ms = all_reduce_max(ms_local) if all_reduce_max else ms_local
    n = all_reduce_sum(n_local) if all_reduce_sum else n_local * world
    return n / (ms / 1e3)
