"""Path sharding for multi-GPU runs: one process per GPU, each owning a contiguous block of simulation paths.

The LSMC path shards by simulation path: the simulator is counter-based (Philox keyed by the GLOBAL
path index, so a shard reproduces exactly the paths the single-process run would have produced), the decision sweep
and the forward pass are independent per path, and the only exchange is the all-reduce of the per-period
normal-equation partials and of the forward sums (done inside liblsmc_b200.so over NCCL).  This module holds the
host-side arithmetic: which paths a rank owns, and how to attach that to a problem description.
"""
from __future__ import annotations

import copy
from dataclasses import dataclass


@dataclass(frozen=True)
class PathShard:
    rank: int
    world: int
    n_global: int
    offset: int     # first global path index of this rank
    count: int      # number of paths of this rank


def shard_paths(n_global: int, world: int, rank: int) -> PathShard:
    """Contiguous, near-equal shards whose boundaries fall on EVEN path indices.

    Antithetic sampling pairs path 2i with path 2i+1 (lsmc_rng.antithetic), and the standard error is computed over pair
    means, so a pair must never be split across ranks: every shard starts at an even index and, except possibly the last
    one when n_global is odd, holds an even number of paths.
    """
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %d outside world of %d" % (rank, world))
    if n_global < 0:
        raise ValueError("n_global must be non-negative")
    pairs = (n_global + 1) // 2
    base, rem = divmod(pairs, world)
    first_pair = rank * base + min(rank, rem)
    n_pairs = base + (1 if rank < rem else 0)
    offset = min(2 * first_pair, n_global)
    count = min(2 * (first_pair + n_pairs), n_global) - offset
    return PathShard(rank, world, n_global, offset, count)


def shard_problem(problem, world: int, rank: int):
    """A copy of `problem` (storage_b200._abi.Problem) restricted to this rank's paths: n_sims = local count,
    n_sims_global / sim_offset as include/lsmc_b200.h documents them.  `problem.n_sims` is the global path count."""
    sh = shard_paths(int(problem.n_sims), world, rank)
    local = copy.copy(problem)
    local.n_sims = sh.count
    local.n_sims_global = sh.n_global
    local.sim_offset = sh.offset
    return local, sh
