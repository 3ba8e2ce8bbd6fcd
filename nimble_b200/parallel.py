"""Multi-GPU execution: one process per GPU (torch.distributed), two ways to split the work.

chain sharding  rank r owns chains [r*K/R, (r+1)*K/R) and a full replica of the shared arrays; nothing is exchanged
                while sampling, mvSamples are gathered once at the end.  The draw stream is keyed by the GLOBAL chain
                index, so the result is bit-identical to a single-process run of all K chains.
data sharding   every rank holds all chains but only a contiguous slice of each large data declaration; each
                per-proposal block sum (and each conjugate contribution vector) is all-reduced over NCCL inside the
                device library (nimb200_enable_data_sharding), after which every rank takes the same decision.

The reference has no counterpart: runMCMC loops over chains in one process (P/R/MCMC_run.R:127-153) and the manual
recommends one separately compiled model per worker (UserManual/src/chapter_MCMC.Rmd:1223-1229)."""
import numpy as np


def chain_range(nchains, rank, world):
    """Chains owned by `rank`: contiguous, sizes differ by at most one."""
    lo = (nchains * rank) // world
    hi = (nchains * (rank + 1)) // world
    return lo, hi


def node_range(n, rank, world, align=128):
    """Node instances [lo, hi) of a sharded declaration owned by `rank` (tile aligned, last rank takes the tail)."""
    per = -(-n // world)
    per = -(-per // align) * align
    lo = min(n, rank * per)
    hi = min(n, (rank + 1) * per)
    return lo, hi


def run_chain_sharded(mcmc_factory, niter, nburnin=0, thin=1, nchains=1, inits=None, seed=0, group=None):
    """Run `nchains` chains split over the ranks of the default (or given) process group.

    mcmc_factory(first_chain, device_index) must return a CompiledMCMC bound to this rank's device.
    inits: None, one dict, or a list of nchains dicts (global chain order).  Returns, on rank 0, the samples of all
    chains [nchains][rows][cols] in global chain order (None on the other ranks)."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    lo, hi = chain_range(nchains, rank, world)
    local_dev = rank % max(1, torch.cuda.device_count()) if torch.cuda.is_available() else 0
    cm = mcmc_factory(lo, local_dev)
    rows = []
    for c in range(lo, hi):
        these = inits[c] if isinstance(inits, (list, tuple)) else inits
        rows.append(cm.model.init_vector(these))
    if hi > lo:
        local = cm.run_chains(np.stack(rows), niter, nburnin, thin, True, seed)
    else:
        local = np.zeros((0, (niter - nburnin) // thin, len(cm.colnames)))
    if world == 1:
        return local
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local, gathered, dst=0, group=group)
    if rank != 0:
        return None
    return np.concatenate([g for g in gathered if g.shape[0] > 0], axis=0)


def hybrid_layout(rank, world, data_shards):
    """(chain group, shard index, number of chain groups) of `rank` in a chain-groups x data-shards layout: the ranks of
    one chain group are consecutive, so on one NVSwitch box a group's all-reduce stays among neighbours."""
    if data_shards < 1 or world % data_shards:
        raise ValueError("world size %d is not a multiple of data_shards %d" % (world, data_shards))
    return rank // data_shards, rank % data_shards, world // data_shards


def run_hybrid(mcmc_factory, niter, nburnin=0, thin=1, nchains=1, inits=None, seed=0, data_shards=1, min_nodes=1 << 16):
    """Chain groups x data shards (third row): the world is cut into world/data_shards chain groups; group g
    owns chains chain_range(nchains, g, groups) and its data_shards ranks each hold a slice of every large data declaration,
    all-reducing block sums inside the group only (one NCCL communicator per group, created inside the device library).
    data_shards = 1 is plain chain sharding, data_shards = world plain data sharding.

    mcmc_factory(first_chain, device_index) as for run_chain_sharded.  Returns on rank 0 the samples of all chains in global
    chain order (None elsewhere)."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    g, shard, groups = hybrid_layout(rank, world, data_shards)
    lo, hi = chain_range(nchains, g, groups)
    if hi == lo:
        raise ValueError("fewer chains (%d) than chain groups (%d)" % (nchains, groups))
    local_dev = rank % max(1, torch.cuda.device_count())
    cm = mcmc_factory(lo, local_dev)
    rows = []
    for c in range(lo, hi):
        these = inits[c] if isinstance(inits, (list, tuple)) else inits
        rows.append(cm.model.init_vector(these))
    if data_shards > 1:
        # every group leader makes an id; all ids travel in one collective and each rank picks its group's
        ids = [None] * world
        dist.all_gather_object(ids, cm.backend.nccl_unique_id() if shard == 0 else None)
        dm = cm._ensure(hi - lo)
        dm.enable_data_sharding(data_shards, shard, ids[g * data_shards], min_nodes=min_nodes)
    local = cm.run_chains(np.stack(rows), niter, nburnin, thin, True, seed)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local if shard == 0 else None, gathered, dst=0)
    if rank != 0:
        return None
    return np.concatenate([x for x in gathered if x is not None], axis=0)
