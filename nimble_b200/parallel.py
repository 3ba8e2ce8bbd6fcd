"""Multi-GPU execution: one process per GPU (torch.distributed), two ways to split the work.

chain sharding  rank r owns chains [r*K/R, (r+1)*K/R) and a full replica of the shared arrays; nothing is exchanged
                while sampling, mvSamples are gathered once at the end.  The draw stream is keyed by the GLOBAL chain
                index, so the result is bit-identical to a single-process run of all K chains.
data sharding   every rank holds all chains but only a contiguous slice of each large data declaration; each
                per-proposal block sum (and each conjugate contribution vector) is all-reduced inside the device
                library -- by its own kernel over NVLink peer memory (nimb200_enable_data_sharding_peer, the default) or
                over NCCL (nimb200_enable_data_sharding) -- after which every rank takes the same decision.

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


def attach_data_sharding(dm, nranks, shard, group_ranks=None, transport="peer", min_nodes=1 << 16, group=None):
    """Turn `dm` (a DeviceModel every rank of the group created alike) into this rank's share of a data-sharded group.

    transport "peer": the all-reduce is a kernel of the device library over NVLink peer memory (every rank exports its exchange
    buffer, the handles are all-gathered -- which is also the barrier the protocol needs -- and attached); "nccl": one NCCL
    communicator per group, all-reduces issued from the host.  group_ranks: the global ranks of this group in shard order
    (default: the whole default group)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    group_ranks = list(range(world)) if group_ranks is None else list(group_ranks)
    assert len(group_ranks) == nranks and group_ranks[shard] == rank
    if transport == "peer":
        mine = dm.peer_export(nranks)
        allh = [None] * world
        dist.all_gather_object(allh, mine, group=group)
        dm.enable_data_sharding_peer(nranks, shard, [allh[r] for r in group_ranks], min_nodes=min_nodes)
    elif transport == "nccl":
        ids = [None] * world
        dist.all_gather_object(ids, dm.backend.nccl_unique_id() if shard == 0 else None, group=group)
        dm.enable_data_sharding(nranks, shard, ids[group_ranks[0]], min_nodes=min_nodes)
    else:
        raise ValueError("transport must be 'peer' or 'nccl'")


def waic_from_partials(parts):
    """WAIC of a data-sharded run from the ranks' nimb200_waic_partial results (MCMC_WAIC.R:276-279)."""
    parts = np.asarray(parts, dtype=np.float64).reshape(-1, 4)
    logsum, ssq, nnodes = parts[:, 0].sum(), parts[:, 1].sum(), parts[:, 2].sum()
    cnt = parts[0, 3]
    if cnt < 2:
        return {"WAIC": float("nan"), "lppd": float("nan"), "pWAIC": float("nan"), "count": cnt}
    lppd = logsum - nnodes * np.log(cnt)
    pwaic = ssq / (cnt - 1.0)
    return {"WAIC": -2.0 * (lppd - pwaic), "lppd": lppd, "pWAIC": pwaic, "count": cnt}


def waic_sharded(dm, group=None):
    """Online WAIC of a data-sharded model: every rank holds the statistics of its slice of the data nodes; their sums over
    the group's ranks give the same lppd and pWAIC as the unsharded run (every rank returns the same dictionary)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    parts = [None] * world
    dist.all_gather_object(parts, dm.waic_partial().tolist(), group=group)
    return waic_from_partials(parts)


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


def run_hybrid(mcmc_factory, niter, nburnin=0, thin=1, nchains=1, inits=None, seed=0, data_shards=1, min_nodes=1 << 16,
               transport="peer"):
    """Chain groups x data shards (third row): the world is cut into world/data_shards chain groups; group g
    owns chains chain_range(nchains, g, groups) and its data_shards ranks each hold a slice of every large data declaration,
    all-reducing block sums inside the group only (transport "peer": the library's own all-reduce kernel over NVLink peer memory,
    captured into the iteration's CUDA graph; "nccl": one NCCL communicator per group, host-issued).
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
        # one exchange (peer buffers, or a NCCL communicator) per chain group; the handles / ids of all groups travel in one
        # collective and each rank picks its own group's
        dm = cm._ensure(hi - lo)
        attach_data_sharding(dm, data_shards, shard, [g * data_shards + k for k in range(data_shards)], transport, min_nodes)
    local = cm.run_chains(np.stack(rows), niter, nburnin, thin, True, seed)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local if shard == 0 else None, gathered, dst=0)
    if rank != 0:
        return None
    return np.concatenate([x for x in gathered if x is not None], axis=0)
