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
