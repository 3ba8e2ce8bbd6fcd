"""Data-sharded linear mixed model (BASELINE config 3) under torchrun: every rank holds all chains and a slice of y."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import nimble_b200 as nb
from nimble_b200 import configs, _capi

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
C = int(sys.argv[1]) if len(sys.argv) > 1 else 8
K = int(sys.argv[2]) if len(sys.argv) > 2 else 200
backend = _capi.default_backend()
m, conf = configs.lmm()
cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local)
dm = cm._ensure(C)
if world > 1:
    ids = [backend.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    dm.enable_data_sharding(world, rank, ids[0], min_nodes=65536)
dm.set_values(np.tile(m.init_vector(), (C, 1)))
dm.run(20, 0, 1, True, 5)
if world > 1:
    dist.barrier()
dm.run(K, 0, 1, False, 5)
ms, launches = dm.last_timing()
ev = configs.evals_per_iteration(cm)
s = dm.samples()
if rank == 0:
    print("lmm data-sharded over %d GPU(s): %d chains, %.3f ms/iter, %d launches/iter, %.3g evals/s, mu mean %.4f"
          % (world, C, ms / K, launches // K, ev * C * K / (ms * 1e-3), s[:, -K // 2:, 0].mean()), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
