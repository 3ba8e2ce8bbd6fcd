"""Scaling sweep of one BASELINE configuration under torchrun, any chain-groups x data-shards layout:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      scripts/scale_config.py <lmm|glmm|logistic|occupancy> <total chains> <iterations> <data_shards> [peer|nccl]
Each chain group owns total_chains / groups chains; time = max over ranks of the device-timed run (CUDA events inside the
library); value = evaluations of all chains / that time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import nimble_b200 as nb
from nimble_b200 import configs, _capi, parallel

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
name = sys.argv[1]
C = int(sys.argv[2])
K = int(sys.argv[3])
shards = int(sys.argv[4]) if len(sys.argv) > 4 else 1
g, shard, groups = parallel.hybrid_layout(rank, world, shards)
lo, hi = parallel.chain_range(C, g, groups)
backend = _capi.default_backend()
m, conf = {"lmm": configs.lmm, "glmm": configs.glmm, "logistic": configs.logistic, "occupancy": configs.occupancy}[name]()
cm = nb.compileNimble(nb.buildMCMC(conf), backend=backend, device=local, first_chain=lo)
dm = cm._ensure(hi - lo)
transport = sys.argv[5] if len(sys.argv) > 5 else "peer"
if shards > 1:
    parallel.attach_data_sharding(dm, shards, shard, [g * shards + k for k in range(shards)], transport, 65536)
dm.set_values(np.tile(m.init_vector(), (hi - lo, 1)))
dm.run(3, 0, 1, True, 5)
if world > 1:
    dist.barrier()
dm.run(K, 0, 1, False, 5)
ms, launches = dm.last_timing()
if world > 1:
    t = torch.tensor([ms], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
ev = configs.evals_per_iteration(cm)
n_ar, replayed = dm.last_allreduces()
if rank == 0:
    print("%s on %d GPU(s) = %d chain group(s) x %d data shard(s) [%s]: %d chains total, %.3f ms/iter, %.3g evals/s, %.1f all-reduces/iter, graph %s"
          % (name, world, groups, shards, transport if shards > 1 else "-", C, ms / K, ev * C * K / (ms * 1e-3), n_ar / K, replayed), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
