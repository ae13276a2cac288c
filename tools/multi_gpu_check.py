"""Under torchrun on N GPUs: libkima_cuda's own NCCL barrier exchange (kc_comm_*) against the
torch.distributed twin, and sharded evaluation against a single-rank evaluation of all lanes."""
import os, sys
sys.path.insert(0, ".")
import numpy as np
import torch
import torch.distributed as dist
from kima_b200 import LikelihoodEngine, ModelFlags
from kima_b200.dist import NcclBarrierExchange, TorchBarrierExchange, env_rank_world, shard_lanes
from kima_b200.synthetic import make_data, make_params

rank, world, local = env_rank_world()
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tex = TorchBarrierExchange(torch.device("cuda", local))

def bcast_bytes(b):
    obj = [b]
    dist.broadcast_object_list(obj, src=0)
    return obj[0]

flags = ModelFlags(npmax=3)
d = make_data(200)
total = 4096
P, nps = make_params(flags, d, total, "mix", seed=1)
lo, hi = shard_lanes(total, rank, world)
with LikelihoodEngine(d, flags, total, device=local) as eng:
    mine = eng.evaluate(P[:, lo:hi], nps[lo:hi])
    nex = NcclBarrierExchange(eng, rank, world, bcast_bytes)
    rng = np.random.default_rng(rank)
    pairs = np.stack([mine[: 5 + 3 * rank], rng.random(5 + 3 * rank)], axis=1)
    a1, c1 = nex.allgather_above(pairs)
    a2, c2 = tex.allgather_above(pairs)
    assert c1 == c2 and np.array_equal(a1, a2), "kc_allgather_above != torch all_gather"
    cnt = (np.arange(8, dtype=np.uint64) + 1) * np.uint64(rank + 1)
    s1 = nex.allreduce_levels(cnt)
    s2 = tex.allreduce_levels(cnt)
    assert np.array_equal(s1, s2), "kc_allreduce_levels != torch all_reduce"
    # sharded == unsharded
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        full = eng.evaluate(P, nps)
        # 2048 lanes per rank take the CTA-per-particle tiling, 4096 the warp-per-particle one:
        # same arithmetic per epoch, different reduction tree -> equal to rounding
        g = np.concatenate(gathered)
        assert np.max(np.abs(g - full) / np.abs(full)) <= 1e-13, "sharded evaluation differs"
        print("multi_gpu_check ok: world", world, "pairs", a1.shape[0], "levels", s1[:3])
dist.barrier()
dist.destroy_process_group()
