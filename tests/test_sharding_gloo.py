"""CPU test of the N > 1 path: independent paths are sharded over ranks with no
data-path collective; only a final gather.  Two gloo ranks; the per-rank "engine"
is stood in for by the CPU oracle (the GPU engine has no CPU mode), so this
checks the host-side partition / gather logic bench.py uses."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, B_per, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_oracle, load_package
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg, po = load_package(), load_oracle()
    wp, vmax, amax = pkg.workloads.joint_paths(B_per * world, 4, 3, seed=1000)
    sl = slice(rank * B_per, (rank + 1) * B_per)
    dur = np.array([po.plan_joint(wp[b], vmax[b], amax[b], 64)["duration"]
                    for b in range(sl.start, sl.stop)])
    dist.barrier()
    gathered = [torch.empty(B_per, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(dur))
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max-over-ranks timing reduction
    if rank == 0:
        ret["dur"] = torch.cat(gathered).numpy()
        ret["tmax"] = float(t.item())
    dist.destroy_process_group()


def test_two_rank_sharding_matches_single_process(po, pkg):
    import torch.multiprocessing as mp
    world, B_per = 2, 3
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, B_per, ret), nprocs=world, join=True)
    wp, vmax, amax = pkg.workloads.joint_paths(B_per * world, 4, 3, seed=1000)
    ref = np.array([po.plan_joint(wp[b], vmax[b], amax[b], 64)["duration"]
                    for b in range(B_per * world)])
    np.testing.assert_array_equal(ret["dur"], ref)
    assert ret["tmax"] == 2.0


def test_shard_range_is_the_contiguous_block_partition(pkg):
    """tpb_shard_range (host code of the product library, no GPU needed): shard g of G owns
    [g * ceil(B / G), (g + 1) * ceil(B / G)); the blocks tile the batch."""
    for B in (0, 1, 7, 4096, 65536, 65537):
        for G in (1, 2, 3, 4, 8, 13):
            per = -(-B // G)
            at = 0
            for g in range(G):
                first, count = pkg.shard_range(B, G, g)
                assert first == min(B, g * per) and count == min(per, B - first)
                assert first == at
                at += count
            assert at == B
    with pytest.raises(pkg.TpbError):
        pkg.shard_range(8, 2, 2)
