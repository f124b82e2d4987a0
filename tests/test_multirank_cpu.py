"""World-size-2 gloo test of the multi-GPU scheme on CPU: frames sharded across ranks, u32 accumulators summed with one
all-reduce, colours identical to the single-rank run (integer sums are associative).  The per-rank worker is the oracle
here (no GPU in this container); the GPU path uses the same sharding helper and pcp_allreduce (NCCL)."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_dir):
    import torch
    import torch.distributed as dist
    from conftest import make_scene
    from oracle import pyoracle as O
    from pointcloudprocessor_b200.shard import frames_of_rank
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = make_scene(20_000, 6, 320, 240)
    oc = O.Colorizer(O.camera(sc["intr"], sc["dist"]), 320, 240, sc["cloud"])
    mine = frames_of_rank(6, rank, world, mode)
    for i in mine:
        oc.frame(sc["poses"][i], sc["images"][i])
    acc = torch.from_numpy(oc.accumulators()[0].astype(np.int64))
    dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    np.save(os.path.join(out_dir, f"acc_{mode}_{rank}.npy"), acc.numpy().astype(np.uint32))
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["interleaved", "blocks"])
def test_two_ranks_equal_one(tmp_path, oracle, mode):
    import torch.multiprocessing as mp
    from conftest import make_scene
    port = _free_port()
    mp.spawn(_worker, args=(2, port, mode, str(tmp_path)), nprocs=2, join=True)
    sc = make_scene(20_000, 6, 320, 240)
    oc = oracle.Colorizer(oracle.camera(sc["intr"], sc["dist"]), 320, 240, sc["cloud"])
    for i in range(6):
        oc.frame(sc["poses"][i], sc["images"][i])
    ref = oc.accumulators()[0]
    a0 = np.load(tmp_path / f"acc_{mode}_0.npy")
    a1 = np.load(tmp_path / f"acc_{mode}_1.npy")
    assert np.array_equal(a0, ref) and np.array_equal(a1, ref)
    assert ref[:, 3].sum() > 0


def test_sharding_is_a_partition():
    from pointcloudprocessor_b200.shard import frames_of_rank
    for mode in ("interleaved", "blocks", "dealt"):
        for world in (1, 2, 3, 4, 8):
            for n in (0, 1, 7, 200):
                got = sorted(sum((frames_of_rank(n, r, world, mode) for r in range(world)), []))
                assert got == list(range(n))
    with pytest.raises(ValueError):
        frames_of_rank(4, 2, 2)


def test_deal_by_cost_balances_and_is_a_partition():
    from pointcloudprocessor_b200.shard import deal_by_cost
    rng = np.random.default_rng(3)
    costs = rng.uniform(3.0, 8.0, 80)
    for world in (1, 2, 4, 8):
        owner = deal_by_cost(costs, world)
        assert len(owner) == 80 and set(owner) == set(range(world))
        loads = [costs[[b for b in range(80) if owner[b] == r]].sum() for r in range(world)]
        assert max(loads) - min(loads) <= costs.min() + 1e-9          # LPT: the spread is below one (small) unit
        assert owner == deal_by_cost(list(costs), world)               # deterministic
    # a rank that is 20 % slower gets ~20 % less work
    owner = deal_by_cost(costs, 4, speed=[1.0, 1.0, 1.0, 0.8])
    loads = [costs[[b for b in range(80) if owner[b] == r]].sum() for r in range(4)]
    assert abs(loads[3] / np.mean(loads[:3]) - 0.8) < 0.05
    with pytest.raises(ValueError):
        deal_by_cost(costs, 2, speed=[1.0])
