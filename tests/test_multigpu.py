"""Two ranks on two GPUs against one: frames sharded, pcp_allreduce (NCCL) — SUM accumulators (all-reduce), mask
accumulators (sum / max planes) and TOP5 slot lists (all-gather + merge kernel) must equal the single-GPU run bit for bit.
Needs two devices: skipped on a one-GPU box (run with `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_POINTS, N_FRAMES, W, H = 60_001, 9, 480, 360


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(gc, sc, frames, mask_frames):
    for i in frames:
        gc.submit_frame(sc["poses"][i], sc["images"][i], mask=sc["masks"][i] if i in mask_frames else None)


def _results(gc, mode, exchange="allreduce"):
    from pointcloudprocessor_b200 import api
    if exchange == "exchange_finalize":  # reduce-scatter + finalize of the slice + all-gather: colours, counts and masks of the job
        rgb, cnt = gc.exchange_finalize()
        mh, mm = gc.mask_accumulators()
        if mode == api.ACCUM_SUM:
            with pytest.raises(api.PcpError):
                gc.accumulators()      # the sums were consumed (TOP5 mode falls back to all-reduce + finalize and keeps them)
        with pytest.raises(api.PcpError):
            gc.exchange_finalize()     # one exchange per pass
        return dict(mh=mh, mm=mm, rgb=rgb, cnt=cnt)
    acc, mh, mm = gc.accumulators()
    rgb, cnt = gc.finalize()
    out = dict(mh=mh, mm=mm, rgb=rgb, cnt=cnt)
    if mode == api.ACCUM_SUM:
        out["acc"] = acc
    return out


def _worker(rank, world, port, mode, sharding, exchange, out_dir):
    import torch
    import torch.distributed as dist
    from conftest import make_scene
    from pointcloudprocessor_b200 import api
    from pointcloudprocessor_b200.shard import frames_of_rank
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    sc = make_scene(N_POINTS, N_FRAMES, W, H, masks=True)
    p = api.default_params(image_size=(W, H), accum_mode=mode)
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), p, device=rank) as gc:
        ids = [gc.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        gc.nccl_init(ids[0], rank, world)
        if exchange == "exchange_finalize":
            # the map crosses the host link once: every rank uploads its slice, the slices are all-gathered (pcp_set_cloud_sharded);
            # float4 records for one sharding, the reference's 32-byte pcl::PointXYZI records for the other; the point count is
            # odd, so the last slice is the shorter one
            cloud = sc["cloud"]
            if sharding == "blocks":
                pcl = np.zeros((cloud.shape[0], 8), np.float32)
                pcl[:, :3], pcl[:, 3], pcl[:, 4] = cloud[:, :3], 1.0, cloud[:, 3]
                cloud = pcl
            gc.set_cloud_sharded(cloud, rank, world)
        else:
            gc.set_cloud(sc["cloud"])
        mine = frames_of_rank(N_FRAMES, rank, world, sharding)
        if sharding == "interleaved":
            gc.set_frame_numbering(rank, world)
        else:
            gc.set_frame_numbering(mine[0] if mine else 0, 1)
        # masks only on frames 0, 2, 4, ... of rank 0's share in the blocks case: rank 1 may then hold no mask accumulator at all
        mask_frames = set(range(0, N_FRAMES, 2)) if sharding == "interleaved" else set(frames_of_rank(N_FRAMES, 0, world, "blocks"))
        _run(gc, sc, mine, mask_frames)
        if exchange == "allreduce":
            gc.allreduce()
            with pytest.raises(api.PcpError):
                gc.allreduce()         # a second exchange would double-count
        res = _results(gc, mode, exchange)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), **res)
    dist.destroy_process_group()


@pytest.mark.parametrize("exchange", ["allreduce", "exchange_finalize"])
@pytest.mark.parametrize("sharding", ["interleaved", "blocks"])
@pytest.mark.parametrize("mode_name", ["sum", "top5"])
def test_two_gpus_equal_one(tmp_path, built_lib, mode_name, sharding, exchange):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from conftest import make_scene
    from pointcloudprocessor_b200 import api
    from pointcloudprocessor_b200.shard import frames_of_rank
    mode = api.ACCUM_SUM if mode_name == "sum" else api.ACCUM_TOP5
    mp.spawn(_worker, args=(2, _free_port(), mode, sharding, exchange, str(tmp_path)), nprocs=2, join=True)
    sc = make_scene(N_POINTS, N_FRAMES, W, H, masks=True)
    mask_frames = set(range(0, N_FRAMES, 2)) if sharding == "interleaved" else set(frames_of_rank(N_FRAMES, 0, 2, "blocks"))
    with api.Colorizer(api.create_camera("pinhole", sc["intr"], sc["dist"]), api.default_params(image_size=(W, H), accum_mode=mode)) as gc:
        gc.set_cloud(sc["cloud"])
        _run(gc, sc, range(N_FRAMES), mask_frames)
        ref = _results(gc, mode)
    assert ref["cnt"].sum() > 0 and ref["mh"].sum() > 0 and (ref["cnt"] > 1).sum() > 100  # points seen from several frames, on both ranks
    for r in range(2):
        got = np.load(tmp_path / f"r{r}.npz")
        for k, v in ref.items():
            if k not in got.files:
                continue  # exchange_finalize hands back colours / counts / masks, not the sums
            assert np.array_equal(got[k], v), (r, k, int((got[k] != v).sum()))
