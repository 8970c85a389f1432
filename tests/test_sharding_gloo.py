"""world_size-2 (and 3) gloo test of the multi-GPU host logic on CPU: landmark partition + per-rank shards, each rank
builds ITS reduced-system contribution (here with the CPU oracle standing in for the GPU kernels), a gloo all-reduce
plays NCCL's role, and the sum must equal the single-rank system; per-rank landmark ranges reassemble to the full
array.  The collective layout mirrors ba_capi.cu's enqueue_linear_solve: [S | rhs] summed, LM scalars summed."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, quadratic, ret):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import bundleadjustment_b200 as bab
    from bundleadjustment_b200 import synthetic as syn
    from oracle import ba_oracle as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pb = syn.flatten(syn.make_scene("t", 12, 7, 150, 3.5, 4, arc_deg=40.0))
        # callers may hand observations in any order: shuffle, then let sharding put them landmark-major
        import dataclasses
        perm = np.random.default_rng(0).permutation(pb.obs_cam.shape[0])
        pb = dataclasses.replace(pb, obs_cam=pb.obs_cam[perm].copy(), obs_pt=pb.obs_pt[perm].copy(), obs_uv=pb.obs_uv[perm].copy())
        full = bab.sharding.landmark_major(pb)
        cut = bab.sharding.partition(full, world, quadratic=quadratic)
        mine = bab.sharding.shard(full, cut, rank)
        Nc, Np = full.cam.shape[0], full.pt.shape[0]
        assert mine.pt.shape[0] == cut[rank + 1] - cut[rank]
        assert mine.obs_pt.size == 0 or (mine.obs_pt.min() >= 0 and mine.obs_pt.max() < mine.pt.shape[0])
        # every observation lands on exactly one rank
        n_obs = torch.tensor([mine.obs_cam.shape[0]], dtype=torch.int64)
        dist.all_reduce(n_obs)
        assert int(n_obs) == full.obs_cam.shape[0]
        # global per-camera observation counts (ba_set_problem all-reduces these to classify cameras)
        cnt = torch.from_numpy(np.bincount(mine.obs_cam, minlength=Nc).astype(np.int64))
        dist.all_reduce(cnt)
        assert np.array_equal(cnt.numpy(), np.bincount(full.obs_cam, minlength=Nc))
        # this rank's reduced-system contribution from ITS shard (local landmark indices 0..n_local)
        # Jacobi scaling off: a camera's column norms need ALL of its observations (the library all-reduces U before
        # computing the scale); without scaling every landmark's contribution is shard-independent
        opt = orc.default_options(jacobi_scaling=0)
        S, rhs, cost = orc.schur_system(mine, 1e4, 0, mine.pt.shape[0], opt)
        buf = torch.from_numpy(np.concatenate([S.reshape(-1), rhs, [cost]]))
        dist.all_reduce(buf)                                  # == ncclAllReduce(sum, f64) of [S | rhs | cost]
        S_sum = buf[:S.size].numpy().reshape(S.shape)
        rhs_sum = buf[S.size:S.size + rhs.size].numpy()
        cost_sum = float(buf[-1])
        So, rhso, costo = orc.schur_system(full, 1e4, 0, Np, opt)
        off = ~np.kron(np.eye(Nc, dtype=bool), np.ones((6, 6), dtype=bool))
        assert np.abs(S_sum[off] - So[off]).max() <= 1e-10 * np.abs(So).max()
        assert abs(cost_sum - costo) <= 1e-12 * costo
        # landmark ranges reassemble
        def all_gather(x):
            out = [None] * world
            dist.all_gather_object(out, x)
            return out
        pts = bab.sharding.gather_points(mine.pt, cut, rank, all_gather)
        assert np.array_equal(pts, full.pt)
        ret[rank] = "ok"
    except BaseException as e:      # noqa: BLE001
        import traceback
        ret[rank] = "".join(traceback.format_exception(e))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,quadratic", [(2, True), (2, False), (3, True)])
def test_landmark_sharding_allreduce(world, quadratic, orc):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, quadratic, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        if p.is_alive():
            p.kill()
            pytest.fail("rank hung")
    for r in range(world):
        assert ret.get(r) == "ok", ret.get(r)
