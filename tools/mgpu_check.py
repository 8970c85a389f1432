"""Multi-GPU parity check through the C ABI (run on a GPU box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py

Every rank solves the SAME problem twice: once alone (no communicator) and once as one of N landmark shards (NCCL
all-reduce of U / b_c / scalars and of S's tiles or the PCG vectors).  The sharded trajectory must match the single-GPU
one to summation-order accuracy, for the dense Schur and the PCG path.  Prints one JSON line per case from rank 0 and
exits non-zero on a mismatch."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bundleadjustment_b200 as bab  # noqa: E402
from bundleadjustment_b200 import synthetic as syn  # noqa: E402


def main() -> int:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    bad = 0
    for solver_name, scale in (("dense", 0.2), ("pcg", 0.2)):
        pb = bab.sharding.landmark_major(syn.flatten(syn.config("K4", scale)))
        kw = dict(max_num_iterations=8)
        if solver_name == "pcg":
            kw.update(linear_solver=bab.capi.ITERATIVE_SCHUR, eta=-1.0, cg_r_tolerance=1e-11, max_linear_solver_iterations=5000)
        opt = bab.default_options(**kw)
        # alone
        s1 = bab.Solver(opt, device=local)
        s1.set_problem(pb)
        sm1 = s1.solve()
        its1 = s1.iterations()
        cam1, pt1 = s1.cameras(), s1.points()
        s1.close()
        # sharded
        s = bab.Solver(opt, device=local)
        uid = torch.zeros(128, dtype=torch.uint8, device=f"cuda:{local}")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(bab.capi.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        s.comm_init(world, rank, bytes(uid.cpu().numpy().tobytes()))
        cut = bab.sharding.partition(pb, world, quadratic=(solver_name == "dense"))
        s.set_problem(bab.sharding.shard(pb, cut, rank))
        sm = s.solve()
        its = s.iterations()
        cam = s.cameras()
        parts = [None] * world
        dist.all_gather_object(parts, np.ascontiguousarray(s.points()))
        pt = np.concatenate(parts, 0)
        torch.cuda.synchronize()
        dist.barrier()
        s.close()
        # the gate bench.py applies before its timed region (bundleadjustment_b200/parity.py): cost trajectory 1e-6, final cost
        # 1e-8, cameras 1e-6, landmarks 1e-6 * max(1, |x|) + the stated conditioning bound -- no percentile waiver
        sol = lambda its_, sm_, cam_, pt_: {"costs": [i["cost"] for i in its_], "final_cost": sm_.final_cost, "cam": cam_, "pt": pt_,
                                            "radius": its_[-1]["trust_region_radius"]}
        par = bab.parity.compare(pb, sol(its, sm, cam, pt), sol(its1, sm1, cam1, pt1))
        flag = torch.tensor([0 if par["ok"] else 1], device=f"cuda:{local}")
        dist.all_reduce(flag)
        bad += int(flag.item())
        if rank == 0:
            print(json.dumps({"case": f"K4 x{scale} {solver_name}", "world": world, "ok": bool(flag.item() == 0), "iterations": len(its),
                              "final_cost_single": sm1.final_cost, "final_cost_sharded": sm.final_cost,
                              **{k: par[k] for k in ("max_cost_rel_diff", "final_cost_rel_diff", "max_cam_diff", "max_pt_diff", "pt_over_1e-6_rel",
                                                     "pt_outside_cond_bound")}}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
