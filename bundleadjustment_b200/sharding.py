"""Landmark sharding for the multi-GPU path: one process per GPU, rank r owns a contiguous
landmark range and ALL of its observations; cameras are replicated.  Host-side logic only (numpy + the C ABI's
ba_partition_landmarks); the collectives live in the library (NCCL) -- see ba_comm_init in include/ba_b200.h."""
from __future__ import annotations

import dataclasses

import numpy as np

from . import capi


def landmark_major(pb):
    """The same problem with observations sorted by (landmark, camera) -- the order ba_partition_landmarks wants."""
    key = pb.obs_pt.astype(np.int64) * np.int64(pb.cam.shape[0]) + pb.obs_cam.astype(np.int64)
    if np.all(key[1:] >= key[:-1]):
        return pb
    order = np.argsort(key, kind="stable")
    return dataclasses.replace(pb, obs_cam=np.ascontiguousarray(pb.obs_cam[order]),
                               obs_pt=np.ascontiguousarray(pb.obs_pt[order]),
                               obs_uv=np.ascontiguousarray(pb.obs_uv[order]))


def partition(pb, n_ranks: int, quadratic: bool) -> np.ndarray:
    """cut[r]..cut[r+1] = rank r's landmarks, balanced by sum t (PCG) or sum t(t+1)/2 (dense Schur work)."""
    pb = landmark_major(pb)
    return capi.partition_landmarks(pb.pt.shape[0], pb.obs_pt, n_ranks, quadratic=quadratic)


def shard(pb, cut, rank: int):
    """Rank's landmark range + its observations (obs_pt rebased to the local range); cameras replicated."""
    a, b = int(cut[rank]), int(cut[rank + 1])
    sel = (pb.obs_pt >= a) & (pb.obs_pt < b)
    return dataclasses.replace(pb, pt=np.ascontiguousarray(pb.pt[a:b]),
                               obs_cam=np.ascontiguousarray(pb.obs_cam[sel]),
                               obs_pt=np.ascontiguousarray(pb.obs_pt[sel] - a),
                               obs_uv=np.ascontiguousarray(pb.obs_uv[sel]))


def gather_points(local_pt: np.ndarray, cut, rank: int, all_gather) -> np.ndarray:
    """Reassemble the full landmark array from the per-rank ranges.  `all_gather(obj) -> list` is the caller's
    control-plane collective (torch.distributed.all_gather_object in bench.py / the tests)."""
    parts = all_gather(np.ascontiguousarray(local_pt))
    out = np.concatenate(parts, 0) if parts else local_pt
    assert out.shape[0] == int(cut[-1])
    return out
