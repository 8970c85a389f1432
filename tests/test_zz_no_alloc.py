"""ba_prune_frames_noalloc (include/ba_b200.h): the allocation-free variant of ba_prune_frames, scratch in a static
__device__ array.  Collected LAST (file name) so that a failure here cannot hide another test under `pytest -x`.

Checks, through the C ABI: its output equals ba_prune_frames' on the same inputs (uint8 flags: bit-exact, the tolerance
every prune test of this suite uses), on two calls in a row that share the scratch (the second with FEWER frames and other
poses, so stale entries of the first call sit in the array while it runs), that device memory in use does not grow across
the calls, and that a batch beyond BA_PRUNE_NOALLOC_MAX_FRAMES is rejected with a clear error.  torch is only the device
allocator of the TEST (inputs and outputs are the caller's)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(seed: int, frames: int, per_frame: int):
    """Random but plausible keypoint data: landmarks in front of near-identity cameras, distance gates around the true range."""
    rng = np.random.default_rng(seed)
    pose = np.tile(np.eye(4, dtype=np.float32), (frames, 1, 1))
    for f in range(frames):
        w = rng.normal(scale=0.05, size=3)
        th = np.linalg.norm(w)
        k = w / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        pose[f, :3, :3] = (np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx).astype(np.float32)
        pose[f, :3, 3] = rng.normal(scale=0.2, size=3).astype(np.float32)
    K = np.tile(np.array([[520.0, 0, 320.0], [0, 520.0, 240.0], [0, 0, 1]], np.float32), (frames, 1, 1))
    n = frames * per_frame
    ptr = (np.arange(frames + 1) * per_frame).astype(np.int32)
    fr = np.repeat(np.arange(frames), per_frame)
    X = np.stack([rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(-1.0, 8.0, n)], 1).astype(np.float32)   # some behind the camera
    E = np.linalg.inv(pose.astype(np.float64))
    pc = np.einsum('nij,nj->ni', E[fr, :3, :3], X.astype(np.float64)) + E[fr, :3, 3]
    pix = np.einsum('nij,nj->ni', K[fr].astype(np.float64), pc)
    pix = pix[:, :2] / np.where(np.abs(pix[:, 2:3]) > 1e-9, pix[:, 2:3], 1.0)
    xy = (pix + rng.normal(scale=4.0, size=(n, 2))).astype(np.float32)              # around the chi-square gate
    d = np.linalg.norm(X - pose[fr, :3, 3], axis=1).astype(np.float32)
    return dict(pose=pose.reshape(frames, 16), K=K.reshape(frames, 9), ptr=ptr, xy=xy,
                octave=rng.integers(0, 6, size=n).astype(np.int32), xyz=X,
                mn=(d * rng.uniform(0.3, 1.05, size=n)).astype(np.float32), mx=(d * rng.uniform(0.95, 2.0, size=n)).astype(np.float32),
                has=(rng.random(n) > 0.1).astype(np.uint8), out=(rng.random(n) > 0.7).astype(np.uint8))


def _existing(bab, d, consider):
    return bab.prune_frames(d["pose"], d["K"], d["ptr"], d["xy"], d["octave"], d["xyz"], d["mn"], d["mx"], d["has"], d["out"],
                            consider_outlier=consider)


def _noalloc(bab, torch, d, consider, stream):
    dev = {k: torch.from_numpy(v).cuda() for k, v in d.items()}          # caller-owned device inputs / in-out flags
    torch.cuda.synchronize()
    F, n = d["ptr"].shape[0] - 1, int(d["ptr"][-1])
    bab.capi.prune_frames_noalloc(F, n, *[dev[k].data_ptr() for k in ("pose", "K", "ptr", "xy", "octave", "xyz", "mn", "mx", "has", "out")],
                                  consider_outlier=consider, device=0, stream=stream.cuda_stream)
    stream.synchronize()
    return dev["out"].cpu().numpy()


def test_prune_noalloc_matches_prune_frames_twice(bab):
    import torch
    stream = torch.cuda.Stream()
    first, second = _frames(11, 300, 160), _frames(12, 37, 211)            # the second call reuses 37 of the 300 scratch entries
    want = [_existing(bab, first, True), _existing(bab, second, False)]
    _noalloc(bab, torch, _frames(13, 8, 16), True, stream)                 # set-up call (cudaGetSymbolAddress) before memory is sampled
    free0 = torch.cuda.mem_get_info()[0]
    got = [_noalloc(bab, torch, first, True, stream), _noalloc(bab, torch, second, False, stream)]
    for g, w, d in zip(got, want, (first, second)):
        assert g.dtype == np.uint8 and g.shape == w.shape
        assert np.array_equal(g, w), f"{int((g != w).sum())} of {g.size} flags differ"
        assert (g != d["out"]).sum() > 0 and g.sum() > 0 and (g == 0).sum() > 0      # the kernel did flag something
    # the calls above allocated only through torch's caching allocator (the test's inputs, freed and re-used); what the
    # DRIVER reports as free must not have dropped by more than those inputs could account for
    torch.cuda.synchronize()
    in_bytes = sum(v.nbytes for v in first.values())
    assert free0 - torch.cuda.mem_get_info()[0] <= 4 * in_bytes + (64 << 20)
    # again, same inputs, the scratch now holding the second call's entries: identical flags
    assert np.array_equal(_noalloc(bab, torch, first, True, stream), want[0])


def test_prune_noalloc_rejects_oversize(bab):
    import torch
    cap = bab.capi.PRUNE_NOALLOC_MAX_FRAMES
    dummy = torch.zeros(64, device="cuda")
    p = dummy.data_ptr()
    with pytest.raises(Exception) as ei:
        bab.capi.prune_frames_noalloc(cap + 1, 10, *([p] * 10))
    assert "BA_PRUNE_NOALLOC_MAX_FRAMES" in str(ei.value)
    bab.capi.prune_frames_noalloc(0, 0, *([0] * 10))                      # empty batch: nothing to do, no error
