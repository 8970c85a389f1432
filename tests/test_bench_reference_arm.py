"""bench.py --impl reference on CPU: the contract of the reference arm (one JSON line, same metric / unit / config keys as the
product arm, `impl`, `cpu_baseline`, a self-referential `e2e`) and the promise that it never loads the product package or
libba_b200.so.  Runs the oracle port on the smallest named workload (K1, a fraction of a second per solve)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

RUNNER = r"""
import json, os, runpy, sys
sys.argv = ["bench.py", "--impl", "reference", "--workload", "K1", "--steps", "2", "--warmup", "1", "--lm-iters", "3"]
try:
    runpy.run_path(os.path.join(%r, "bench.py"), run_name="__main__")
except SystemExit as e:
    assert not e.code, e.code
maps = open("/proc/self/maps").read()
sys.stderr.write("PRODUCT_LOADED=%%d\n" %% int("libba_b200" in maps or "bundleadjustment_b200" in sys.modules))
""" % ROOT


def _run(env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    p = subprocess.run([sys.executable, "-c", RUNNER], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    return p


def test_reference_arm_contract_and_isolation():
    p = _run()
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, p.stdout                      # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "lm_iters_per_sec" and d["unit"] == "LM iterations/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    cfg = d["config"]
    assert cfg["workload"] == "K1" and cfg["max_num_iterations"] == 3 and cfg["linear_solver"] == "dense"
    assert cfg["cams"] == 30 and cfg["observations"] > 50000
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert d["gpu_launches"] == 0
    assert "PRODUCT_LOADED=0" in p.stderr, p.stderr[-500:]   # only oracle/ ran


def test_reference_arm_other_ranks_exit_quietly():
    """Under torchrun (N > 1) rank 0 alone runs the CPU arm; the other ranks print nothing and exit 0."""
    p = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p.stdout.strip() == ""


def test_product_arm_fails_loudly_without_a_gpu():
    """No CPU fallback: on a box without a CUDA device the product arm of bench.py exits non-zero and prints no result line,
    and the C ABI itself reports BA_ERR_CUDA from ba_create."""
    import pytest
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "K1", "--no-cpu", "--no-k3", "--no-k5", "--steps", "1", "--warmup", "0"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode != 0
    assert p.stdout.strip() == ""
    sys.path.insert(0, ROOT)
    import bundleadjustment_b200 as bab
    with pytest.raises(Exception) as ei:
        bab.Solver(bab.default_options())
    assert "no CPU fallback" in str(ei.value) or "CUDA" in str(ei.value)
