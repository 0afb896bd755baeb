"""CPU tests of the bench contract: `bench.py --impl reference` (the reference's kernel source on the host cores, oracle/_ref, or the C port)
prints ONE JSON line with the keys the driver reads, ranks other than 0 exit silently, and the scene table holds BASELINE.json's configurations."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, args=()):
    env = dict(os.environ)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                           "--cpu-points", "1", "--config", "cfg1", *args], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


def test_reference_arm_prints_the_contract_line():
    p = _run()
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["n_gpus"] == 1 and d["gpu_launches"] == 0
    assert d["unit"] == "Gaussian-bin evals/s" and d["value"] > 0 and d["ms_per_step"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and "scan points" in cb["sample"]
    assert d["config"]["workload"].startswith("cfg1: 1000 Gaussians, 16x16 confocal scan")


def test_reference_arm_other_ranks_do_nothing():
    p = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, args=("--gpus", "2"))
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_scene_table_holds_the_baseline_configurations():
    from nlos_gaussian_renderer_b200 import scenes
    want = {"cfg1": (1000, 16, 16, 128), "cfg2": (50000, 64, 64, 512), "cfg3": (200000, 64, 64, 512),
            "cfg4": (500000, 128, 128, 1024), "cfg5": (1000000, 256, 256, 1024), "north_star": (1000000, 128, 128, 1024)}
    for k, v in want.items():
        assert tuple(scenes.CONFIGS[k][:4]) == v, k
    cfgs = json.load(open(os.path.join(ROOT, "BASELINE.json")))["configs"]
    assert re.search(r"50k Gaussians, 64.64 confocal scan, 512 bins", cfgs[1])
    assert re.search(r"500k Gaussians to a synthetic 128.128.1024 transient", cfgs[3])   # the bench default (configs[3], the 1/2/4/8-GPU fitting loop)
    import bench
    sys.argv = ["bench.py"]
    a = bench.parse()
    assert a.config == "cfg4" and a.scaling == "strong" and a.gpus == 1 and a.mode == "analytic"


def test_reference_arm_uses_every_core_under_torchrun_env():
    """torchrun exports OMP_NUM_THREADS=1: the reference arm must ask for all host cores itself (round-1 SCALE ran it on one core at N >= 2)."""
    p = _run({"OMP_NUM_THREADS": "1"})
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == (os.cpu_count() or 1)
