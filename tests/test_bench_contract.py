"""bench.py's output contract, checked on the reference (CPU) arm - no GPU needed."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra_env=None, *args):
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    env.update(extra_env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "labnet",
                          "--steps", "2", "--warmup", "1", *args], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


def test_reference_arm_prints_one_contract_line():
    stdout = _run(None, "--gpus", "1")
    lines = [ln for ln in stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, stdout                      # exactly one line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "train samples/sec" and d["unit"] == "samples/s"
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["dtype"] == "f32" and d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["gpu_launches"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_two_gloo_ranks_and_silent_other_ranks():
    d = json.loads(_run(None, "--gpus", "2").strip())
    assert d["n_gpus"] == 2 and d["config"]["parallelism"] == "dp2" and "gloo x2" in d["arm"]["where"] and d["value"] > 0
    # under torchrun only rank 0 works and prints; the others exit 0 without a word
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2") == ""


def test_reference_arm_survives_torch_distributed_run():
    """Launched exactly as the driver launches it for N > 1.  The gloo workers must not inherit
    the elastic agent's store settings (round 1: they did, and init_process_group never returned)."""
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", os.path.join(ROOT, "bench.py"),
                          "--impl", "reference", "--gpus", "2", "--model", "labnet", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip().startswith("{")]
    assert len(lines) == 1, out.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["steps"] == 2 and d["warmup"] == 1 and d["value"] > 0


def test_both_arms_spell_the_workload_identically():
    import bench
    assert bench.workload_config("resnet18", 128, 4) == bench.workload_config("resnet18", 128, 4)
    cfg = bench.workload_config("resnet18", 128, 4)
    assert cfg["params"] == 11181642 and cfg["tensors"] == 62 and cfg["global_batch"] == 512
    assert "model" not in cfg and "workload" in cfg
