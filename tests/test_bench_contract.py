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


def test_parity_check_inputs_have_an_exactly_computable_mean():
    """bench.parity_check compares the multi-GPU result BIT FOR BIT with a mean every rank computes
    locally; that only works if the synthetic gradients sum exactly in fp32 in any order."""
    import itertools
    import torch
    import bench
    numels = [1, 33, 4099]
    world = 8
    per_rank = [bench._exact_grads(torch, numels, r, 3, "cpu")[0] for r in range(world)]
    ints = [bench._exact_grads(torch, numels, r, 3, "cpu")[1] for r in range(world)]
    for t in range(len(numels)):
        want = (sum(ints[r][t] for r in range(world)).to(torch.float64) * 2.0 ** -20 / world)
        assert all(int(v.abs().max()) <= 4096 for v in (ints[r][t] for r in range(world)))
        for order in (range(world), reversed(range(world)), [3, 7, 0, 5, 1, 6, 2, 4]):
            acc = torch.zeros(numels[t])
            for r in order:
                acc = acc + per_rank[r][t]
            assert torch.equal((acc / world).to(torch.float64), want)          # exact, whatever the order
        # pre-scaling by 1/N in the pack (what power-of-two worlds do) gives the same bits
        acc = torch.zeros(numels[t])
        for r in range(world):
            acc = acc + per_rank[r][t] * (1.0 / world)
        assert torch.equal(acc.to(torch.float64), want)
    assert not torch.equal(per_rank[0][2], per_rank[1][2])                     # ranks really differ


def test_nccl_mode_parsing(monkeypatch):
    import b200ddp
    for value, want in ((None, "lazy"), ("1", "1"), ("0", "0"), ("lazy", "lazy"), ("nonsense", "lazy")):
        if value is None:
            monkeypatch.delenv("B200DDP_NCCL", raising=False)
        else:
            monkeypatch.setenv("B200DDP_NCCL", value)
        assert b200ddp.runtime.nccl_mode() == want


def test_reference_arm_task4_runs_replicas_of_the_whole_model():
    """configs[3] on the host cores: N GPUs = N / 2 replicas, so the CPU counterpart is N / 2 gloo ranks of
    the whole lab CNN with plain SGD (codes/task4/model.py:126)."""
    d = json.loads(_run(None, "--gpus", "4", "--model", "task4").strip())       # the later --model wins over _run's
    assert d["impl"] == "reference" and d["n_gpus"] == 4 and "gloo x2" in d["arm"]["where"]
    assert d["config"]["parallelism"] == "dp2" and "task4" in d["config"]["workload"] and d["value"] > 0
