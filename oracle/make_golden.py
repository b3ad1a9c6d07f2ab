"""Generate tests/golden/*.npz by running the REFERENCE itself.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Run in the build
container, where /root/reference is mounted:

    python -m oracle.make_golden

What is frozen:
  * ddp_labnet_ws{2,4}.npz  - the reference's hot loop (codes/task2/model.py:49-63)
    on the lab CNN with synthetic batches: reference `Net` (codes/task3/model.py:12-37),
    reference `init_parameters` + `allreduce_average_gradients`
    (codes/task2/dist_utils.py:33-42) imported by path, process group from the
    reference's gloo init (codes/task4/dist_utils.py:6-15), torch.optim.SGD(lr=.01,
    momentum=.9) (codes/task2/model.py:131).  Holds every rank's LOCAL gradients
    (the inputs of the sync step) and the averaged gradients / parameters /
    momentum the reference produced from them.
  * ddp_grid_<variant>_ws<N>.npz - reference sync + torch SGD on a ragged tensor
    set with injected gradients (oracle/synth.py), several SGD flag sets.
  * sampler_*.npz - torch DistributedSampler (codes/task2/model.py:124) outputs.
  * opt_{gd,adam,adam_b}.npz - the lab's own GdOptimizer / AdamOptimizer
    (codes/task1/pytorch/MyOptimizer.py:18-43, imported by path) on injected gradients.
  * bf16_hook_ws{2,8}.npz - torch's bf16_compress_hook under DDP/gloo on injected gradients
    (the semantics bf16 buckets are compared with; the reference itself is fp32 only).

While generating, every oracle restatement (oracle/ref_ddp.py,
oracle/sampler_oracle.py) is run on the same inputs and must agree bit for bit
with the reference, otherwise the script aborts.
"""
from __future__ import annotations

import hashlib
import importlib.util
import os
import sys
import tempfile

import numpy as np
import torch
import torch.multiprocessing as mp

REF = "/root/reference/codes"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _load(name, path, extra_sys_path=None):
    if extra_sys_path:
        sys.path.insert(0, extra_sys_path)
    try:
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if extra_sys_path:
            sys.path.remove(extra_sys_path)
    return mod


def _flat(tensors):
    return np.concatenate([t.detach().cpu().numpy().reshape(-1) for t in tensors]).astype(np.float32)


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


# ------------------------------------------------------------------ workers
def _labnet_worker(rank, ws, port, tmp, steps):
    from oracle import ref_ddp
    torch.set_num_threads(1)
    t2 = _load("ref_t2_dist_utils", f"{REF}/task2/dist_utils.py")
    t3 = _load("ref_t3_dist_utils", f"{REF}/task3/dist_utils.py")
    t4 = _load("ref_t4_dist_utils", f"{REF}/task4/dist_utils.py")
    t4.dist_init(ws, rank, "127.0.0.1", str(port))                 # gloo
    # reference Net; its module imports `dist_utils` and `sampler` from its own dir
    sys.modules.setdefault("matplotlib", type(sys)("matplotlib"))
    m3 = _load("ref_t3_model", f"{REF}/task3/model.py", f"{REF}/task3")

    def run(net_ctor, initp, sync):
        torch.manual_seed(0 if rank == 0 else 100 + rank)          # ranks differ until the broadcast
        model = net_ctor(1, 10)
        opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
        loss_fn = torch.nn.CrossEntropyLoss()
        initp(model)
        rec = {"p0": _flat(model.parameters())}
        for s in range(steps):
            x, y = ref_ddp.synthetic_batch(rank, s, 32, (1, 28, 28), 10)
            loss = loss_fn(model(x), y)
            opt.zero_grad()
            loss.backward()
            rec[f"local_s{s}"] = _flat([p.grad for p in model.parameters()])
            sync(model)
            rec[f"avg_s{s}"] = _flat([p.grad for p in model.parameters()])
            opt.step()
            rec[f"p_s{s}"] = _flat(model.parameters())
            rec[f"m_s{s}"] = _flat([opt.state[p]["momentum_buffer"] for p in model.parameters()])
        return rec

    ref = run(m3.Net, t2.init_parameters, t2.allreduce_average_gradients)
    ref3 = run(m3.Net, t3.init_parameters, t3.average_gradients)
    mine = run(ref_ddp.LabNet, ref_ddp.init_parameters, ref_ddp.allreduce_average_gradients)
    mine3 = run(ref_ddp.LabNet, ref_ddp.init_parameters, ref_ddp.average_gradients)
    for other in (ref3, mine, mine3):
        for k in ref:
            assert np.array_equal(ref[k], other[k]), f"oracle restatement differs from the reference at {k}"
    np.savez(os.path.join(tmp, f"r{rank}.npz"), **ref)
    torch.distributed.destroy_process_group()


def _grid_worker(rank, ws, port, tmp, variant, steps, kind):
    from oracle import ref_ddp, synth
    torch.set_num_threads(1)
    t2 = _load("ref_t2_dist_utils", f"{REF}/task2/dist_utils.py")
    t4 = _load("ref_t4_dist_utils", f"{REF}/task4/dist_utils.py")
    t4.dist_init(ws, rank, "127.0.0.1", str(port))
    gen = synth.grid_grad if kind == "grid" else synth.dense_grad
    hp = synth.SGD_VARIANTS[variant]

    def run(sync):
        params = torch.nn.ParameterList(
            [torch.nn.Parameter(torch.from_numpy(synth.init_param(n, t))) for t, n in enumerate(synth.RAGGED_NUMELS)])
        opt = torch.optim.SGD(params.parameters(), **hp)
        rec = {}
        for s in range(steps):
            opt.zero_grad()
            for t, p in enumerate(params):
                p.grad = torch.from_numpy(gen(p.numel(), rank, s, t))
            sync(params)
            opt.step()
        rec["avg"] = _flat([p.grad for p in params])
        rec["p"] = _flat(params)
        if hp.get("momentum", 0) != 0:
            rec["m"] = _flat([opt.state[p]["momentum_buffer"] for p in params])
        return rec

    ref = run(t2.allreduce_average_gradients)
    mine = run(ref_ddp.allreduce_average_gradients)
    for k in ref:
        assert np.array_equal(ref[k], mine[k]), k
    # the numpy restatement of the update, fed the reference's averaged grads
    if rank == 0:
        p = [synth.init_param(n, t) for t, n in enumerate(synth.RAGGED_NUMELS)]
        buf = [None] * len(p)
        for s in range(steps):
            for t, n in enumerate(synth.RAGGED_NUMELS):
                per_rank = [torch.from_numpy(gen(n, r, s, t)) for r in range(ws)]
                g = ref_ddp.average_over_ranks(per_rank).numpy()
                p[t], buf[t] = ref_ddp.sgd_step_numpy(p[t], g, buf[t], **hp)
        got = np.concatenate(p)
        if kind == "grid":   # order-independent inputs: must be bit-exact at any ws
            assert np.array_equal(got, ref["p"]), "sgd_step_numpy differs from torch.optim.SGD"
        np.savez(os.path.join(tmp, "r0.npz"), **ref)
    torch.distributed.destroy_process_group()


def _bf16_hook_worker(rank, ws, port, tmp):
    """torch's own bf16_compress_hook under DDP/gloo (the semantics `bf16 buckets` are held to):
    a module whose gradients are exactly the injected tensors."""
    from oracle import ref_ddp, synth
    from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.set_num_threads(1)
    t4 = _load("ref_t4_dist_utils", f"{REF}/task4/dist_utils.py")
    t4.dist_init(ws, rank, "127.0.0.1", str(port))

    class Inject(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.ps = torch.nn.ParameterList([torch.nn.Parameter(torch.zeros(n)) for n in synth.RAGGED_NUMELS])

        def forward(self, xs):
            return sum((p * x).sum() for p, x in zip(self.ps, xs))      # d/dp = x

    model = DDP(Inject())
    model.register_comm_hook(None, default_hooks.bf16_compress_hook)
    xs = [torch.from_numpy(synth.dense_grad(n, rank, 0, t)) for t, n in enumerate(synth.RAGGED_NUMELS)]
    model(xs).backward()
    got = _flat([p.grad for p in model.module.ps])
    mine = torch.cat([ref_ddp.bf16_wire_average(
        [torch.from_numpy(synth.dense_grad(n, r, 0, t)) for r in range(ws)]) for t, n in enumerate(synth.RAGGED_NUMELS)]).numpy()
    if ws <= 2:
        assert np.array_equal(got, mine), "bf16_wire_average differs from torch's bf16_compress_hook"
    else:       # gloo's bf16 summation order is not rank order
        assert np.max(np.abs(got - mine)) <= 2e-2 * np.max(np.abs(got))
    if rank == 0:
        np.savez(os.path.join(tmp, "r0.npz"), avg=got)
    torch.distributed.destroy_process_group()


def make_bf16_hook(ws):
    r0 = _spawn(_bf16_hook_worker, ws)["r0"]
    r0["world"] = np.int64(ws)
    np.savez_compressed(os.path.join(OUT, f"bf16_hook_ws{ws}.npz"), **r0)
    print("bf16 hook ws", ws, "ok")


def _spawn(fn, ws, *args):
    port = 29600 + (os.getpid() + _spawn.n) % 300
    _spawn.n += 1
    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(fn, args=(ws, port, tmp) + args, nprocs=ws, join=True)
        return {f[:-4]: dict(np.load(os.path.join(tmp, f))) for f in sorted(os.listdir(tmp))}


_spawn.n = 0


# ------------------------------------------------------------------ drivers
def make_labnet(ws, steps=2):
    ranks = _spawn(_labnet_worker, ws, steps)
    r0 = ranks["r0"]
    out = {"p0": r0["p0"], "world": np.int64(ws), "steps": np.int64(steps)}
    for s in range(steps):
        for r in range(ws):
            out[f"local_r{r}_s{s}"] = ranks[f"r{r}"][f"local_s{s}"]
            # every rank must hold the same averaged state (A1: observable on every rank)
            for k in (f"avg_s{s}", f"p_s{s}", f"m_s{s}"):
                assert np.array_equal(ranks[f"r{r}"][k], r0[k]), (k, r)
        out[f"avg_s{s}"] = r0[f"avg_s{s}"]
    out["p_final"] = r0[f"p_s{steps - 1}"]
    out["m_final"] = r0[f"m_s{steps - 1}"]
    np.savez_compressed(os.path.join(OUT, f"ddp_labnet_ws{ws}.npz"), **out)
    print("labnet ws", ws, "ok")


def make_grid(variant, ws, steps=3, kind="grid"):
    r0 = _spawn(_grid_worker, ws, variant, steps, kind)["r0"]
    r0["world"] = np.int64(ws)
    r0["steps"] = np.int64(steps)
    np.savez_compressed(os.path.join(OUT, f"ddp_{kind}_{variant}_ws{ws}.npz"), **r0)
    print(kind, variant, "ws", ws, "ok")


def make_sampler():
    from torch.utils.data.distributed import DistributedSampler
    from oracle import sampler_oracle as so

    class DS:
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n

    def torch_shard(n, N, r, seed, epoch, shuffle, drop_last):
        s = DistributedSampler(DS(n), num_replicas=N, rank=r, shuffle=shuffle, seed=seed, drop_last=drop_last)
        s.set_epoch(epoch)
        return np.asarray(list(s), dtype=np.int64)

    small = {}
    for n in (0, 1, 2, 5, 10, 17, 1000):
        for N in (1, 2, 3, 4, 8):
            for drop_last in (False, True):
                for epoch in (0, 1):
                    for shuffle in (True, False):
                        for r in range(N):
                            ref = torch_shard(n, N, r, 3, epoch, shuffle, drop_last)
                            assert np.array_equal(ref, so.distributed_sampler_indices(n, N, r, 3, epoch, shuffle, drop_last))
                            assert np.array_equal(ref, so.c_distributed_sampler_indices(n, N, r, 3, epoch, shuffle, drop_last))
                            small[f"n{n}_N{N}_d{int(drop_last)}_e{epoch}_s{int(shuffle)}_r{r}"] = ref.astype(np.int32)
    np.savez_compressed(os.path.join(OUT, "sampler_small.npz"), **small)

    # the reference's own configuration: MNIST-length dataset, 2 ranks, seed 0, epoch 0
    mnist = {f"r{r}": torch_shard(60000, 2, r, 0, 0, True, False).astype(np.int32) for r in range(2)}
    for r in range(2):
        assert np.array_equal(mnist[f"r{r}"], so.c_distributed_sampler_indices(60000, 2, r))
        assert np.array_equal(mnist[f"r{r}"], so.distributed_sampler_indices(60000, 2, r))
    np.savez_compressed(os.path.join(OUT, "sampler_mnist_ws2.npz"), **mnist)

    # larger cases as digests + head/tail
    big = {}
    for (n, N, seed, epoch, drop_last) in ((50000, 8, 0, 0, False), (60000, 8, 0, 1, False),
                                           (1281167, 8, 0, 0, False), (1281167, 8, 0, 3, True),
                                           (100003, 4, 11, 2, False)):
        for r in (0, N - 1):
            ref = torch_shard(n, N, r, seed, epoch, True, drop_last)
            assert np.array_equal(ref, so.c_distributed_sampler_indices(n, N, r, seed, epoch, True, drop_last))
            key = f"n{n}_N{N}_seed{seed}_e{epoch}_d{int(drop_last)}_r{r}"
            big[key + "_sha"] = np.frombuffer(bytes.fromhex(sha(ref)), np.uint8)
            big[key + "_head"] = ref[:16]
            big[key + "_tail"] = ref[-16:]
            big[key + "_len"] = np.int64(len(ref))
    np.savez_compressed(os.path.join(OUT, "sampler_big.npz"), **big)
    print("sampler ok")


def make_lab_optimizers(steps=4):
    """The lab's own GdOptimizer / AdamOptimizer (codes/task1/pytorch/MyOptimizer.py), imported by path
    and run on the ragged tensor set with injected gradients; the numpy restatements must agree bit
    for bit.  Steps 0..steps-1 use dense gradients; one tensor also sees zeros (sqrt(0) + eps path)."""
    from oracle import ref_ddp, synth
    ref = _load("ref_t1_myoptimizer", f"{REF}/task1/pytorch/MyOptimizer.py")
    numels = synth.RAGGED_NUMELS

    def grad(n, s, t):
        g = synth.dense_grad(n, 0, s, t)
        if t == 11:
            g[::3] = 0.0
        return g

    for name, ctor, hp in (("gd", ref.GdOptimizer, dict(lr=0.05)),
                           ("adam", ref.AdamOptimizer, dict(lr=0.002, b1=0.9, b2=0.999, epsilon=1e-8)),
                           ("adam_b", ref.AdamOptimizer, dict(lr=0.01, b1=0.8, b2=0.95, epsilon=1e-6))):
        params = [torch.nn.Parameter(torch.from_numpy(synth.init_param(n, t))) for t, n in enumerate(numels)]
        opt = ctor(params, **hp)
        kw = dict(lr=hp["lr"]) if name == "gd" else dict(lr=hp["lr"], beta1=hp["b1"], beta2=hp["b2"], eps=hp["epsilon"])
        p_np = [synth.init_param(n, t) for t, n in enumerate(numels)]
        m_np = [np.zeros(n, np.float32) for n in numels]
        v_np = [np.zeros(n, np.float32) for n in numels]
        p_t = [torch.from_numpy(a.copy()) for a in p_np]
        m_t = [torch.zeros(n) for n in numels]
        v_t = [torch.zeros(n) for n in numels]
        for s in range(steps):
            for t, p in enumerate(params):
                p.grad = torch.from_numpy(grad(p.numel(), s, t))
            opt.step()
            for t, n in enumerate(numels):
                g = grad(n, s, t)
                if name == "gd":
                    p_np[t] = ref_ddp.gd_step_numpy(p_np[t], g, **kw)
                    p_t[t] = ref_ddp.gd_step_torch(p_t[t], torch.from_numpy(g), **kw)
                else:
                    p_np[t], m_np[t], v_np[t] = ref_ddp.adam_step_numpy(p_np[t], g, m_np[t], v_np[t], **kw)
                    p_t[t], m_t[t], v_t[t] = ref_ddp.adam_step_torch(p_t[t], torch.from_numpy(g), m_t[t], v_t[t], **kw)
        out = {"p": _flat(params), "steps": np.int64(steps), "hp": np.asarray(list(hp.values()), np.float64)}
        assert np.array_equal(out["p"], _flat(p_t)), f"{name}: torch restatement differs from the reference"

        def close(a, b):
            return np.max(np.abs(a - b)) <= 1e-6 * np.max(np.abs(a))
        if name == "gd":
            assert np.array_equal(out["p"], np.concatenate(p_np)), "gd: numpy restatement differs from the reference"
        else:
            out["m"], out["v"] = _flat(opt.momentums), _flat(opt.velocities)
            assert np.array_equal(out["m"], _flat(m_t)) and np.array_equal(out["v"], _flat(v_t))
            assert np.array_equal(out["m"], np.concatenate(m_np)) and np.array_equal(out["v"], np.concatenate(v_np))
            assert close(out["p"], np.concatenate(p_np)), "adam: numpy restatement is off by more than 1e-6"
        np.savez_compressed(os.path.join(OUT, f"opt_{name}.npz"), **out)
        print("lab optimizer", name, "ok")


def lab_optimizer_grad(n, s, t):
    """The gradients make_lab_optimizers injects (shared with the tests)."""
    from oracle import synth
    g = synth.dense_grad(n, 0, s, t)
    if t == 11:
        g[::3] = 0.0
    return g


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--lab-optimizers-only" in sys.argv:
        return make_lab_optimizers()
    make_sampler()
    make_labnet(2)
    make_labnet(4)
    for v in ("A", "B", "C", "D"):
        make_grid(v, 2)
    for ws in (1, 4, 8):
        make_grid("A", ws)
    make_grid("A", 4, kind="dense")
    make_grid("A", 8, kind="dense")
    make_bf16_hook(2)
    make_bf16_hook(8)
    make_lab_optimizers()


if __name__ == "__main__":
    main()
