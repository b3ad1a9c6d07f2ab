"""CPU oracle for the data-parallel gradient-sync hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the shipped package
(`distributed-machine-learning-experiment-document_b200/`, alias `b200ddp`)
imports from here.  The only permitted importers are `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py`, and there only as the checker or the timed CPU arm, never as the
product path.

Parity status: the reference ships no tests and no golden vectors
(SURVEY.md §4), so the oracle is pinned the other way the task allows:
against outputs of the reference itself, run in the build container —
`oracle/make_golden.py` imports `/root/reference/codes/task{2,3,4}/dist_utils.py`
by path, runs them under gloo, and writes `tests/golden/*.npz`; the sampler
restatement is pinned against `torch.utils.data.DistributedSampler` (the
object the reference constructs at `codes/task2/model.py:124`).
`tests/test_oracle_golden.py` replays both.
"""
