"""The C-ABI library loads and exports exactly what include/b200ddp.h declares.
No compute calls: this runs without a GPU."""
import ctypes
import os
import re

import pytest

import b200ddp
from b200ddp import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b200ddp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ddp_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(built_lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(built_lib, n), f"{n} is declared in include/b200ddp.h but not exported"


def test_binding_table_matches_header(built_lib):
    assert sorted(_lib._SIGNATURES) == _declared()


def test_header_cites_reference_lines():
    src = open(os.path.join(ROOT, "include", "b200ddp.h")).read()
    for cite in ("codes/task2/dist_utils.py:39-42", "codes/task2/model.py:131,62", "codes/task2/model.py:124",
                 "codes/task3/sampler.py:5-25", "codes/task2/dist_utils.py:6-30"):
        assert cite in src


def test_version_and_error_channel(built_lib):
    assert built_lib.b200ddp_version() >= 100
    h = ctypes.c_void_p()
    rc = built_lib.b200ddp_create(3, 2, 0, 1 << 20, ctypes.byref(h))   # rank outside the world: rejected before any CUDA call
    assert rc == -1 and h.value is None
    assert "rank 3" in _lib.last_error()
    rc = built_lib.b200ddp_create(0, 9, 0, 1 << 20, ctypes.byref(h))
    assert rc == -1 and "world size 9" in _lib.last_error()


@pytest.mark.parametrize("n,world,drop_last", [(0, 1, False), (1, 8, False), (5, 8, True), (60000, 2, False),
                                               (10007, 4, True), (10007, 4, False), (1281167, 8, False),
                                               (3, 8, True), (16, 8, True)])
def test_sampler_sizes_match_torch(built_lib, n, world, drop_last):
    from torch.utils.data.distributed import DistributedSampler

    class DS:
        def __len__(self):
            return n

    ref = DistributedSampler(DS(), num_replicas=world, rank=0, drop_last=drop_last)
    assert b200ddp.sampler_sizes(n, world, drop_last) == (ref.num_samples, ref.total_size)


def test_no_cpu_fallback_without_a_gpu(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(b200ddp.B200DDPError):
        b200ddp.Context(0, 1)
    with pytest.raises(b200ddp.B200DDPError):
        b200ddp.default_context()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "distributed-machine-learning-experiment-document_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower(), f"{f} mentions the oracle"


def test_header_is_plain_c():
    """The boundary is a C ABI: the header must compile as C99 (no torch, no C++ in the signatures)."""
    import subprocess
    hdr = os.path.join(ROOT, "include", "b200ddp.h")
    for cmd in (["gcc", "-x", "c", "-std=c99", "-Wall", "-pedantic", "-Werror", "-fsyntax-only", hdr],
                ["g++", "-x", "c++", "-std=c++11", "-fsyntax-only", hdr]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    text = open(hdr).read()
    assert "torch" not in re.sub(r"/\*.*?\*/", "", text, flags=re.S)
