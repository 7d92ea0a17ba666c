"""CPU tier: the C-ABI library builds, loads and exports every symbol include/billiard_cuda.h declares; the
product refuses to run without a CUDA device (no CPU fallback); nothing under billiard_ai_b200/ touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "billiard_cuda.h")


@pytest.fixture(scope="module")
def lib_path():
    from billiard_ai_b200.build import build
    return build()


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bcu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(lib_path):
    names = _declared()
    assert len(names) >= 15
    out = subprocess.check_output(["nm", "-D", "--defined-only", lib_path]).decode()
    exported = set(re.findall(r" T (bcu_[a-z0-9_]+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, missing
    lib = C.CDLL(lib_path)
    for n in names:
        getattr(lib, n)


def test_library_contains_sm100a_code(lib_path):
    out = subprocess.check_output(["cuobjdump", "-lelf", lib_path]).decode()
    assert "sm_100a" in out


def test_no_device_means_error_not_fallback(lib_path, table):
    import torch
    from billiard_ai_b200 import Engine, BcuError
    from oracle_lib import SNOOKER_SCORES
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(BcuError, match="NO_DEVICE"):
        Engine.from_table(table, SNOOKER_SCORES)


def test_missing_library_raises(tmp_path):
    from billiard_ai_b200 import engine
    with pytest.raises(engine.BcuError, match="missing"):
        engine.load_library(str(tmp_path / "libbilliard_cuda.so"))


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "billiard_ai_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(d, f), errors="ignore").read()
                assert "oracle_lib" not in text and "libport_oracle" not in text and "libref_oracle" not in text, f
                assert "/root/reference" not in text, f


def test_workload_generators_are_deterministic():
    from billiard_ai_b200 import workloads as W
    p1, t1 = W.snooker22()
    assert p1.shape == (22, 2) and (t1 == 1).sum() == 15 and (t1 == 0).sum() == 1
    d = np.hypot(*(p1[:, None, :] - p1[None, :, :]).transpose(2, 0, 1)) + np.eye(22) * 1e9
    assert d.min() >= 52.3
    a, _ = W.random_layout(16, 16)
    b, _ = W.random_layout(16, 16)
    assert np.array_equal(a, b)
    dirs, speeds = W.grid(1024, 64)
    assert dirs.shape == (1024, 2) and speeds[0] == np.float32(78.125) and speeds[-1] == np.float32(5000.0)
    vx, vy = W.grid_velocities(dirs, speeds)
    assert vx.shape == (65536,) and vx[64 * 3 + 5] == np.float32(speeds[5] * dirs[3, 0])
    assert W.shard_range(10, 3, 4) == (9, 10) and W.shard_range(1 << 20, 7, 8) == (7 << 17, 1 << 20)
