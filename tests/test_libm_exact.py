"""CPU tier: the product's per-lane arithmetic (billiard_ai_b200/csrc/bcu_{libm,math,quartic,physics}.cuh), compiled
for the host by tests/host_harness.cpp with g++ -ffp-contract=off, must reproduce

  * glibc 2.39's exp / log / log1p / pow / atan2 / sincos / hypot  BIT FOR BIT (the routines the reference's
    std::complex quartic resolves to), and
  * the reference's own physics::timeToCollision / solveQuarticFormula (oracle/_ref) bit for bit,

without ever leaving the restated code (fallback counter == 0).  The same headers are what the CUDA kernel compiles
with -fmad=false, so agreement here + IEEE device arithmetic is what makes the GPU results bit-identical.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle_lib import _ttc_batch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host_harness.cpp")
OUT = os.path.join(ROOT, "tests", "_build", "libhost_harness.so")


def _load_harness():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    deps = [SRC] + [os.path.join(ROOT, "billiard_ai_b200", "csrc", f)
                    for f in os.listdir(os.path.join(ROOT, "billiard_ai_b200", "csrc")) if f.endswith((".cuh", ".inc"))]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-mfma", "-fPIC", "-shared", "-x", "c++",
                               SRC, "-o", OUT])
    lib = C.CDLL(OUT)
    lib.hh_fallback_count.restype = C.c_longlong
    lib.hh_fallback_count(1)
    return lib


@pytest.fixture(scope="module")
def hh():
    return _load_harness()


def _P(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _both(lib, which, x, y=None):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x) if y is None else np.ascontiguousarray(y, dtype=np.float64)
    res = []
    for fn in (lib.hh_unary_batch, lib.hh_glibc_batch):
        o, o2 = np.zeros_like(x), np.zeros_like(x)
        fn(C.c_int(which), C.c_int64(len(x)), _P(x), _P(y), _P(o), _P(o2))
        res.append((o, o2))
    return res


N = 400_000
_rng = np.random.RandomState(1)
_f32 = lambda a: a.astype(np.float32).astype(np.float64)
CASES = [
    ("exp wide", 0, lambda r: (r.uniform(-40, 40, N), None)),
    ("exp small", 0, lambda r: (r.uniform(-1, 1, N) * 10.0 ** r.uniform(-20, 0, N), None)),
    ("log wide", 1, lambda r: (10.0 ** r.uniform(-30, 30, N), None)),
    ("log near 1", 1, lambda r: (1 + r.uniform(-0.07, 0.07, N), None)),
    ("log1p", 2, lambda r: (r.uniform(-0.9, 3, N), None)),
    ("log1p small", 2, lambda r: (r.uniform(-1, 1, N) * 10.0 ** r.uniform(-12, 0, N), None)),
    ("log1p large", 2, lambda r: (10.0 ** r.uniform(-3, 6, N), None)),
    ("pow cube root", 3, lambda r: (10.0 ** r.uniform(-30, 30, N), np.full(N, 1.0 / 3.0))),
    ("pow cube of float", 3, lambda r: (_f32(10.0 ** r.uniform(-6, 3, N)) * r.choice([-1.0, 1.0], N), np.full(N, 3.0))),
    ("pow general", 3, lambda r: (10.0 ** r.uniform(-5, 5, N), r.uniform(-5, 5, N))),
    ("atan2 scales", 4, lambda r: (r.uniform(-1, 1, N) * 10.0 ** r.uniform(-10, 10, N),
                                   r.uniform(-1, 1, N) * 10.0 ** r.uniform(-10, 10, N))),
    ("atan2 square", 4, lambda r: (r.uniform(-1, 1, N), r.uniform(-1, 1, N))),
    ("sincos third of (-pi, pi]", 5, lambda r: (r.uniform(-np.pi / 3, np.pi / 3, N), None)),
    ("sincos to 1.43", 5, lambda r: (r.uniform(-1.43, 1.43, N), None)),
    ("sincos small", 5, lambda r: (r.uniform(-1, 1, N) * 10.0 ** r.uniform(-10, -0.5, N), None)),
    ("hypot", 6, lambda r: (r.uniform(-1, 1, N) * 10.0 ** r.uniform(-10, 10, N),
                            r.uniform(-1, 1, N) * 10.0 ** r.uniform(-10, 10, N))),
]


def check_libm_functions(n):
    """Every case above on n arguments each (used by the GPU tier to check the box's own glibc)."""
    global N
    lib = _load_harness()
    saved, N = N, n
    try:
        for name, which, gen in CASES:
            x, y = gen(np.random.RandomState(77 + which * 17 + len(name)))
            lib.hh_fallback_count(1)
            (a, a2), (b, b2) = _both(lib, which, x, y)
            assert lib.hh_fallback_count(1) == 0, name
            assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), name
            if which == 5:
                assert np.array_equal(a2.view(np.uint64), b2.view(np.uint64)), name
    finally:
        N = saved


@pytest.mark.parametrize("name,which,gen", CASES, ids=[c[0] for c in CASES])
def test_elementary_functions_match_glibc_bit_for_bit(hh, name, which, gen):
    x, y = gen(np.random.RandomState(1000 + which * 17 + len(name)))
    hh.hh_fallback_count(1)
    (a, a2), (b, b2) = _both(hh, which, x, y)
    assert hh.hh_fallback_count(1) == 0, "left the restated code"
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    if which == 5:
        assert np.array_equal(a2.view(np.uint64), b2.view(np.uint64))


def test_pow_special_bases(hh):
    x = np.array([0.0, -0.0, -2.0, -1e-6, -1.5, 3.0, 0.0], dtype=np.float64)
    y = np.array([3.0, 3.0, 3.0, 3.0, 2.0, 3.0, 1 / 3.0], dtype=np.float64)
    (a, _), (b, _) = _both(hh, 3, x, y)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def _pairs(rng, n, near):
    p1 = np.stack([rng.uniform(-900, 900, n), rng.uniform(-440, 440, n)], 1).astype(np.float32)
    if near:
        d = rng.uniform(53, 400, n)
        a = rng.uniform(0, 2 * np.pi, n)
        p2 = (p1 + np.stack([d * np.cos(a), d * np.sin(a)], 1)).astype(np.float32)
    else:
        p2 = np.stack([rng.uniform(-900, 900, n), rng.uniform(-440, 440, n)], 1).astype(np.float32)
    dirv = (p2 - p1).astype(np.float64)
    ang = np.arctan2(dirv[:, 1], dirv[:, 0]) + rng.normal(0, 0.15, n)
    sp = 10.0 ** rng.uniform(0.5, 3.7, n)
    v1 = np.stack([sp * np.cos(ang), sp * np.sin(ang)], 1).astype(np.float32)
    moving2 = rng.rand(n) < 0.4
    a2 = rng.uniform(0, 2 * np.pi, n)
    sp2 = 10.0 ** rng.uniform(0.5, 3.7, n) * moving2
    v2 = np.stack([sp2 * np.cos(a2), sp2 * np.sin(a2)], 1).astype(np.float32)
    r1 = (rng.rand(n) < 0.6).astype(np.int32)
    r2 = (rng.rand(n) < 0.6).astype(np.int32)
    return p1, v1, r1, p2, v2, r2


@pytest.mark.parametrize("near", [True, False])
def test_time_to_collision_matches_reference_bit_for_bit(hh, ref, near):
    """physics::timeToCollision (billiard_physics.cpp:283-381) incl. the complex<double> quartic."""
    args = _pairs(np.random.RandomState(5 + near), 500_000, near)
    tr, hr = ref.time_to_collision_batch(*args, 52.3)
    hh.hh_fallback_count(1)
    th, hhas = _ttc_batch(hh.hh_time_to_collision_batch, *args, 52.3)
    assert hh.hh_fallback_count(1) == 0, "left the restated libm"
    assert hr.sum() > 30_000
    assert np.array_equal(hr, hhas)
    assert np.array_equal(tr[hr == 1].view(np.uint32), th[hr == 1].view(np.uint32))


def test_quartic_known_answers_and_random_roots(hh, ref):
    """solveQuarticFormula KATs (billiard_physics_test.cpp:84-143) and quantised random coefficient sets: all four
    complex roots equal the reference's, bit for bit (as sets: the reference sorts them)."""
    rng = np.random.RandomState(3)
    n = 200_000
    coef = np.floor(rng.uniform(-1, 1, (n, 5)) * 10.0 ** rng.uniform(-3, 1, (n, 5)) * 1e6) / 1e6
    coef[:, 0] = np.abs(coef[:, 0]) + 1e-6
    coef = coef.astype(np.float32).astype(np.float64)
    coef[:3] = [[1, -4, 0, 8, 1], [1, -4, 3, 8, 1], [0.009604, -0.0274841, 0.0327719, -0.018757, 0.00174089]]
    want, cnt = ref.quartic_batch(coef)
    # degenerate sets (Q == 0 -> division by zero -> NaN/inf roots, no near-real root on either side) are excluded
    nan = ~np.isfinite(want).all(axis=(1, 2))
    assert nan.mean() < 0.01
    coef, want, cnt = np.ascontiguousarray(coef[~nan]), want[~nan], cnt[~nan]
    n = len(coef)
    got = np.zeros((n, 4, 2))
    gcnt = np.zeros(n, np.int32)
    hh.hh_fallback_count(1)
    hh.hh_quartic_batch(C.c_int64(n), _P(coef), _P(got), gcnt.ctypes.data_as(C.POINTER(C.c_int32)))
    assert hh.hh_fallback_count(1) == 0
    assert np.array_equal(cnt, gcnt)
    key = lambda a: np.sort(a.view(np.uint64).reshape(n, 4, 2)[..., 0] ^ (a.view(np.uint64).reshape(n, 4, 2)[..., 1] * 3), axis=1)
    assert np.array_equal(key(got), key(want))
    assert abs(want[2, 0, 0] - 0.11314) < 1e-4
