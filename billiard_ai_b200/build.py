"""Builds billiard_ai_b200/libbilliard_cuda.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.

    python -m billiard_ai_b200.build          # or: from billiard_ai_b200.build import build; build()

-fmad=false is not an optimisation choice: the reference's results are only reproducible with
un-contracted IEEE arithmetic.  Explicit fused operations in the code use fma().
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libbilliard_cuda.so")
SOURCES = ["bcu_kernels.cu", "bcu_pipeline.cu", "bcu_topk.cu", "bcu_merge.cu", "bcu_plan.cu", "bcu_api.cu", "bcu_rules.cpp", "bcu_search.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-shared", "-ldl",
]


def nvcc_path():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


OBJDIR = os.path.join(HERE, "_build")


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "billiard_cuda.h")]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(d) > t for d in _deps())


def build(force=False, verbose=False, extra=()):
    """One object per translation unit (compiled in parallel, rebuilt only when a dependency is newer), then one link."""
    if not force and not needs_build():
        return OUT
    os.makedirs(OBJDIR, exist_ok=True)
    newest = max(os.path.getmtime(d) for d in _deps() if not d.endswith((".cu", ".cpp")))
    flags = [f for f in NVCC_FLAGS if f not in ("-shared", "-ldl")]
    procs = []
    objs = []
    for src in SOURCES:
        obj = os.path.join(OBJDIR, os.path.splitext(src)[0] + ".o")
        objs.append(obj)
        spath = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(newest, os.path.getmtime(spath)):
            continue
        cmd = [nvcc_path()] + flags + list(extra) + ["-c", spath, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, pr in procs:
        if pr.wait() != 0:
            raise subprocess.CalledProcessError(pr.returncode, cmd)
    subprocess.check_call([nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT] + objs + ["-ldl"])
    return OUT


# Guard variants of the reach bound (csrc/bcu_items.cuh): each threshold x 10 and / 10, and builds with one guard removed.
# tests/test_gpu_prune_regressions.py runs the committed regression vectors through them: the weakened builds must FAIL on the
# vectors (so the vectors demonstrably exercise the guards), the default and the x / 10 builds must pass (the margin).
VARIANT_DIR = os.path.join(HERE, "_variants")
GUARD_VARIANTS = {
    "qrel0": ["-DPRUNE_MIN_QREL_V=0.0f"],
    "restall": ["-DPRUNE_REST_PASS_ONLY_V=0"],
    "qrel_x10": ["-DPRUNE_MIN_QREL_V=1e-1f"], "qrel_d10": ["-DPRUNE_MIN_QREL_V=1e-3f"],
    "sin_x10": ["-DPRUNE_MIN_SIN_SQ_V=1e-3f"], "sin_d10": ["-DPRUNE_MIN_SIN_SQ_V=1e-5f"],
    "lead_x10": ["-DPRUNE_MIN_LEADING_V=1e-4f"], "lead_d10": ["-DPRUNE_MIN_LEADING_V=1e-6f"],
}


def variant_path(name):
    return os.path.join(VARIANT_DIR, f"lib_{name}.so")


def build_variants(names=None, force=False):
    """Only bcu_pipeline.cu depends on the guard macros: one object per variant (in parallel), linked with the default
    build's other objects.  -> {name: path}"""
    build(force=False)
    os.makedirs(VARIANT_DIR, exist_ok=True)
    names = list(GUARD_VARIANTS) if names is None else list(names)
    flags = [f for f in NVCC_FLAGS if f not in ("-shared", "-ldl")]
    src = os.path.join(CSRC, "bcu_pipeline.cu")
    newest = max(os.path.getmtime(d) for d in _deps())
    procs = []
    for name in names:
        obj = os.path.join(OBJDIR, f"variant_{name}.o")
        if not force and os.path.exists(variant_path(name)) and os.path.getmtime(variant_path(name)) > newest:
            continue
        cmd = [nvcc_path()] + flags + GUARD_VARIANTS[name] + ["-c", src, "-o", obj]
        procs.append((name, obj, cmd, subprocess.Popen(cmd, stderr=subprocess.DEVNULL)))
    for name, obj, cmd, pr in procs:
        if pr.wait() != 0:
            raise subprocess.CalledProcessError(pr.returncode, cmd)
        others = [os.path.join(OBJDIR, os.path.splitext(s)[0] + ".o") for s in SOURCES if s != "bcu_pipeline.cu"]
        subprocess.check_call([nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", variant_path(name), obj] + others + ["-ldl"])
    return {n: variant_path(n) for n in names}


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(OUT)
    if "--variants" in sys.argv:
        print(build_variants(force="--force" in sys.argv))
