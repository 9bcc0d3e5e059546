"""
Build libswarm_b200.so (hand-written CUDA, sm_100a only) in-tree with nvcc.

    python -m swarmrl_b200.build [--force]

The shared object is git-ignored but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import pathlib
import shutil
import subprocess
import sys

PKG = pathlib.Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libswarm_b200.so"
SOURCES = ["swarm_engine.cu", "swarm_obs.cu", "swarm_dd.cu", "swarm_policy.cu"]
HEADERS = sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "swarm_b200.h"]   # every header: a stale library is worse than a slow build

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and pathlib.Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: libswarm_b200.so cannot be built")


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = [CSRC / s for s in SOURCES if (CSRC / s).exists()] + HEADERS
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False, timing: bool = False, variant: bool = False) -> pathlib.Path:
    """timing=True builds libswarm_b200_timing.so with -DSW_TIMING (clock64 phase counters in the resident
    kernel; select it with SWARM_B200_LIB=<path>) -- a profiling variant, never the default.  variant=True builds
    libswarm_b200_var_<tag>.so with the -D flags of the command line only (A/B runs of a compile-time choice)."""
    if timing or variant:
        extra = [a for a in sys.argv if a.startswith("-D")]
        tag = "".join(a[2:].lower() for a in extra)
        out = PKG / (f"libswarm_b200_timing{('_' + tag) if tag else ''}.so" if timing else f"libswarm_b200_var_{tag}.so")
        srcs = [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
        res = subprocess.run([_nvcc(), *NVCC_FLAGS, *(["-DSW_TIMING"] if timing else []), *extra, "-o", str(out), *srcs], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + res.stderr[-4000:])
        return out
    if not force and not needs_build():
        return LIB
    srcs = [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", str(LIB), *srcs]
    res = subprocess.run(cmd, capture_output=True, text=True)
    (PKG / "build.log").write_text(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stderr[-4000:])
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True, timing="--timing" in sys.argv, variant="--variant" in sys.argv)
    print("built", p)
