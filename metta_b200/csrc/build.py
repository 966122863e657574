"""Builds libmettagrid_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch dependency)."""

from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "libmettagrid_b200.so")
SOURCES = ["mg_abi.cu", "mg_host_order.cpp"]
DEPS = SOURCES + ["mg_kernels.cuh", "mg_device.cuh", "mg_obs.cuh", "mg_rng.cuh", "mg_state.h", "mg_platform.h",
                  os.path.join("..", "..", "include", "mg_config.h")]
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--fmad=false",  # float rewards must round like the reference's scalar code (hard part 3)
    "-Xcompiler", "-fPIC", "-shared",
]  # fmt: skip


def stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(HERE, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, out: str = OUT, extra: list[str] | None = None) -> str:
    if out == OUT and not force and not stale():
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc] + FLAGS + (extra or []) + (["-Xptxas", "-v"] if verbose else []) + ["-o", out] + SOURCES
    subprocess.check_call(cmd, cwd=HERE)
    return out


if __name__ == "__main__":
    print(build(force=True, verbose=True))
