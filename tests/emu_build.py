"""Builds tests/_emu/libmg_emu.so: the simulator's kernel sources compiled for the host on top of
tests/simt_emu (a fiber-based SIMT emulator). TEST INFRASTRUCTURE: lets the CPU suite debug
kernel logic without a GPU. The package never loads this library."""

from __future__ import annotations

import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "metta_b200", "csrc")
OUT_DIR = os.path.join(HERE, "_emu")
OUT = os.path.join(OUT_DIR, "libmg_emu.so")


def build_emu(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".cpp"))]
    deps += [os.path.join(HERE, "simt_emu", f) for f in ("simt_emu.h", "simt_emu.cpp")]
    deps.append(os.path.join(ROOT, "include", "mg_config.h"))
    if not force and os.path.exists(OUT) and all(os.path.getmtime(d) <= os.path.getmtime(OUT) for d in deps):
        return OUT
    cmd = [
        "g++", "-std=c++17", "-O1", "-g", "-fPIC", "-shared", "-DMG_SIMT_EMU", "-ffp-contract=off",
        "-I", os.path.join(HERE, "simt_emu"), "-I", CSRC,
        "-x", "c++", os.path.join(CSRC, "mg_abi.cu"), os.path.join(CSRC, "mg_host_order.cpp"),
        os.path.join(HERE, "simt_emu", "simt_emu.cpp"), "-o", OUT,
    ]  # fmt: skip
    subprocess.check_call(cmd)
    return OUT


def load_emu() -> ctypes.CDLL:
    from metta_b200 import capi

    return capi._declare(ctypes.CDLL(build_emu()))
