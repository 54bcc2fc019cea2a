"""In-tree build of the native pieces (no cmake, no JIT cache):

  libopenbnsl_b200.so            nvcc, sm_100a -- CUDA kernels + the extern "C" layer
  openbnsllib<ext>.so            g++ + pybind11 -- C++ host drivers behind the reference's
                                 Python surface (base / score / citest / structure_learning)

Both land next to this file so that they travel to the GPU box with the tree.
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libopenbnsl_b200.so")
EXT = sysconfig.get_config_var("EXT_SUFFIX")
MOD = os.path.join(HERE, "openbnsllib" + EXT)

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def build_cuda(force=False, verbose=False):
    deps = (glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) +
            glob.glob(os.path.join(CSRC, "*.inl")) + glob.glob(os.path.join(CSRC, "*.hpp")) +
            [os.path.join(ROOT, "include", "openbnsl_b200.h")])
    if not force and not _stale(LIB, deps):
        return LIB
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    cmd = [_nvcc()] + NVCC_FLAGS + ["-o", LIB] + srcs
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return LIB


def build_module(force=False, verbose=False):
    host_dir = os.path.join(CSRC, "host")
    srcs = sorted(glob.glob(os.path.join(host_dir, "*.cpp")))
    if not srcs:
        return None
    deps = srcs + glob.glob(os.path.join(host_dir, "*.hpp")) + [os.path.join(ROOT, "include", "openbnsl_b200.h"), LIB]
    if not force and not _stale(MOD, deps):
        return MOD
    import pybind11
    cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-Wall",
           "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
           "-I", os.path.join(ROOT, "include"), "-I", host_dir] + srcs + [
           "-L", HERE, "-lopenbnsl_b200", "-Wl,-rpath,$ORIGIN", "-o", MOD]
    if verbose:
        print(" ".join(cmd))
    subprocess.check_call(cmd)
    return MOD


def build_all(force=False, verbose=False):
    build_cuda(force, verbose)
    build_module(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
