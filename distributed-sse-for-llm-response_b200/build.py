"""Builds libssegpu.so (CUDA kernels + C ABI) in-tree for sm_100a with nvcc.

python distributed-sse-for-llm-response_b200/build.py [--force]
The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libssegpu.so")
SOURCES = ["sse_kernels.cu", "sse_stream.cu", "sse_wire.cu", "sse_chunk.cu", "sse_encode.cu", "ssegpu_api.cu", "ssegpu_static.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "--shared", "-cudart", "shared", "--threads", "4"]


def _deps():
    d = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    d.append(os.path.join(os.path.dirname(HERE), "include", "ssegpu.h"))
    d.append(os.path.abspath(__file__))
    return d


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


LOADLIB = os.path.join(HERE, "libsseload.so")
LOADSRC = os.path.join(HERE, "host", "loadgen.cpp")


def build_host(force=False):
    """host-side C++ (no CUDA): the synthetic load generator"""
    if force or not os.path.exists(LOADLIB) or os.path.getmtime(LOADSRC) > os.path.getmtime(LOADLIB):
        subprocess.check_call([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared", "-pthread",
                               "-o", LOADLIB, LOADSRC])
    return LOADLIB


HOSTLIB = os.path.join(HERE, "libssehost.so")
HOSTSRC = os.path.join(HERE, "host", "adapter_host.cpp")


def build_adapter_host(force=False):
    """host-side C++ mirror of the adapter's per-connection flow; links against libssegpu.so (same directory)"""
    deps = [HOSTSRC, os.path.join(os.path.dirname(HERE), "include", "ssegpu.h"), LIB]
    if force or not os.path.exists(HOSTLIB) or any(os.path.getmtime(d) > os.path.getmtime(HOSTLIB) for d in deps):
        subprocess.check_call([os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared", "-o", HOSTLIB, HOSTSRC,
                               "-L" + HERE, "-lssegpu", "-Wl,-rpath,$ORIGIN"])
    return HOSTLIB


def build(force=False, verbose=False):
    build_host(force)
    if not force and not needs_build():
        build_adapter_host(force)
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    build_adapter_host(True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
