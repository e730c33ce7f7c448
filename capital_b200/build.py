"""Build recipe: nvcc -> capital_b200/libcapital_b200.so (sm_100a only, in-tree so it travels with gpurun)."""
import os
import subprocess
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcapital_b200.so")
SOURCES = ["capi.cu", "comm.cu"]
def _headers():
    hs = [f for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    return hs + [os.path.join("..", "..", "include", "capital_b200.h")]


def nccl_dirs():
    site = sysconfig.get_paths()["purelib"]
    inc, lib = os.path.join(site, "nvidia", "nccl", "include"), os.path.join(site, "nvidia", "nccl", "lib")
    if os.path.exists(os.path.join(lib, "libnccl.so.2")):
        return inc, lib
    return "/usr/include", "/usr/lib/x86_64-linux-gnu"


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + _headers())


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    inc, lib = nccl_dirs()
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
           "-Xcompiler", "-fPIC", "-I" + inc] + [os.path.join(CSRC, s) for s in SOURCES] + \
          ["-o", LIB, "-L" + lib, "-l:libnccl.so.2", "-Xlinker", "-rpath=" + lib]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB


HOST_BIN = os.path.join(HERE, "host", "bin")
HOST_MAINS = ["cacqr", "cholinv", "summa_gemm", "potrf"]


def build_host(force=False):
    """C++ host drivers (argv-compatible with the reference's bench mains) linked against the C-ABI library."""
    os.makedirs(HOST_BIN, exist_ok=True)
    hdr = [os.path.join(HERE, "host", "capital.hpp"), os.path.join(HERE, "host", "bench", "common.hpp"), LIB]
    for name in HOST_MAINS:
        src, out = os.path.join(HERE, "host", "bench", name + ".cpp"), os.path.join(HOST_BIN, name)
        if not force and os.path.exists(out) and all(os.path.getmtime(f) <= os.path.getmtime(out) for f in hdr + [src]):
            continue
        subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-o", out, src, "-L" + HERE, "-l:libcapital_b200.so", "-lpthread",
                        "-Wl,-rpath,$ORIGIN/../..", "-Wl,-rpath," + nccl_dirs()[1], "-Wl,-rpath,/usr/local/cuda/lib64",
                        "-Wl,--allow-shlib-undefined"], check=True)
    return HOST_BIN


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
