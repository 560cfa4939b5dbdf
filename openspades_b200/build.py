"""Builds openspades_b200/lib/libosb200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU, so this runs in the CPU container; the .so travels to the GPU
box with the repo snapshot.  Usage: python -m openspades_b200.build [--force] [--verbose]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
OUT_DIR = os.path.join(HERE, "lib")
OBJ_DIR = os.path.join(HERE, "lib", "obj")
LIB = os.path.join(OUT_DIR, "libosb200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# --fmad=false: the reference is x86-64 code without FMA contraction; parity needs separate mul/add.
CUFLAGS = ["-O3", "-std=c++17", "-lineinfo", "--fmad=false", "-Xcompiler", "-fPIC,-fvisibility=hidden",
           "-ccbin", "/usr/bin/g++"]
CUFLAGS += os.environ.get("OSB200_NVCC_EXTRA", "").split()
if os.environ.get("OSB200_LIB_OUT"):  # experiment builds: a differently named library next to the default one
    LIB = os.environ["OSB200_LIB_OUT"]
    OBJ_DIR = LIB + ".obj"
CU_SOURCES = ["api.cu", "castray2.cu", "terrain.cu", "radiosity.cu", "ambient.cu", "models.cu", "vxl.cu", "links.cu", "group.cu", "sse_calibrate.cpp"]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def build(force=False, verbose=False):
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "osb200.h"))
    objs = []
    procs = []
    for s in CU_SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ_DIR, os.path.splitext(s)[0] + ".o")
        objs.append(obj)
        if force or _newer([src] + headers, obj):
            cmd = [NVCC] + ARCH + CUFLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd), flush=True)
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"--- nvcc failed on {s} ---\n{out}\n")
        elif verbose or out.strip():
            print(out)
    if failed:
        raise RuntimeError("nvcc compilation failed")
    if force or procs or _newer(objs, LIB):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-ccbin", "/usr/bin/g++", "-ldl"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    build_host(force, verbose)
    return LIB


HOST_LIB = os.path.join(OUT_DIR, "libosb200_host.so")


def build_host(force=False, verbose=False):
    """C++ host layer (VoxelMap, B200Renderer + C wrappers) on top of libosb200.so."""
    srcs = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST)) if f.endswith(".cpp")]
    deps = srcs + [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".h")]
    deps.append(os.path.join(HERE, "..", "include", "osb200.h"))
    if force or _newer(deps + [LIB], HOST_LIB):
        cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-pthread",
               "-o", HOST_LIB] + srcs + ["-L" + OUT_DIR, "-losb200", "-Wl,-rpath,$ORIGIN"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return HOST_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
