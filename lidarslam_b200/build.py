"""In-tree build of liblidarslam_b200.so (CUDA kernels + C++ host driver + C ABI).

nvcc cross-compiles for sm_100a without a GPU.  -fmad=false / -ffp-contract=off:
the kernels must execute the same IEEE operation sequence as the CPU oracle.
Objects are cached under build/ and rebuilt when a source or header changes."""
import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "_lib", "liblidarslam_b200.so")
OBJ = os.path.join(ROOT, "build", "obj")

SOURCES = ["cuda/extract.cu", "cuda/knn.cu", "cuda/match.cu", "cuda/map.cu", "cuda/sort32.cu", "cuda/slam_kernels.cu", "cuda/lcp.cu", "cuda/convert.cu", "cuda/peaks.cu",
           "host/slam.cu", "host/capi.cu", "host/capi_stages.cu", "host/defaults.cpp", "host/pcd_io.cpp"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
         "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-fvisibility=hidden", "-Xcudafe", "--diag_suppress=177"]


def _headers():
    hs = [os.path.join(ROOT, "include", "lidarslam_b200.h")]
    for d, _, fs in os.walk(CSRC):
        hs += [os.path.join(d, f) for f in fs if f.endswith((".h", ".cuh"))]
    return sorted(hs)


def _digest(paths):
    h = hashlib.sha256(" ".join(FLAGS).encode())
    for p in paths:
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _compile(src, hdr_digest, verbose):
    path = os.path.join(CSRC, src)
    obj = os.path.join(OBJ, src.replace("/", "_") + ".o")
    stamp = obj + ".sha"
    dig = _digest([path]) + hdr_digest
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj
    cmd = [NVCC] + FLAGS + (["-x", "cu"] if src.endswith(".cpp") else []) + ["-c", path, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    with open(stamp, "w") as f:
        f.write(dig)
    return obj


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    hd = _digest(_headers()) + ("force%d" % os.getpid() if force else "")
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = list(ex.map(lambda s: _compile(s, hd, verbose), SOURCES))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-lpthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
