#!/usr/bin/env python
"""In-tree build of the native core.

Produces (git-ignored, but shipped to the GPU box by gpurun):
  ytk_mp4j_b200/_lib/libmp4j.so   - C++ control/host plane + sm_100a CUDA kernels, C ABI of include/mp4j.h
  ytk_mp4j_b200/_lib/mp4j-master  - stand-alone master binary (CommMaster.main equivalent)

Host sources go through g++, .cu sources through
`nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`; objects are cached under build/ and
rebuilt when a source or any header is newer. No torch headers are involved: the Python
package binds the C ABI with ctypes, so a full rebuild takes well under a minute.
"""
import concurrent.futures as cf
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(ROOT, "ytk_mp4j_b200", "_lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

HOST_SRCS = [
    "csrc/core/ops.cc",
    "csrc/control/net.cc",
    "csrc/control/master.cc",
    "csrc/control/client.cc",
    "csrc/host/process_comm.cc",
    "csrc/host/thread_group.cc",
    "csrc/capi/c_api.cc",
]
CXXFLAGS = ["-O2", "-g", "-fPIC", "-std=c++17", "-Wall", "-Wno-unused-function", "-pthread",
            "-I" + os.path.join(ROOT, "include")]
NVCCFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
             "-Xcompiler", "-fPIC", "-Xcompiler", "-pthread", "--expt-relaxed-constexpr",
             "-I" + os.path.join(ROOT, "include")]


def _cuda_srcs():
    d = os.path.join(ROOT, "csrc", "gpu")
    return sorted("csrc/gpu/" + f for f in os.listdir(d) if f.endswith(".cu")) if os.path.isdir(d) else []


def _newest_header():
    newest = 0.0
    for base in ("csrc", "include"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for f in fns:
                if f.endswith((".h", ".cuh", ".hpp")):
                    newest = max(newest, os.path.getmtime(os.path.join(dp, f)))
    return newest


def _obj(src):
    return os.path.join(OBJ_DIR, src.replace("/", "_") + ".o")


def _compile(src, hdr_mtime, verbose):
    obj = _obj(src)
    sp = os.path.join(ROOT, src)
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(sp), hdr_mtime):
        return obj, None
    if src.endswith(".cu"):
        cmd = [NVCC] + NVCCFLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", obj]
    else:
        cmd = [CXX] + CXXFLAGS + ["-c", sp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("compile failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return obj, (r.stderr if verbose else None)


def build(verbose=False, force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr = _newest_header()
    if force:
        hdr = float("inf")
    srcs = HOST_SRCS + _cuda_srcs()
    objs = []
    with cf.ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 4)) as ex:
        for obj, log in ex.map(lambda s: _compile(s, hdr, verbose), srcs):
            objs.append(obj)
            if log:
                sys.stderr.write(log)
    lib = os.path.join(OUT_DIR, "libmp4j.so")
    if force or not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in objs):
        cmd = [NVCC, "-shared", "-o", lib] + objs + ["-Xcompiler", "-pthread", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stderr))
    # stand-alone master (host objects only)
    exe = os.path.join(OUT_DIR, "mp4j-master")
    main_src = os.path.join(ROOT, "csrc/tools/mp4j_master_main.cc")
    host_objs = [_obj(s) for s in ("csrc/control/net.cc", "csrc/control/master.cc")]
    if force or not os.path.exists(exe) or os.path.getmtime(exe) < max(
            [os.path.getmtime(main_src)] + [os.path.getmtime(o) for o in host_objs]):
        cmd = [CXX] + CXXFLAGS + [main_src] + host_objs + ["-o", exe, "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stderr))
    return lib


if __name__ == "__main__":
    v = "-v" in sys.argv
    f = "-f" in sys.argv
    print(build(verbose=v, force=f))
