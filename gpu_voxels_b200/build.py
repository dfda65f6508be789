"""Builds libgvx.so (the C-ABI library) in-tree with nvcc for sm_100a.

    python -m gpu_voxels_b200.build [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgvx.so")
SOURCES = ["gvx_api.cu", "gvx_api_list.cu", "gvx_api_batch.cu", "gvx_collide.cu", "gvx_insert.cu", "gvx_list.cu",
           "gvx_batch.cu", "gvx_kinematics.cu", "gvx_sweptvolume.cu", "gvx_listlist.cu", "gvx_api_io.cu",
           "gvx_dispatch.cu", "gvx_api_dispatch.cu", "gvx_api_files.cu", "gvx_collide_tma.cu", "gvx_api_shard.cu", "gvx_noalloc.cu", "gvx_api_lanes.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--ftz=false", "--prec-div=true", "--prec-sqrt=true", "--fmad=true",
              "-Xcompiler", "-ffp-contract=off"]  # host code: no FMA contraction (file loaders, DH re-computation)


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "gvx.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    objs = []
    for s in SOURCES:
        o = os.path.join(CSRC, s.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, s), "-o", o]
        subprocess.check_call(cmd)
        objs.append(o)
    # the arch on the link line too: without it nvcc adds an (empty) element for its default architecture to the fatbinary
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
