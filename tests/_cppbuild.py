"""Build helper of the C++ test programs under tests/cpp: an executable is up to date when the HASH of its sources (and
the flags) equals the stamp written beside it. Modification times are not used: a snapshot copied to another machine
need not preserve them, and a stale test program would then silently run in place of the current one."""
import hashlib
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def facade_headers():
    out = []
    for d, _, fs in os.walk(os.path.join(ROOT, "include")):  # gvx.h, gpu_voxels/, icl_core_config/
        out += [os.path.join(d, f) for f in fs]
    return sorted(out)


def build_cpp(src, exe, deps=(), opt="-O1", link_gvx=True):
    """compile src -> exe unless exe was built from exactly these files with exactly this command; returns exe"""
    cmd = [CXX, "-std=c++14", opt, "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"), src]
    if link_gvx:
        from gpu_voxels_b200 import build as gvx_build
        gvx_build.build()
        cmd += ["-L" + os.path.join(ROOT, "gpu_voxels_b200"), "-lgvx", "-Wl,-rpath," + os.path.join(ROOT, "gpu_voxels_b200"),
                "-Wl,-rpath,$ORIGIN/../../gpu_voxels_b200", "-lpthread"]
    cmd += ["-o", exe]
    h = hashlib.sha256(" ".join(c.replace(ROOT, "") for c in cmd).encode())
    for f in [src] + sorted(deps):
        h.update(os.path.relpath(f, ROOT).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    digest, stamp = h.hexdigest(), exe + ".stamp"
    if os.path.exists(exe) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        return exe
    subprocess.check_call(cmd)
    with open(stamp, "w") as fh:
        fh.write(digest + "\n")
    return exe
