"""What the built library contains, read from its SASS with cuobjdump (no GPU needed): every device binary is sm_100a,
and the kernels carry the Blackwell-specific instructions DESIGN.md says they use -- TMA bulk copies with mbarrier
waits in the TMA-fed collide, programmatic-dependent-launch instructions in the two kernels of a sensor step, 256-bit
global loads in the voxel-streaming collide, the saturating float -> voxel conversion of the reference in the scatter
kernels, warp-level REDUX in the reductions. Skipped where the CUDA binary utilities are not installed."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not os.path.exists(CUOBJDUMP), reason="cuobjdump not installed")


@pytest.fixture(scope="module")
def kernels():
    from gpu_voxels_b200 import build as gvx_build
    lib = gvx_build.build()
    elfs = subprocess.run([CUOBJDUMP, "-lelf", lib], capture_output=True, text=True, check=True).stdout.split()
    archs = {m.group(1) for m in (re.search(r"\.(sm_\w+)\.cubin", e) for e in elfs) if m}
    sass = subprocess.run([CUOBJDUMP, "-sass", lib], capture_output=True, text=True, check=True).stdout
    out = {}
    for body in re.split(r"\n\s*Function : ", sass)[1:]:
        name, _, text = body.partition("\n")
        out[name.strip()] = set(re.findall(r"\b([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_]+)+|[A-Z]{4,})\b", text))
    return archs, out


def _of(kernels, fragment):
    found = {n: ops for n, ops in kernels[1].items() if fragment in n}
    assert found, fragment
    return found


def _has(ops, prefix):
    return any(o.startswith(prefix) for o in ops)


def test_every_device_binary_is_sm_100a(kernels):
    assert kernels[0] == {"sm_100a"}
    assert len(kernels[1]) > 100


def test_tma_collide_uses_bulk_copies_and_mbarriers(kernels):
    for name, ops in _of(kernels, "k_collide_tma").items():
        assert _has(ops, "UBLKCP") and _has(ops, "SYNCS.ARRIVE") and _has(ops, "SYNCS.PHASECHK"), name


def test_step_kernels_carry_programmatic_dependent_launch(kernels):
    for fragment in ("k_update_logodds", "k_collide_summary2"):
        for name, ops in _of(kernels, fragment).items():
            if "dispatch" in name:
                continue  # the peer-dispatch form is launched plainly
            assert "ACQBULK" in ops and "PREEXIT" in ops, name
    # and only kernels that wait may be launched early: no other kernel triggers its dependents
    others = [n for n, ops in kernels[1].items() if "PREEXIT" in ops and "ACQBULK" not in ops]
    assert not others, others


def test_streaming_collide_loads_whole_voxels_and_reduces_in_the_warp(kernels):
    wide = [n for n, ops in _of(kernels, "k_collide_voxelwise").items() if _has(ops, "LDG.E.NA.ENL2.256") or any(".256" in o for o in ops)]
    assert wide, "no 256-bit global loads in the voxel-streaming collide"
    for name, ops in _of(kernels, "k_collide_voxelwise").items():
        assert _has(ops, "REDUX"), name


def test_scatter_kernels_use_the_reference_float_to_voxel_conversion(kernels):
    for fragment in ("k_insert", "k_update_logodds", "k_raycast_free", "k_list_keys"):
        for name, ops in _of(kernels, fragment).items():
            assert "F2I.U32.FLOOR.NTZ" in ops, name


def test_multi_frame_fold_stages_points_with_bulk_copies_and_has_no_cta_barrier(kernels):
    """k_update_logodds_frames: per-warp rings of point chunks filled by cp.async.bulk (UBLKCP) behind mbarriers, hits
    queued in shared memory (STS/LDS) and applied with CAS -- and no BAR.SYNC: a warp that drains its queue must not
    stall the other warps of its CTA"""
    for name, ops in _of(kernels, "k_update_logodds_frames").items():
        assert _has(ops, "UBLKCP") and _has(ops, "SYNCS.ARRIVE") and _has(ops, "SYNCS.PHASECHK"), name
        assert _has(ops, "ATOMG.E.CAS") or _has(ops, "ATOM.E.CAS") or any("CAS" in o for o in ops), name
        assert not _has(ops, "BAR.SYNC"), name
        assert _has(ops, "MATCH.ANY"), name


def test_noalloc_scratch_is_a_static_device_array():
    """gvx_noalloc.cu: 16 MiB of .global, 256-byte aligned, inside the library (not a cudaMalloc)"""
    from gpu_voxels_b200 import build as gvx_build
    lib = gvx_build.build()
    ptx_or_elf = subprocess.run([CUOBJDUMP, "-elf", lib], capture_output=True, text=True).stdout
    hit = [l for l in ptx_or_elf.splitlines() if "g_noalloc_scratch" in l and "0x1000000" in l]
    assert hit, "the 16 MiB scratch symbol is missing from the device ELF"
