"""The C-ABI library loads and exports every symbol include/gvx.h declares. No compute calls (no GPU needed)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "gvx.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gvx_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from gpu_voxels_b200 import _lib
    L = _lib.load()
    names = _declared()
    assert len(names) >= 26
    for n in names:
        assert hasattr(L, n), "libgvx.so does not export " + n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_no_gpu_fails_loudly():
    """Without a GPU gvx_init must return an error (never exit(), never fall back to a CPU path)."""
    import torch
    from gpu_voxels_b200 import _lib
    if torch.cuda.is_available():
        return
    L = _lib.load()
    rc = L.gvx_init(-1)
    assert rc != 0 and b"CUDA" in L.gvx_last_error()
    h = _lib.C.c_uint64(0)
    assert L.gvx_map_create(_lib.C.byref(h), 0, 8, 8, 8, 1.0) == _lib.ERR_NOT_INITIALIZED


def test_product_never_imports_oracle():
    """The product package must not reference oracle/ in any way."""
    pkg = os.path.join(ROOT, "gpu_voxels_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "gvoracle" not in src and "oracle." not in src and "libgvref" not in src, f


def test_python_constants_equal_the_header():
    """every enumerator of include/gvx.h that the Python binding names has the header's value (status codes, memory
    kinds, map types, options)"""
    from gpu_voxels_b200 import _lib, api
    txt = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "gvx.h")).read(), flags=re.S)
    enum = {k: int(v) for k, v in re.findall(r"\b(GVX_[A-Z0-9_]+)\s*=\s*(-?\d+)", txt)}
    assert len(enum) >= 18
    checked = 0
    for name, value in enum.items():
        short = name[4:]
        for mod in (_lib, api):
            for attr in (name, short):
                if hasattr(mod, attr):
                    assert getattr(mod, attr) == value, (mod.__name__, attr)
                    checked += 1
    assert checked >= 14
    # the facade's map types are the reference's on-disk tags (GV/helpers/common_defines.h:53-74)
    assert (enum["GVX_MT_BITVECTOR_VOXELMAP"], enum["GVX_MT_BITVECTOR_VOXELLIST"], enum["GVX_MT_PROBAB_VOXELMAP"]) == (0, 1, 4)
