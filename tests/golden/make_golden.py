#!/usr/bin/env python
"""tests/golden/make_golden.py -- generates the golden vectors under tests/golden/ from the REFERENCE'S OWN CODE:
oracle/_ref/libgvref_cpu.so is the reference's voxel / collider / BitVector / KinematicLink sources compiled unmodified
for the host from /root/reference (oracle/ref_cpu_driver.cpp, oracle/Makefile). Run in the authoring container (where
/root/reference is mounted):

    make -C oracle && python tests/golden/make_golden.py

The .npz files are committed; tests/test_golden_fixtures.py checks the CPU oracle against them without the reference
(`-m "not gpu"`), and tests/test_gpu_golden.py checks the CUDA path against them on the GPU box, where neither
/root/reference nor (necessarily) oracle/_ref exists. Inputs are seeded; the script is deterministic."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
BIT, PROB = 0, 4


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def random_maps(seed, n):
    rng = np.random.Generator(np.random.PCG64(seed))
    bit = np.zeros((n, 32), np.uint8)
    for k in np.nonzero(rng.uniform(size=n) < 0.35)[0]:
        for b in rng.integers(0, 256, size=rng.integers(1, 4)):
            bit[k, b >> 3] |= 1 << (b & 7)
    bit[rng.uniform(size=n) < 0.1, 0] |= 1
    prob = rng.integers(-128, 128, size=n).astype(np.int8)
    return bit.reshape(-1), prob


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgvref_cpu.so"))
    ref.gvrefcpu_collide_map_map.restype = C.c_uint64
    ref.gvrefcpu_collide_map_map.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, C.c_float, C.c_int64, C.c_int]
    ref.gvrefcpu_collide_bit_prob_types.restype = C.c_uint64
    ref.gvrefcpu_collide_bit_prob_types.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_float, C.c_int, C.c_void_p]
    ref.gvrefcpu_float_to_probability.argtypes = [C.c_float]
    ref.gvrefcpu_map_to_voxels.argtypes = [C.c_float, C.c_void_p, C.c_void_p]
    ref.gvrefcpu_list_entry_hit.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p]

    # 1. colliders: DefaultCollider on every map-type pair, thresholds, offsets; collideWithTypes (SVCollider)
    dims = (16, 12, 10)
    n = dims[0] * dims[1] * dims[2]
    bit_a, prob_a = random_maps(1, n)
    bit_b, prob_b = random_maps(2, n)
    out = {"dims": np.array(dims, np.uint32), "bit_a": bit_a, "prob_a": prob_a, "bit_b": bit_b, "prob_b": prob_b}
    cases = []
    for ta, tb in ((BIT, PROB), (PROB, BIT), (PROB, PROB), (BIT, BIT)):
        for thr in (1.0, 0.5, 0.1, 0.0):
            for off in (0, 1, -dims[0], dims[0] * dims[1] * 3 + 5):
                a = (bit_a if ta == BIT else prob_a.view(np.uint8)).copy()
                b = bit_b if tb == BIT else prob_b.view(np.uint8)
                cnt = ref.gvrefcpu_collide_map_map(vp(a), ta, vp(b), tb, n, thr, off, 1)
                key = "collide_%d_%d_%g_%d" % (ta, tb, thr, off)
                out[key + "_marked"] = a
                cases.append((key, ta, tb, thr, off, cnt))
    out["collide_cases"] = np.array([(k, ta, tb, thr, off, cnt) for k, ta, tb, thr, off, cnt in cases], dtype=object).astype(str)
    for thr in (1.0, 0.5, 0.3):
        a = bit_a.copy()
        mask = np.zeros(32, np.uint8)
        cnt = ref.gvrefcpu_collide_bit_prob_types(vp(a), vp(prob_b), n, thr, 1, vp(mask))
        out["types_%g_count" % thr] = np.array([cnt], np.uint64)
        out["types_%g_mask" % thr] = mask
        out["types_%g_marked" % thr] = a
    np.savez_compressed(os.path.join(HERE, "colliders.npz"), **out)

    # 2. scalar conversions
    thr = np.concatenate([np.linspace(-0.5, 1.5, 401), [0.1, 0.3, 0.5, 0.999, 1.0]]).astype(np.float32)
    f2p = np.array([ref.gvrefcpu_float_to_probability(float(t)) for t in thr], np.int32)
    rng = np.random.Generator(np.random.PCG64(3))
    sides = np.array([0.01, 0.025, 0.0173, 1.0, 0.3], np.float32)
    pts = rng.uniform(0, 3.0, size=(1500, 3)).astype(np.float32)
    kk = rng.integers(0, 200, size=(300, 3)).astype(np.float32)
    vox = {}
    for s in sides:
        p = np.concatenate([pts, kk * s, np.nextafter(kk * s, np.float32(-1)).clip(0, None)]).astype(np.float32)
        o = np.zeros((len(p), 3), np.uint32)
        for i in range(len(p)):
            ref.gvrefcpu_map_to_voxels(float(s), vp(p[i]), vp(o[i]))
        vox["pts_%g" % s] = p
        vox["vox_%g" % s] = o
    np.savez_compressed(os.path.join(HERE, "conversions.npz"), thr=thr, f2p=f2p, sides=sides, **vox)

    # 3. swept-volume window check and shift (GV/helpers/BitVector.h:361-471)
    m = 1500
    v1 = np.zeros((m, 32), np.uint8); v2 = np.zeros((m, 32), np.uint8)
    margin = np.zeros(m, np.uint32); svo = np.zeros(m, np.uint32); shift = np.zeros(m, np.uint32)
    coll_in = rng.integers(0, 256, size=(m, 32)).astype(np.uint8)
    coll_out = coll_in.copy(); res = np.zeros(m, np.uint8); shifted = np.zeros((m, 32), np.uint8)
    coll_fresh = np.zeros((m, 32), np.uint8); res_fresh = np.zeros(m, np.uint8)  # from an empty result vector
    for i in range(m):
        d = rng.choice([0.02, 0.1, 0.5])
        v1[i] = np.packbits(rng.uniform(size=256) < d, bitorder="little")
        v2[i] = np.packbits(rng.uniform(size=256) < d, bitorder="little")
        margin[i] = rng.integers(0, 32)
        svo[i] = rng.integers(0, 8 * (32 + margin[i])) if i % 3 == 0 else rng.integers(0, 24)
        res[i] = ref.gvrefcpu_bit_margin_check(vp(v1[i]), vp(v2[i]), vp(coll_out[i]), int(margin[i]), int(svo[i]))
        res_fresh[i] = ref.gvrefcpu_bit_margin_check(vp(v1[i]), vp(v2[i]), vp(coll_fresh[i]), int(margin[i]), int(svo[i]))
        shift[i] = rng.integers(0, 57)
        shifted[i] = v1[i]
        ref.gvrefcpu_shift_left(vp(shifted[i]), int(shift[i]))
    np.savez_compressed(os.path.join(HERE, "sweptvolume.npz"), v1=v1, v2=v2, margin=margin, sv_offset=svo, coll_in=coll_in,
                        coll_out=coll_out, result=res, coll_fresh=coll_fresh, result_fresh=res_fresh, shift=shift, shifted=shifted)

    # 4. DH chain (GV/robot/dh_robot/KinematicLink.cu:30-84, KinematicChain.cu:104-124, cuda_matrices.h:772-803)
    nl, nc = 7, 300
    dh = rng.uniform(-0.5, 0.5, size=(nl, 4)).astype(np.float32)
    dh[:, 3] = rng.choice(np.array([0, np.pi / 2, -np.pi / 2, np.pi, 1.1], np.float32), nl)
    jt = np.array([0, 0, 1, 0, 0, 1, 0], np.int32)
    jv = rng.uniform(-7, 7, size=(nc, nl)).astype(np.float32)
    jv[:10] = np.array([0.0, -0.0, np.pi / 2, -np.pi / 2, np.pi, 1e-30, 1e5, -12345.678, 1e-45, 0.78539816], np.float32)[:, None]
    tr = np.empty((nc, nl, 4, 4), np.float32)
    ref.gvrefcpu_dh_chain(vp(dh), vp(jt), nl, vp(jv), nc, vp(tr))
    np.savez_compressed(os.path.join(HERE, "dh_chain.npz"), dh4=dh, joint_types=jt, joint_values=jv, transforms_bits=tr.view(np.uint32))

    # 5. list-entry predicates of the list x map kernels (GV/voxellist/kernels/VoxelListOperations.hpp:139-150,207-216,282-292)
    nle = 3000
    lv, _ = random_maps(5, nle)
    mb, mp = random_maps(6, nle)
    mask = np.zeros(32, np.uint8); mask[1] = 0xF0; mask[4] = 0xFF; mask[20] = 0x0F
    hits = {}
    for mt, mm, vb in ((PROB, mp.view(np.uint8), 1), (BIT, mb, 32)):
        for thr_ in (1.0, 0.5, 30.0, -5.0):
            for mode in (0, 1, 2):
                h = np.array([ref.gvrefcpu_list_entry_hit(mode, vp(lv[32 * i:]), vp(mm[vb * i:]), mt, thr_, vp(mask))
                              for i in range(nle)], np.uint8)
                hits["hit_%d_%g_%d" % (mt, thr_, mode)] = h
    np.savez_compressed(os.path.join(HERE, "list_predicates.npz"), list_voxels=lv, map_bit=mb, map_prob=mp, mask=mask, **hits)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
