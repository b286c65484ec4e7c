"""Generates the golden vectors in this directory from the reference's own shader source
(oracle/_ref = RTrace/raytrace.metal + sampling.metal compiled as C++; needs /root/reference
at build time).  Run from the repo root:  python tests/golden/make_golden.py
Outputs (float32 / uint32, bit-exact expectations):
  pathtrace_cornell_64x48_ref400spp.npz   image of the reference kernel at its literals
                                          (400 spp, 3 bounces), offsets seed 3, + the inputs
  pathtrace_cornell_80x60_moved_ref400spp.npz   the same kernel seen by a camera inside the room: off-axis direction,
                                          rolled up vector, 100 degrees wide; offsets seed 5
  drawtriangle_cornell_64x48_ref.npz      RGBA8 + float3 buffers of the reference MIS kernel (6 x 300)
  halton_ref.npz                          halton(i, d) of sampling.metal:107-122 for 4096 (i, d)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O   # noqa: E402  (fixtures only: scene bytes + offsets)
import ref_lib as R      # noqa: E402

ref = R.load()
assert ref is not None, "oracle/_ref cannot be built here"
W, H = 64, 48
sc = O.cornell(W, H)
off = O.offsets(W, 0, H, seed=3)
img = R.pathtrace(ref, sc, W, H, off)
np.savez_compressed(os.path.join(HERE, "pathtrace_cornell_64x48_ref400spp.npz"), image=img, offsets=off,
                    camera=sc["camera"], materials=sc["materials"], lights=sc["lights"], vertices=sc["vertices"])
# a second camera: inside the room, looking down and to the left, rolled, wide
W2, H2 = 80, 60
sc2 = O.cornell(W2, H2)
cam = sc2["camera"].view(np.float32)
cam[0:3] = (0.8, 0.6, 2.1); cam[4:7] = (-0.45, -0.35, -1.0); cam[8:11] = (0.15, 1.0, 0.05); cam[14] = 1.75
off2 = O.offsets(W2, 0, H2, seed=5)
img2 = R.pathtrace(ref, sc2, W2, H2, off2)
np.savez_compressed(os.path.join(HERE, "pathtrace_cornell_80x60_moved_ref400spp.npz"), image=img2, offsets=off2,
                    camera=sc2["camera"], materials=sc2["materials"], lights=sc2["lights"], vertices=sc2["vertices"])
# mode B: the reference MIS kernel at its literals (6 camera rays, 300 MIS samples), 1.5 x 1.5 light
scb = O.cornell(W, H, light_size=1.5)
pxb, txb = R.drawtriangle(ref, scb, W, H)
np.savez_compressed(os.path.join(HERE, "drawtriangle_cornell_64x48_ref.npz"), pixels=pxb, text=txb,
                    camera=scb["camera"], materials=scb["materials"], lights=scb["lights"], vertices=scb["vertices"])
rng = np.random.default_rng(11)
i = np.concatenate([np.arange(64), rng.integers(0, 1 << 21, 4096 - 64 - 2), [0xffffffff, 1 << 24]]).astype(np.uint32)
d = rng.integers(0, 24, i.size).astype(np.uint32)
h = np.array([ref.rt_ref_halton(int(a), int(b)) for a, b in zip(i, d)], np.float32)
np.savez_compressed(os.path.join(HERE, "halton_ref.npz"), i=i, d=d, value=h)
print("wrote", os.listdir(HERE))
