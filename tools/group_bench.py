"""rtGroupRender (one process, N devices, peer-to-peer gather behind the C ABI) on a BASELINE config, beside what bench.py's
torch + NCCL path gives for the same frame:  python tools/group_bench.py N [config] [steps]
Prints one JSON line: Mpaths/s over wall time of the whole call (uploads, render, gather, read-back), the slowest member's
device time, and parity of the frame with the oracle on bench.py's sample rows."""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import gpuraytracer_b200 as G  # noqa: E402
from gpuraytracer_b200 import _lib as L  # noqa: E402
import oracle_lib as O  # noqa: E402
import bench  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1
cfg = sys.argv[2] if len(sys.argv) > 2 else "2"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
sys.argv = ["bench.py", "--config", cfg]
a = bench.parse()
W, H = a.W, a.H
sc = O.spheres(W, H) if a.scene == "spheres" else O.cornell(W, H)
off = O.offsets(W, 0, H)
g = C.c_void_p()
L.check(L.lib.rtGroupCreate((C.c_int * N)(*range(N)), N, C.byref(g)), "rtGroupCreate")
p = G.default_params(W, H, spp=a.spp, numTriangles=sc["T"], numSpheres=sc["S"], numLights=sc["L"])
out = np.zeros((H, W, 4), np.float32)
ms = C.c_double()


def render():
    L.check(L.lib.rtGroupRender(g, b"pathTrace", O.vp(sc["camera"]), O.vp(sc["materials"]), sc["materials"].nbytes,
                                O.vp(sc["lights"]), sc["lights"].nbytes, O.vp(sc["vertices"]), sc["vertices"].nbytes,
                                O.vp(sc["spheres"]), 0 if sc["spheres"] is None else sc["spheres"].nbytes,
                                O.vp(off), C.byref(p), O.vp(out), C.byref(ms)), "rtGroupRender")


for _ in range(2):
    render()
t0 = time.perf_counter()
dev_ms = []
for _ in range(steps):
    render()
    dev_ms.append(ms.value)
wall = (time.perf_counter() - t0) / steps
ref, cpaths, ct, cores, rows = bench.oracle_sample(a)
equal = all(np.array_equal(out[r0:r0 + rows], tile) for r0, tile in ref.items())
print(json.dumps({"what": "rtGroupRender", "n_gpus": N, "config": bench.workload(a)["workload"], "steps": steps,
                  "e2e_mpaths_s": W * H * a.spp / wall / 1e6, "ms_per_step_wall": wall * 1e3,
                  "slowest_member_kernel_ms": sum(dev_ms) / len(dev_ms),
                  "parity": {"rows_checked": len(ref) * rows, "bit_equal": bool(equal)}}))
L.lib.rtGroupRelease(g)
