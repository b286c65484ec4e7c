"""Would a bank-aware order of the block sort pay?    CPU simulation on the key distribution of a
Cornell frame (oracle/rt_groups.c checker): python tools/sim_bank_sort.py
  a  the kernel's sort: keys descending, arbitrary order inside a key
  b  one sorted list per residue class (slot % 32), chunk c = the c-th entry of every class: conflict-free, but ...
  c  the kernel's sort with the entries of a key laid out round-robin over the residues
For each: chunks a bounce, passes of the quad loop (the union of a chunk's quad lists), wavefronts of one slot-indexed access."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O

W, H, spp = 480, 270, 4
kd = O.cullcheck(O.cornell(W, H), W, H, spp)["keys_dual"][:4].astype(float)
kd /= kd.sum()
faces = np.array([0.17, 0.17, 0.2, 0.12, 0.2, 0.0, 0.1, 0.04]); faces /= faces.sum()    # five walls, -, none, all
boxq = np.array([0, 6, 6, 12])
rng = np.random.default_rng(1)


def quads(keys):
    g = np.bitwise_or.reduce(keys >> 3); f = set(keys & 7)
    return 1 + (5 if 7 in f else len([x for x in f if x < 5])) + boxq[g]


def waves(res):
    return np.bincount(res, minlength=32).max()


def run(n=600, trials=300):
    out = np.zeros((3, 3))
    for _ in range(trials):
        k = (rng.choice(4, size=n, p=kd) << 3) | rng.choice(8, size=n, p=faces)
        res = rng.integers(0, 32, size=n)
        o = np.argsort(-k, kind="stable")
        for c in range(0, n, 32):
            out[0] += (1, quads(k[o[c:c + 32]]), waves(res[o[c:c + 32]]))
        lists = [np.sort(k[res == r])[::-1] for r in range(32)]
        for c in range(max(len(l) for l in lists)):
            out[1] += (1, quads(np.array([l[c] for l in lists if len(l) > c])), 1)
        occ = np.zeros(n, int)
        for key in np.unique(k):
            for r in range(32):
                m = np.nonzero((k == key) & (res == r))[0]; occ[m] = np.arange(len(m))
        o = np.lexsort((res, occ, -k))
        for c in range(0, n, 32):
            out[2] += (1, quads(k[o[c:c + 32]]), waves(res[o[c:c + 32]]))
    return out / trials


r = run()
for name, (ch, q, w) in zip("abc", r):
    print(f"{name}: {ch:5.1f} chunks a bounce, {q:6.1f} quad passes, {w / ch:4.2f} wavefronts per slot-indexed access")
