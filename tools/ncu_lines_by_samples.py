"""Stall samples per CUDA source line of one captured kernel, most-sampled first (complements tools/ncu_lines.py, which sorts by
executed instructions):  python tools/ncu_lines_by_samples.py rep.ncu-rep [N]"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur, hdr, agg = None, None, {}
for r in csv.reader(io.StringIO(raw)):
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) >= len(hdr) and r[2] == "-":
        try:
            agg[(cur, int(r[0]))] = (r[1].strip()[:90], int(r[7] or 0), int(r[6] or 0))
        except ValueError:
            pass
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
t, ts = sum(v[1] for v in agg.values()), sum(v[2] for v in agg.values())
byfile = {}
for k, v in agg.items():
    byfile.setdefault(k[0], [0, 0]); byfile[k[0]][0] += v[1]; byfile[k[0]][1] += v[2]
for f, (i, s) in sorted(byfile.items(), key=lambda kv: -kv[1][1]):
    print(f"FILE {f:28s} inst {i / t * 100:5.2f}% samples {s / ts * 100:5.2f}%")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:n]:
    print(f"{k[0]:20s}:{k[1]:4d} samples {v[2] / ts * 100:5.2f}% inst {v[1] / t * 100:5.2f}%  {v[0]}")
