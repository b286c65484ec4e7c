"""Executed instructions and stall samples per CUDA source line of one captured kernel (needs -lineinfo and --import-source on):
    python tools/ncu_lines.py rep.ncu-rep [N]            top N lines
    python tools/ncu_lines.py a.ncu-rep b.ncu-rep        lines whose executed-instruction counts differ most between two captures"""
import csv, io, subprocess, sys


def lines(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(io.StringIO(raw)):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) > 5 and r[0] == "Line No":
            hdr = r
        elif hdr and len(r) >= len(hdr) and r[2] == "-":
            try:
                agg[(cur, int(r[0]))] = (r[1].strip()[:80], int(r[7] or 0), int(r[6] or 0))
            except ValueError:
                pass
    return agg


a = lines(sys.argv[1])
if len(sys.argv) > 2 and sys.argv[2].endswith(".ncu-rep"):
    b = lines(sys.argv[2])
    ta, tb = sum(v[1] for v in a.values()), sum(v[1] for v in b.values())
    print(f"total executed: {ta} vs {tb} ({(ta / tb - 1) * 100:+.2f} %)")
    keys = sorted(set(a) | set(b), key=lambda k: -abs(a.get(k, ("", 0, 0))[1] - b.get(k, ("", 0, 0))[1]))
    for k in keys[:25]:
        x, y = a.get(k, ("", 0, 0)), b.get(k, ("", 0, 0))
        print(f"{k[0]:20s}:{k[1]:4d} {x[1] / ta * 100:6.2f}% vs {y[1] / tb * 100:6.2f}%  diff {(x[1] - y[1]) / tb * 100:+6.2f}% of b   {x[0] or y[0]}")
else:
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    t, ts = sum(v[1] for v in a.values()), sum(v[2] for v in a.values())
    for k, v in sorted(a.items(), key=lambda kv: -kv[1][1])[:n]:
        print(f"{k[0]:20s}:{k[1]:4d} inst {v[1] / t * 100:5.2f}% samples {v[2] / ts * 100:5.2f}%  {v[0]}")
