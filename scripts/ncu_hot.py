"""Print the top stall-sample SASS lines of an ncu report's source page.
usage: python scripts/ncu_hot.py report.ncu-rep [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ci["# Samples"]] or 0) for r in body)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot)
idx = sorted(range(len(body)), key=lambda k: -int(body[k][ci["# Samples"]] or 0))[:N]
for k in sorted(idx):
    r = body[k]
    s = int(r[ci["# Samples"]])
    top = sorted(((int(r[ci[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print(f"{k:5d} {100*s/tot:5.1f}% {r[ci['Source']].strip()[:70]:70s} {top}")
if len(sys.argv) > 3:
    # cumulative share by ranges given as comma separated boundaries, and a SASS window
    b = [int(x) for x in sys.argv[3].split(",")]
    b = [0] + b + [len(body)]
    for lo, hi_ in zip(b[:-1], b[1:]):
        s = sum(int(body[k][ci["# Samples"]] or 0) for k in range(lo, hi_))
        print(f"lines [{lo},{hi_}) {100*s/tot:5.1f}%")
if len(sys.argv) > 4:
    lo, hi_ = [int(x) for x in sys.argv[4].split(":")]
    for k in range(lo, hi_):
        print(k, body[k][ci["# Samples"]], body[k][ci["Source"]].strip()[:100])
