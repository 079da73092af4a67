"""Top stall-sample lines of every kernel section of an ncu report's source page.
usage: python scripts/ncu_hot.py report.ncu-rep [N] [section-index]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
only = int(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
for si, (a, b) in enumerate(zip(starts[:-1], starts[1:])):
    if only is not None and si != only:
        continue
    hdr = rows[a + 1]; body = [r for r in rows[a + 2:b] if len(r) == len(hdr)]
    ci = {h: i for i, h in enumerate(hdr)}
    tot = sum(int(r[ci["# Samples"]] or 0) for r in body)
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    print(f"=== section {si}: {rows[a][1][:60]}  lines {len(body)}  samples {tot}")
    agg = {h: sum(int(r[ci[h]] or 0) for r in body) for h in stalls}
    print("   by reason:", sorted(((v, k[6:]) for k, v in agg.items() if v), reverse=True)[:6])
    idx = sorted(range(len(body)), key=lambda k: -int(body[k][ci["# Samples"]] or 0))[:N]
    for k in sorted(idx):
        r = body[k]; s = int(r[ci["# Samples"]] or 0)
        top = sorted(((int(r[ci[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
        print(f"{k:5d} {100*s/max(tot,1):5.1f}% {r[ci['Source']].strip()[:90]:90s} {top}")
