"""One line per profiled launch of an ncu report (--set full): duration, DRAM read / write bytes, tensor-pipe and
SM-busy percentages, L2 -> SM traffic, registers, grid.
usage: python scripts/ncu_summary.py report.ncu-rep [label ...]     (labels name the launches in order)"""
import csv, io, subprocess, sys

rep = sys.argv[1]
labels = sys.argv[2:]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, body = rows[0], rows[1], rows[2:]
ci = {h: i for i, h in enumerate(hdr)}


def col(r, name, scale=1.0):
    i = ci.get(name)
    if i is None or r[i] in ("", "n/a"):
        return None
    v = float(r[i].replace(",", ""))
    u = units[i]
    mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
            "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}.get(u, 1.0)
    return v * mult * scale


def fmt(v, f):
    return "n/a" if v is None else f % v


for k, r in enumerate(body):
    name = r[ci["Kernel Name"]].split("(")[0][-48:]
    t = col(r, "gpu__time_duration.sum")
    rd, wr = col(r, "dram__bytes_read.sum"), col(r, "dram__bytes_write.sum")
    tens = col(r, "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active")
    if tens is None:
        tens = col(r, "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active")
    tpc = None
    for h in hdr:
        if "pipe_tensor" in h and "pct_of_peak_sustained_elapsed" in h and ci[h] < len(r) and r[ci[h]] not in ("", "n/a"):
            tpc = (h, float(r[ci[h]].replace(",", "")))
            break
    l2 = col(r, "lts__t_bytes.sum")
    smb = col(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed")
    label = labels[k] if k < len(labels) else name
    print(f"{label:52s} time {fmt(t and t * 1e3, '%7.3f')} ms  dram read {fmt(rd and rd / 1e9, '%6.3f')} GB write "
          f"{fmt(wr and wr / 1e9, '%6.3f')} GB  L2 bytes {fmt(l2 and l2 / 1e9, '%7.3f')} GB  sm throughput {fmt(smb, '%5.1f')} %  "
          f"tensor {fmt(tpc[1] if tpc else tens, '%5.1f')} %  grid {r[ci['Grid Size']]}  block {r[ci['Block Size']]}  "
          f"regs {r[ci['Registers Per Thread']] if 'Registers Per Thread' in ci else r[ci.get('launch__registers_per_thread', 0)]}")
