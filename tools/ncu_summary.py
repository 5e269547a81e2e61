"""Condense an .ncu-rep (read here, no GPU) into the numbers DESIGN.md / profiles/ quote:
key raw metrics plus a per-role breakdown of the stream kernel from the source page."""
import csv
import io
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic"]
print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
for h, u, v in zip(hdr, units, vals):
    if h in want or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")
                     and float(v or 0) > 0.3):
        print("  %-85s %-8s %s" % (h, u, v))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def f(r, c):
    try:
        return float(r[ix[c]])
    except Exception:
        return 0.0


# roles are separated by EXIT instructions in program order: prologue | producer | role W | role H | tail
cuts = [i for i, r in enumerate(data) if r[ix["Source"]].strip().endswith("EXIT") and "@" not in r[ix["Source"]]]
tot = sum(f(r, "# Samples") for r in data)
print("samples", int(tot), "unconditional EXIT at", cuts)
start = 0
for ci, c in enumerate(cuts + [len(data)]):
    seg = data[start:c + 1]
    start = c + 1
    if not seg:
        continue
    ie = sum(f(r, "Instructions Executed") for r in seg)
    sm = sum(f(r, "# Samples") for r in seg)
    fp = sum(f(r, "Instructions Executed") for r in seg
             if any(t in r[ix["Source"]] for t in ("FFMA2", "FMUL2", "FADD2", "DFMA", "DMUL", "DADD")))
    wf = sum(f(r, "L1 Wavefronts Shared") for r in seg)
    waits = sum(f(r, "# Samples") for i, r in enumerate(seg)
                if "BRA" in r[ix["Source"]] and i > 0 and "TRYWAIT" in seg[i - 1][ix["Source"]])
    ops = Counter()
    for r in seg:
        t = [o for o in r[ix["Source"]].split() if not o.startswith("@")]
        if t:
            ops[t[0].split(".")[0]] += f(r, "Instructions Executed")
    top = ", ".join("%s %.0f%%" % (k, 100 * v / max(ie, 1)) for k, v in ops.most_common(8))
    print("segment %d: inst %.1fM (packed/fp64 math %.1fM), samples %.1f%% (mbarrier-wait %.1f%%), smem wavefronts %.1fM | %s"
          % (ci, ie / 1e6, fp / 1e6, 100 * sm / tot, 100 * waits / tot, wf / 1e6, top))
