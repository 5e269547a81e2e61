"""Opcode histogram (and the head of the hottest loop) of the product library's kernels, from `cuobjdump -sass`
of cnmf_b200/libcnmf_b200.so -- the Blackwell-native evidence (UBLKCP = cp.async.bulk through the TMA engine,
SYNCS = mbarrier, FFMA2 / DFMA, cluster / DSMEM ops) without needing the binary.

    python tools/sass_hist.py            ->  profiles/sass_hist_<name>.txt for the kernels listed below
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cnmf_b200", "libcnmf_b200.so")
OUT = os.path.join(ROOT, "profiles")

# name -> regex on the mangled function name
KERNELS = {
    "mma_f64_k6": r"mma_kernelILi6ELi6ELi512E",
    "mma_f64_k8": r"mma_kernelILi8ELi6ELi512E",
    "stream_f64_k6": r"stream_kernelIdLi6ELi8ELi2ELi1ELi384ELi0E",
    "stream_f32_k6": r"stream_kernelIfLi6ELi8ELi2ELi2ELi384ELi0E",
    "stream_f32_k8": r"stream_kernelIfLi8ELi8ELi1ELi2ELi384ELi0E",
    "stream_f64_k4": r"stream_kernelIdLi4ELi8ELi2ELi1ELi384ELi0E",
    "resident_f64_k2": r"resident_kernelIdLi2E",
    "resident_f32_k3": r"resident_kernelIfLi3E",
    "batched_f32_k3": r"batched_kernelIfLi3E",
    "exchange_finalize_f64": r"exchange_finalize_kernelIdE",
    "reduce_partials": r"reduce_partials_kernel",
}


def functions(sass):
    cur, name = [], None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                yield name, cur
            name, cur = m.group(1), []
        elif name and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
            cur.append(line)
    if name:
        yield name, cur


def opcode(line):
    body = re.sub(r"/\*[0-9a-f]+\*/", "", line).strip().rstrip(";").strip()
    toks = [t for t in body.split() if not t.startswith("@")]
    return (toks[0] if toks else "?"), body


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fns = list(functions(sass))
    for label, pat in KERNELS.items():
        hits = [(n, b) for n, b in fns if re.search(pat, n)]
        if not hits:
            print("no kernel matches", label, file=sys.stderr)
            continue
        name, body = hits[0]
        ops = collections.Counter()
        full = collections.Counter()
        for line in body:
            op, _ = opcode(line)
            ops[op.split(".")[0]] += 1
            full[op] += 1
        # hottest loop: the backward branch of at most 260 instructions that spans the most FP instructions
        addr = {}
        for i, line in enumerate(body):
            m = re.match(r"\s+/\*([0-9a-f]+)\*/", line)
            addr[int(m.group(1), 16)] = i
        best = (0, 0, 0)
        for i, line in enumerate(body):
            m = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", line)
            if not m:
                continue
            t = addr.get(int(m.group(1), 16))
            if t is None or t >= i:
                continue
            fp = sum(1 for l in body[t:i] if re.search(r"\b(DFMA|DMUL|DADD|FFMA2|FMUL2|FADD2|FFMA)\b", l))
            if fp > best[0] and i - t <= 260:   # innermost arithmetic loops, not the tile loop around them
                best = (fp, t, i)
        with open(os.path.join(OUT, "sass_hist_%s.txt" % label), "w") as f:
            f.write("%s\n%d SASS instructions (cuobjdump -sass cnmf_b200/libcnmf_b200.so)\n\n" % (name, len(body)))
            f.write("opcode histogram (static counts):\n")
            for op, n in ops.most_common(40):
                f.write("  %-14s %6d\n" % (op, n))
            marks = ["UBLKCP", "SYNCS", "FFMA2", "DFMA", "UTMALDG", "UTCHMMA", "LDTM", "STTM", "LDS", "STS", "ATOMS", "REDUX",
                     "LDCU", "CCTL", "MEMBAR", "UCGABAR", "ERRBAR", "ELECT", "SHFL", "MAPA"]
            f.write("\nmarkers: " + ", ".join("%s %d" % (m, sum(n for o, n in full.items() if o.split(".")[0] == m)) for m in marks) + "\n")
            fp, t, i = best
            f.write("\nhottest loop (%d FP instructions between the backward branch at line %d and its target %d), first 40 lines:\n" % (fp, i, t))
            for line in body[t:t + 40]:
                f.write("  " + opcode(line)[1] + "\n")
        print("wrote profiles/sass_hist_%s.txt (%d instructions)" % (label, len(body)))


if __name__ == "__main__":
    main()
