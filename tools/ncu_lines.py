"""Attribute an ncu report's per-SASS-instruction counters to CUDA source lines.

    python tools/ncu_lines.py <report.ncu-rep> <kernel regex> <cubin> <function substring> <source file> [top]

Uses `ncu --page source --csv` (SASS view, one row per instruction, in order) and `nvdisasm -g` line info of the
same cubin (extract with `cuobjdump -xelf all <lib.so>`).  The library must be the build the report was taken from.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def main():
    rep, kre, cubin, fn, src = sys.argv[1:6]
    top = int(sys.argv[6]) if len(sys.argv) > 6 else 30
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    first = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    block = rows[first[0] + 1: first[1] if len(first) > 1 else len(rows)]
    hdr, data = block[0], block[1:]
    ie, si = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
    start = [i for i, l in enumerate(dis) if l.startswith(".text.") and fn in l][0]
    seq, cur = [], None
    for l in dis[start + 1:]:
        if l.startswith("\t.section") and seq:
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
            seq.append(cur)
    if len(seq) != len(data):
        print("warning: %d SASS instructions in the cubin, %d rows in the report" % (len(seq), len(data)))
    ex, st = collections.Counter(), collections.Counter()
    for ln, r in zip(seq, data):
        ex[ln] += int(r[ie] or 0)
        st[ln] += int(r[si] or 0)
    te, ts = sum(ex.values()), max(sum(st.values()), 1)
    text = open(src).read().split("\n")
    base = src.split("/")[-1]
    print("instructions executed %d, stall samples %d" % (te, ts))
    for ln, c in sorted(ex.items(), key=lambda x: -x[1])[:top]:
        code = text[ln[1] - 1].strip()[:80] if ln and ln[0] == base else str(ln)
        print("%5s %6.2f%% instr %6.2f%% stall  %s" % (ln[1] if ln else "?", 100.0 * c / te, 100.0 * st[ln] / ts, code))


if __name__ == "__main__":
    main()
