"""Per-kernel counts of the SASS mnemonics that prove what the library runs on (B200_PROFILING.md): tcgen05 MMAs
(UTCHMMA / UTCQMMA ...), TMEM loads (LDTM), TMA loads (UTMALDG), bulk copies (UBLKCP), L2 reductions (REDG), packed
fp32 FMAs (FFMA2), cluster barriers (UCGABAR), mbarrier ops (SYNCS).  Runs on the CPU container (cuobjdump only).

    python tools/sass_summary.py [ovmono3d_b200/libovmono3d_lift.so] > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATTERNS = ["UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "REDG", "ATOMG", "RED.",
            "FFMA2", "FFMA", "HMMA", "UCGABAR", "SYNCS", "LDS", "STS", "LDG", "STG", "SHFL", "MUFU", "ACQBULK", "ST.E", "STAS"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ovmono3d_b200", "libovmono3d_lift.so")
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", out)))
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        kernels[cur]["_total"] += 1
        for p in PATTERNS:
            if p.endswith("."):
                if op.startswith(p) or op == p[:-1]:
                    kernels[cur][p] += 1
            elif op == p or op.startswith(p + "."):
                if p == "FFMA" and op.startswith("FFMA2"):
                    continue
                if p == "UTCHMMA" and ".2CTA" in op:
                    kernels[cur]["UTCHMMA.2CTA"] += 1
                kernels[cur][p] += 1
    print("library: %s" % os.path.relpath(lib, ROOT))
    print("cubin architectures: %s" % ", ".join(arch))
    demangle = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    cols = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UTMALDG", "UBLKCP", "UTCBAR", "UCGABAR", "SYNCS", "REDG", "ATOMG", "FFMA2", "FFMA",
            "MUFU", "SHFL", "_total"]
    print("%-72s " % "kernel" + " ".join("%8s" % c.replace("UTCHMMA.2CTA", ".2CTA").replace("_total", "instrs") for c in cols))
    tot = collections.Counter()
    for (k, c), name in zip(kernels.items(), demangle):
        name = re.sub(r"ovm::\(anonymous namespace\)::", "", name)
        name = re.sub(r"\((int|bool|unsigned int)\)", "", name)
        name = re.sub(r"\(.*", "", name).replace("void ", "").replace("ovm::<unnamed>::", "")
        print("%-72s " % name[:72] + " ".join("%8d" % c[x] for x in cols))
        tot.update(c)
    print("%-72s " % "TOTAL" + " ".join("%8d" % tot[x] for x in cols))


if __name__ == "__main__":
    main()
