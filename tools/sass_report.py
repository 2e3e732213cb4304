"""Per-kernel count of the Blackwell-specific SASS instructions in libddpb200.so (no GPU needed).

Evidence that the hot kernels really use the sm_100a machinery (B200_PROFILING.md names the mnemonics):
  UTCHMMA  tcgen05.mma          LDTM  tcgen05.ld          UTCBAR  tcgen05.commit
  UTMALDG  TMA tensor load      UTMASTG / UTMAREDG  TMA store / reduce-add      UBLKCP  bulk copy
  SYNCS    mbarrier ops         UCGABAR_*  cluster barrier        UTCATOMSWS  TMEM alloc / dealloc
  LDGMC / STG.MC? multimem (NVLS)     ACQBULK / PREEXIT  griddepcontrol.wait / launch_dependents
Usage: python tools/sass_report.py [> profiles/sass_summary_r1.txt]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs744-distributed-data-parallel_b200", "_native", "libddpb200.so")
WATCH = ["UTCHMMA", "LDTM", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "UCGABAR_ARV",
         "UTCATOMSWS", "LDGMC", "MULTIMEM", "ACQBULK", "PREEXIT", "REDG", "ATOMG", "HMMA", "FFMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    demangle = {}
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    names = list(kernels)
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    for n, d in zip(names, dem):
        demangle[n] = d
    print(f"# {os.path.relpath(LIB, ROOT)}: {len(kernels)} kernels; columns = instruction counts in the SASS")
    print("# " + " ".join(f"{w:>9s}" for w in WATCH) + "  kernel")
    for n, c in kernels.items():
        if not any(c[w] for w in WATCH[:13]):
            continue            # plain SIMT kernel
        short = re.sub(r"\(.*", "", demangle.get(n, n)).replace("void ddpb::", "").replace("ddpb::", "")
        print("  " + " ".join(f"{c[w]:9d}" for w in WATCH) + "  " + short)
    tot = collections.Counter()
    for c in kernels.values():
        tot.update(c)
    print("# totals: " + ", ".join(f"{w} {tot[w]}" for w in WATCH if tot[w]))


if __name__ == "__main__":
    sys.exit(main())
