"""Per-kernel counts of the Blackwell-native instructions in libodex.so (cuobjdump -sass): UTCHMMA / UTCQMMA (tcgen05.mma),
LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA load / store), UTCBAR (tcgen05.commit), SYNCS (mbarrier), plus the
ptxas register / spill lines of the build logs.  No GPU needed.    usage: python tools/sass_summary.py > profiles/r02_sass_summary.md"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ode_rl_b200", "libodex.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCBAR", "SYNCS", "HMMA", "FFMA", "MUFU"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", name).replace("void ", "")
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            for o in OPS:
                if op.startswith(o):
                    counts[cur][o] += 1
    print("# r02 - instruction evidence from `cuobjdump -sass ode_rl_b200/libodex.so` (sm_100a)\n")
    print("Regenerate with `python tools/sass_summary.py` after `make` (the library itself is git-ignored).  UTCHMMA = tcgen05.mma,")
    print("LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG / UTMASTG = TMA tensor load / store, UTCBAR = tcgen05.commit, SYNCS = mbarrier.\n")
    print("| kernel | " + " | ".join(OPS) + " |\n|---|" + "---|" * len(OPS))
    for k, c in counts.items():
        if not any(c[o] for o in OPS[:8]):
            continue
        print("| `%s` | " % k[:70] + " | ".join(str(c[o]) if c[o] else "" for o in OPS) + " |")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print("| **all %d kernels** | " % len(counts) + " | ".join(str(tot[o]) for o in OPS) + " |\n")
    print("## ptxas (registers, spills) of the tensor-core kernels\n\n```")
    for f in sorted(glob.glob(os.path.join(ROOT, "ode_rl_b200", "csrc", "build", "*.ptxas.log"))):
        txt = open(f).read()
        for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*\n\s+(\d+ bytes stack frame, \d+ bytes spill stores, \d+ bytes spill loads)\nptxas info\s+: Used (\d+) registers", txt):
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name).replace("void ", "")
            if any(t in name for t in ("convnet", "conv16", "conv_tc", "wgrad_tc", "wgrad_f16")):
                print("%-60s %3s registers, %s" % (name[:60], m.group(3), m.group(2)))
    print("```")


if __name__ == "__main__":
    main()
