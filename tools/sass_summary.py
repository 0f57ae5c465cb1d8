#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of heatb200/libheat.so (cuobjdump -sass) -> profiles/rNN_sass_summary.txt.

    python tools/sass_summary.py [out.txt]

The columns are the instructions that prove what the kernels are made of: UTMALDG (TMA tensor loads), SYNCS
(mbarrier), DFMA/DADD/DMUL and FFMA/FADD/FMUL (contracted vs individually rounded arithmetic), 128-bit global
stores, shared-memory loads, shuffles, ACQBULK/PDL markers.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cuda-hip-mpi-heat-equation-test_b200", "heatb200", "libheat.so")
COLS = ["UTMALDG", "SYNCS", "DFMA", "DADD", "DMUL", "FFMA", "FADD", "FMUL", "STG.E.128", "STG.E.64", "STG.E",
        "LDG", "LDS", "SHFL", "ATOM", "RED", "BAR", "ACQBULK", "total"]


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_summary.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    fn, counts = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and fn:
            op = m.group(1)
            c = counts[fn]
            c["total"] += 1
            for col in COLS[:-1]:
                if op == col or op.startswith(col + "."):
                    c[col] += 1
    demangled = {}
    try:
        names = subprocess.run(["cu++filt"] + list(counts), capture_output=True, text=True, check=True).stdout.splitlines()
        demangled = dict(zip(counts, names))
    except Exception:
        pass
    with open(out, "w") as f:
        f.write("# cuobjdump -sass %s   (arch: %s)\n" % (os.path.relpath(LIB, ROOT), ", ".join(arch)))
        f.write("# regenerate: python tools/sass_summary.py\n")
        f.write("# STG.E counts every global store; STG.E.128 / STG.E.64 are the subsets of that width\n")
        f.write("\t".join(["kernel"] + COLS) + "\n")
        for k, c in counts.items():
            name = demangled.get(k, k)
            i = name.find(">(")                      # drop the parameter list, keep the template arguments
            name = (name[:i + 1] if i >= 0 else re.sub(r"\(.*", "", name)).replace("void ", "").replace("heat::", "")
            name = name.replace("(int)", "").replace("(bool)", "").replace("(anonymous namespace)::", "")
            f.write("\t".join([name] + [str(c[col]) for col in COLS]) + "\n")
    print("wrote", out, "-", len(counts), "kernels; arch", arch)


if __name__ == "__main__":
    main()
