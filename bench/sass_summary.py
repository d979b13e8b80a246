#!/usr/bin/env python
"""Regenerates docs/sass/summary.txt: per kernel of libdtf_b200_cuda.so, the count of the SASS
mnemonics that prove which hardware path it uses (tcgen05 -> UTCHMMA / LDTM / UTCBAR, TMA ->
UTMALDG, mbarrier -> SYNCS, peer atomics -> REDG / ATOMG ... .SYS). Runs on the CPU box."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedtensorflow_b200", "lib", "libdtf_b200_cuda.so")
KEY = re.compile(r"^(UTC|UTMA|LDTM|STTM|SYNCS|REDG|ATOMG|ATOMS|MEMBAR|UCGABAR|ACQBULK|ERRBAR|CCTL|HMMA|REDUX|ELECT|UBLKCP)")
GLOBAL = re.compile(r"^(LDG|STG)")


def main() -> None:
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, name = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            kernels[name] = [0, collections.Counter(), collections.Counter()]
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]*)", line)
        if m and name:
            op = m.group(1)
            kernels[name][0] += 1
            if KEY.match(op):
                kernels[name][1][op] += 1
            elif GLOBAL.match(op):
                kernels[name][2][op] += 1
    lines = ["# SASS evidence per kernel (cuobjdump -sass libdtf_b200_cuda.so, sm_100a); regenerate with bench/sass_summary.py",
             "# tcgen05.mma -> UTCHMMA (.2CTA = cta_group::2), tcgen05.ld -> LDTM, TMA -> UTMALDG, tcgen05.commit -> UTCBAR,",
             "# mbarrier -> SYNCS.*, cluster barrier -> UCGABAR; legacy mma.sync would show as HMMA (none expected)", ""]
    hmma = 0
    for k, (n, key, glob) in kernels.items():
        hmma += sum(v for o, v in key.items() if o.startswith("HMMA"))
        lines.append(f"{k}  [{n} instr]")
        lines.append("    " + ", ".join(f"{o} x{c}" for o, c in key.most_common()))
        lines.append("    global: " + ", ".join(f"{o} x{c}" for o, c in glob.most_common()))
    lines.append("")
    lines.append(f"# kernels: {len(kernels)}; HMMA (mma.sync) instructions: {hmma}")
    path = os.path.join(ROOT, "docs", "sass", "summary.txt")
    open(path, "w").write("\n".join(lines) + "\n")
    print(path, len(kernels), "kernels, HMMA:", hmma)


if __name__ == "__main__":
    main()
