#!/usr/bin/env python
"""SASS opcode histogram per kernel of the built library: the evidence that the kernels are Blackwell-native
(UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA tensor load / store, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk,
FFMA2 = packed fp32 FMA).      tools/sass_histogram.py [library.so] > profiles/r02_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
KEY = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UBLKCP", "UTCBAR", "SYNCS", "FFMA2", "FFMA", "FMUL", "FADD",
       "MUFU", "LDS", "STS", "LDG", "STG", "ATOM", "RED", "BAR", "SHFL", "VOTE", "S2R", "BRA", "HMMA", "IMMA"]


def main():
    lib = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "simpleraytracer_b200" / "lib" / "libModelRunner.so"
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    print(f"# SASS opcode histogram of {lib.name} (cuobjdump -sass, sm_100a); columns: selected opcodes, then the total")
    print("# UTCHMMA = tcgen05.mma, UTMALDG/UTMASTG = TMA tensor load/store, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk, FFMA2 = fma.rn.f32x2")
    for (name, hist), nice in zip(kernels.items(), demangle):
        nice = re.sub(r"\(anonymous namespace\)::", "", nice)
        nice = re.sub(r"\(ML::RenderLaunch\)|\(ML::GridBuild\)", "", nice)[:110]
        sel = " ".join(f"{k}={hist[k]}" for k in KEY if hist[k])
        print(f"{nice}\n    total={sum(hist.values())} {sel}")


if __name__ == "__main__":
    main()
