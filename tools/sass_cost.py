#!/usr/bin/env python3
"""Static cost of SASS kernels under the issue model measured on B200 (profiles/r2_01_ubench_pipes.txt):
IMAD.WIDE / IMAD.HI hold the sub-partition 4 clocks and nothing pairs with them; every other IMAD form, DFMA/DADD/DMUL and
the ALU-pipe instructions (IADD3, LOP3, SHF, SEL, LEA, ISETP, PRMT, MOV) take 2 clocks each on their own pipe and pair across
pipes; I2F.F64 takes 8 clocks on the conversion pipe.
    T = 4.04 * WIDE + max(2.03 * IMADother + 2.2 * FP64, 2.03 * ALU, 8.06 * I2F)
usage: tools/sass_cost.py file.cubin|file.so [regex-of-kernel-names]"""
import collections
import re
import subprocess
import sys


def cost(c):
    wide = sum(v for k, v in c.items() if k.startswith(("IMAD.WIDE", "IMAD.HI")))
    imado = sum(v for k, v in c.items() if k.startswith("IMAD") and not k.startswith(("IMAD.WIDE", "IMAD.HI")))
    fp64 = sum(v for k, v in c.items() if k.split(".")[0] in ("DFMA", "DADD", "DMUL"))
    alu = sum(v for k, v in c.items() if k.split(".")[0] in ("IADD3", "LOP3", "SHF", "SEL", "LEA", "ISETP", "PRMT", "MOV", "VIADD", "PLOP3", "IABS", "FSEL"))
    i2f = sum(v for k, v in c.items() if k.split(".")[0] in ("I2F", "F2I"))
    t = 4.04 * wide + max(2.03 * imado + 2.2 * fp64, 2.03 * alu, 8.06 * i2f)
    return dict(total=sum(c.values()), WIDE=wide, IMADo=imado, FP64=fp64, ALU=alu, I2F=i2f, T=round(t))


def main():
    path, pat = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else ".")
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    for f in re.split(r"Function : ", txt)[1:]:
        name = f.split("\n")[0]
        if not re.search(pat, name):
            continue
        c = collections.Counter()
        for line in f.split("\n"):
            m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_.]+)", line)
            if m:
                c[m.group(1)] += 1
        print(name[:60], cost(c))
        if "-v" in sys.argv:
            print("    ", c.most_common(16))


if __name__ == "__main__":
    main()
