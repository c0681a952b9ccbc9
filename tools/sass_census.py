#!/usr/bin/env python3
"""SASS opcode census of libzkgpu.so's kernels (cuobjdump -sass): per kernel, the number of instructions by pipe class.
Used to derive the per-permutation / per-element instruction budgets quoted in DESIGN.md section 3 and in bench.py's
arithmetic roofs; the committed outputs live under profiles/ (r2_sass_census_*.txt).

  python tools/sass_census.py [substring-of-kernel-name ...]
"""
import collections, pathlib, re, subprocess, sys

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "zkevm_prover_b200" / "libzkgpu.so"

CLASSES = [
    ("IMAD.WIDE", r"^IMAD\.WIDE"), ("IMAD.HI", r"^IMAD\.HI"), ("IMAD (32-bit mul)", r"^IMAD(\.U32|\.SHL)?($|\.X$)?$|^IMAD$"),
    ("IMAD.MOV/IADD/X (flexible)", r"^IMAD\.(MOV|IADD|X)"), ("IADD3", r"^IADD3"), ("LOP3/SHF/SEL/PRMT/SGXT/LEA", r"^(LOP3|SHF|SEL|PRMT|SGXT|LEA|IABS|BMSK|POPC|FLO)"),
    ("ISETP/PLOP3", r"^(ISETP|PLOP3|PSETP)"), ("MOV/CS2R/S2R", r"^(MOV|CS2R|S2R|S2UR)"), ("DFMA/DADD/DMUL", r"^(DFMA|DADD|DMUL)"),
    ("I2F/F2I (XU)", r"^(I2F|F2I|I2FP|F2F)"), ("LDG/STG", r"^(LDG|STG|LD\.|ST\.)"), ("LDS/STS", r"^(LDS|STS)"), ("LDC/ULDC", r"^(LDC|ULDC)"),
    ("LDGSTS/cp.async", r"^(LDGSTS|LDGDEPBAR|DEPBAR)"), ("LDL/STL (spill)", r"^(LDL|STL)"), ("BAR/branch/exit", r"^(BAR|BRA|BSSY|BSYNC|EXIT|RET|CALL|WARPSYNC|NOP|JMP|BRX)"),
    ("uniform (U*)", r"^U[A-Z]"),
]


def main():
    want = sys.argv[1:]
    lib = LIB
    if want and want[0] == "--file":
        lib, want = pathlib.Path(want[1]), want[2:]
    txt = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
    cur, per = None, collections.OrderedDict()
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = name
            per[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            per[cur][m.group(1)] += 1
    for name, cnt in per.items():
        if want and not any(w in name for w in want):
            continue
        total = sum(cnt.values())
        print("== %s : %d SASS instructions" % (name[:150], total))
        left = collections.Counter(cnt)
        for label, rx in CLASSES:
            n = sum(v for k, v in cnt.items() if re.match(rx, k) and k in left)
            for k in list(left):
                if re.match(rx, k):
                    del left[k]
            if n:
                print("   %-34s %6d" % (label, n))
        if left:
            print("   other: " + ", ".join("%s=%d" % kv for kv in left.most_common(12)))


if __name__ == "__main__":
    main()
