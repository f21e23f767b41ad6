#!/usr/bin/env python
"""SASS opcode histogram of one kernel, whole and per loop body.

usage: sass_hist.py <mangled-name-substring> [lib-or-cubin]
Committed output under profiles/ is the evidence that rhs_dly is Blackwell-native:
LDTM/STTM = tcgen05.ld/st (tensor memory), UTMALDG = cp.async.bulk.tensor (TMA),
SYNCS = mbarrier, USETMAXREG = setmaxnreg (warp specialisation)."""
import collections
import re
import sys

from sassparse import load, loops

CLASSES = [
    ("fp64", r"^D(FMA|ADD|MUL|SETP|MNMX)"),
    ("mufu/cvt", r"^(MUFU|F2F|F2I|I2F|FRND)"),
    ("mov", r"^(MOV|UMOV|IMAD\.MOV|PRMT|SEL|FSEL|SHFL|CS2R|S2R|S2UR|R2UR|MOVM)"),
    ("lds", r"^LDS"), ("sts", r"^STS"),
    ("tmem", r"^(LDTM|STTM|UTC)"),
    ("tma/mbarrier", r"^(UTMA|SYNCS|BAR|FENCE|MEMBAR|ERRBAR|WARPSYNC|DEPBAR|NANOSLEEP|ELECT|VOTE|UBLKCP|USETMAXREG)"),
    ("ldg/stg/ldc", r"^(LDG|STG|LD$|ST$|LDL|STL|LDC|ULDC|LDCU|ATOM|RED|CCTL)"),
    ("int", r"^(IADD|IMAD|LEA|LOP|SHF|ISETP|UIADD|UIMAD|ULEA|ULOP|USHF|UISETP|VIADD|UP2UR|PLOP|UPLOP|IABS|IMNMX|VIMNMX|POPC|FLO|BREV|USEL|UPRMT|P2R|R2P)"),
    ("fp32", r"^(FFMA|FADD|FMUL|FSETP|FMNMX|FCHK|HADD|HFMA|HMUL)"),
    ("branch", r"^(BRA|BSSY|BSYNC|EXIT|RET|CALL|BRX|JMP|NOP|BPT|YIELD|BREAK|KILL|BMOV)"),
]


def classify(op):
    for name, pat in CLASSES:
        if re.match(pat, op):
            return name
    return "other:" + op.split(".")[0]


def show(title, body):
    h = collections.Counter(classify(q["op"]) for q in body)
    print(f"{title}: {len(body)} instructions")
    for k, v in sorted(h.items(), key=lambda kv: -kv[1]):
        print(f"   {k:14s} {v:6d}  {100.0 * v / len(body):5.1f} %")
    ops = collections.Counter(q["op"] for q in body)
    print("   top opcodes:", ", ".join(f"{o}={n}" for o, n in ops.most_common(28)))


def main():
    key = sys.argv[1]
    name, ins = load(key, sys.argv[2] if len(sys.argv) > 2 else None)
    if not ins:
        sys.exit("kernel not found: " + key)
    print("kernel:", name)
    show("whole kernel", ins)
    native = collections.Counter(q["op"].split(".")[0] for q in ins
                                 if re.match(r"^(LDTM|STTM|UTMALDG|UTMASTG|SYNCS|USETMAXREG|UTCBAR|UTCATOMSWS)", q["op"]))
    print("Blackwell-native opcodes:", dict(native))
    for lo, hi in loops(ins):
        show(f"loop 0x{lo:x}..0x{hi:x}", [q for q in ins if lo <= q["addr"] <= hi])


if __name__ == "__main__":
    main()
