#!/usr/bin/env python
"""Static issue-time model of a kernel's loops from the SASS control words.

usage: sass_stalls.py <mangled-name-substring> [lib-or-cubin]
For every loop body (backward BRA) with >= 300 instructions: instruction count and the sum of the
compiler's stall counts = the cycles ONE warp needs to issue the body when no scoreboard wait occurs.
Unrolled mbarrier try_wait spin blocks (SYNCS.PHASECHK + BRA pairs after the first) are skipped: they
do not execute when the barrier is ready.  rhs_dly is issue bound, so this number tracks its run time
closely enough to rank code variants without GPU time."""
import collections
import re
import sys

from sassparse import load, loops


def main():
    key = sys.argv[1]
    name, ins = load(key, sys.argv[2] if len(sys.argv) > 2 else None)
    if not ins:
        sys.exit("kernel not found: " + key)
    print("kernel:", name, " instructions:", len(ins))
    for lo, hi in loops(ins):
        body = [q for q in ins if lo <= q["addr"] <= hi]
        keep, run = [], 0
        for q in body:
            if q["op"].startswith("SYNCS.PHASECHK"):
                run += 1
                if run > 1:
                    continue
            elif q["op"] == "BRA" and run > 1:
                continue
            elif not q["op"].startswith("BRA"):
                run = 0
            keep.append(q)
        cyc = sum(max(q["stall"], 1) for q in keep)
        per, cnt = collections.Counter(), collections.Counter()
        for q in keep:
            k = "fp64" if re.match(r"^D(FMA|ADD|MUL|SETP|MNMX)", q["op"]) else q["op"].split(".")[0]
            per[k] += max(q["stall"], 1)
            cnt[k] += 1
        waits = sum(1 for q in keep if q["wait"])
        print(f"loop 0x{lo:x}..0x{hi:x}: {len(body)} instr, {len(keep)} on the no-spin path, "
              f"static stall sum {cyc} cycles, {waits} instr wait on a scoreboard")
        print("   " + ", ".join(f"{k}: n={cnt[k]} cyc={v}" for k, v in per.most_common(14)))


if __name__ == "__main__":
    main()
