"""Shared SASS reader for tools/sass_hist.py and tools/sass_stalls.py.

load(key, lib) -> list of dicts for the first kernel whose mangled name contains `key`:
  addr, text (predicate + opcode + operands), op (opcode without predicate),
  stall (compiler stall count, control bits 41..44 of the upper word), yld, wr, rd (scoreboards
  set), wait (scoreboard wait mask)."""
import os
import re
import subprocess

DEFAULT_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "pencil-code_b200", "libpcb200.so")


def load(key, lib=None):
    lib = lib or DEFAULT_LIB
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.splitlines()
    cur, name, ins = None, None, []
    i = 0
    while i < len(txt):
        line = txt[i]
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
        elif cur and key in cur and (name is None or name == cur):
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]+) \*/", line)
            if m and i + 1 < len(txt):
                m2 = re.search(r"/\* (0x[0-9a-f]+) \*/", txt[i + 1])
                if m2:
                    name = cur
                    c = int(m2.group(1), 16) >> 41
                    text = m.group(2).strip()
                    op = re.sub(r"^@!?U?P\d+\s+", "", text).split()[0]
                    ins.append(dict(addr=int(m.group(1), 16), text=text, op=op, pred=text.startswith("@"),
                                    stall=c & 15, yld=(c >> 4) & 1, wr=(c >> 5) & 7, rd=(c >> 8) & 7, wait=(c >> 11) & 63))
                    i += 1
        i += 1
    return name, ins


def loops(ins, min_len=300):
    """(lo, hi) address pairs of backward branches whose body has at least min_len instructions"""
    out = set()
    for q in ins:
        if q["op"].startswith("BRA"):
            m = re.search(r"0x([0-9a-f]+)", q["text"])
            if m and int(m.group(1), 16) < q["addr"]:
                lo, hi = int(m.group(1), 16), q["addr"]
                if sum(1 for r in ins if lo <= r["addr"] <= hi) >= min_len:
                    out.add((lo, hi))
    return sorted(out)
