"""Decode the scheduling control fields of a kernel's SASS (sm_100a 128-bit encoding): stall count, the scoreboard an
instruction sets for its result (wr) / for its source operands (rd), and the scoreboards it waits on.  Tuning aid: shows
which instruction carries the wait for a batch of loads (e.g. a temporary that reuses the register of a dead load result).
python -m harness.sass_ctl build/p2g.o _Z5k_p2gILb1EEv7P2GArgs"""
import re
import subprocess
import sys


def decode(obj, fun):
    txt = subprocess.check_output(["cuobjdump", "-sass", "-fun", fun, obj], text=True).split("\n")
    out, i = [], 0
    while i < len(txt):
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", txt[i])
        if m and i + 1 < len(txt):
            m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", txt[i + 1])
            if m2:
                hi = int(m2.group(1), 16)
                out.append((m.group(1), m.group(2).strip(), (hi >> 41) & 0xF, (hi >> 46) & 7, (hi >> 49) & 7, (hi >> 52) & 0x3F))
                i += 2
                continue
        i += 1
    return out


if __name__ == "__main__":
    for a, t, stall, wr, rd, wait in decode(sys.argv[1], sys.argv[2]):
        print(f"{a} st={stall:2d} wr={'-' if wr == 7 else wr} rd={'-' if rd == 7 else rd} wait={wait:06b} {t}")
