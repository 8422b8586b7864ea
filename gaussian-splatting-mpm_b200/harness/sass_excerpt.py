"""Per-kernel SASS evidence from the built library (cuobjdump, no GPU needed): resource usage, opcode histogram and
the memory / reduction / warp-collective instructions with their modifiers -- the vector RED of p2g's tile flush, the
LDGSTS (cp.async) staging of g2p, and the absence of local-memory (spill) traffic.
    python -m harness.sass_excerpt > ../profiles/rNN_sass_excerpts.txt"""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libmpmk.so")
KERNELS = ["k_p2g<true>", "k_p2g<false>", "k_g2p<true>", "k_g2p<false>", "k_grid(", "k_cs_flag_scan", "k_sort_onesweep<8, 256, 8>",
           "k_cs_place_movers", "k_cs_place_stayers", "k_tile_runs", "k_build_items<1>", "k_halo_push", "k_halo_wait_add"]
MEM = re.compile(r"^(RED|REDG|ATOM|ATOMG|ATOMS|LDGSTS|LDL|STL|MATCH|SHFL|LDS|STS|LD|ST|LDG|STG|VOTE|REDUX|UBLKCP|UTMALDG)(\.|$)")


def main():
    res = subprocess.run(["cuobjdump", "--dump-resource-usage", LIB], capture_output=True, text=True).stdout.splitlines()
    usage = {}
    for i, line in enumerate(res):
        m = re.search(r"Function (\S+):", line)
        if m and i + 1 < len(res):
            usage[m.group(1)] = res[i + 1].strip()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)[1:]
    print(f"# SASS excerpts of {os.path.relpath(LIB)} (cuobjdump -sass, sm_100a)")
    for blk in blocks:
        mangled = blk.split("\n", 1)[0].strip()
        name = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
        if not any(k in name for k in KERNELS):
            continue
        ops, mem = collections.Counter(), collections.Counter()
        for line in blk.splitlines():
            m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
            if not m:
                continue
            op = m.group(1)
            ops[op.split(".")[0]] += 1
            if MEM.match(op):
                mem[op] += 1
        print(f"\n== {name}\n   {usage.get(mangled, '')}")
        print("   opcodes: " + "; ".join(f"{n} {o}" for o, n in ops.most_common(20)))
        print("   memory / reduction / collective: " + "; ".join(f"{n} {o}" for o, n in sorted(mem.items(), key=lambda kv: -kv[1])))
        print(f"   local-memory (spill) instructions: {sum(n for o, n in mem.items() if o.startswith(('LDL', 'STL')))}")


if __name__ == "__main__":
    main()
