"""Top stall lines of one kernel from an ncu report: python -m harness.ncu_hot file.ncu-rep [N]"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    H = rows[hdr]
    ci, ce, cs = H.index("Warp Stall Sampling (All Samples)"), H.index("Instructions Executed"), H.index("Source")
    body = [r for r in rows[hdr + 1:] if len(r) > ci and r[ci].isdigit()]
    tot = sum(int(r[ci]) for r in body) or 1
    print("total samples", tot, " instructions", sum(int(r[ce]) for r in body if r[ce].isdigit()))
    idx = {id(r): i for i, r in enumerate(body)}
    for r in sorted(body, key=lambda r: -int(r[ci]))[:top]:
        print(f"{100 * int(r[ci]) / tot:5.1f}%  #{idx[id(r)]:4d}  exec={r[ce]:>8}  {r[cs].strip()[:100]}")


if __name__ == "__main__":
    main()
