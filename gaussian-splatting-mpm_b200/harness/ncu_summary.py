"""Column subset of ncu --set full reports, one row per captured launch:
    python -m harness.ncu_summary out.csv a.ncu-rep [b.ncu-rep ...]"""
import csv
import subprocess
import sys

COLS = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        # atomic / reduction traffic (north_star: "atomic throughput"): L2 sectors of RED and ATOM operations, and the
        # L1 requests that carried global reductions (k_p2g's tile flush is red.global.add.v4.f32)
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "lts__t_sectors_srcunit_tex_op_red.sum",
        "smsp__inst_executed_op_global_red.sum", "smsp__sass_inst_executed_op_global_red.sum",
        "sm__cycles_elapsed.max"]


def main():
    out, reps = sys.argv[1], sys.argv[2:]
    rows, units = [], None
    for rep in reps:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], capture_output=True, text=True).stdout
        r = list(csv.reader(txt.splitlines()))
        H = r[0]
        idx = [H.index(c) if c in H else None for c in COLS]
        if units is None:
            units = [r[1][i] if i is not None else "" for i in idx]
        for row in r[2:]:
            rows.append([row[i] if i is not None else "" for i in idx])
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(COLS)
        w.writerow(units)
        w.writerows(rows)
    print(f"{len(rows)} launches -> {out}")


if __name__ == "__main__":
    main()
