#!/usr/bin/env python
"""Extract the metrics we cite from an .ncu-rep into a small text summary (profiles/*.txt).
usage: tools/ncu_summary.py report.ncu-rep [kernel-substring] > profiles/xxx.txt
       tools/ncu_summary.py report.ncu-rep direct_force_kernel --traffic-json BODIES profiles/direct_traffic.json
         (DRAM bytes of one force-kernel launch, read by bench.py for roofline.traffic)"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg", "smsp__inst_executed.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def main():
    rep = sys.argv[1]
    sub = sys.argv[2] if len(sys.argv) > 2 else ""
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kcol = hdr.index("Kernel Name")
    if "--traffic-json" in sys.argv:
        import json
        k = sys.argv.index("--traffic-json")
        bodies, dst = int(sys.argv[k + 1]), sys.argv[k + 2]
        for r in data:
            if sub in r[kcol]:
                def val(name):
                    i = hdr.index(name)
                    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
                    return float(r[i].replace(",", "")) * mult
                rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
                with open(dst, "w") as f:
                    json.dump({"bodies": bodies, "dram_bytes_per_launch": int(rd + wr), "dram_bytes_read": int(rd),
                               "dram_bytes_written": int(wr), "kernel": r[kcol][:80],
                               "source": f"ncu --set full capture {rep.split('/')[-1]} (tools/ncu_summary.py --traffic-json)"}, f)
                    f.write("\n")
                print(open(dst).read())
                return
        raise SystemExit("kernel not found")
    print(f"# {rep}")
    for r in data:
        if sub and sub not in r[kcol]:
            continue
        print(f"\n## {r[kcol][:150]}")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"{w:90s} {r[i]:>18s} {units[i]}")


if __name__ == "__main__":
    main()
