"""Turn an `ncu --set full` capture of the bucket-accumulation launches of one 64-ballot chunk into the evidence bench.py
quotes (profiles/ncu_accumulate_evidence.json) and a readable summary (profiles/<name>.txt).

  ncu -i gpurun_out/r02_acc.ncu-rep --page raw --csv > /tmp/acc_raw.csv
  python tools/ncu_evidence.py /tmp/acc_raw.csv profiles/r02_ncu_full_msm_accumulate.txt

The dominant launch (the H_query multiexp: the longest G1 one) supplies traffic and pipe utilisation."""
import csv
import json
import os
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]


def to_bytes(v: str, unit: str) -> float:
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main(raw_csv: str, out_txt: str, what: str = ""):
    rows = list(csv.reader(open(raw_csv)))
    hdr, units, launches = rows[0], rows[1], rows[2:]
    u = dict(zip(hdr, units))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(out_txt, "w") as f:
        f.write("# ncu --set full --clock-control none; %s\n" % (what or raw_csv))
        for i, r in enumerate(launches):
            d = dict(zip(hdr, r))
            name = d["Kernel Name"].replace("void unnamed>::", "").split("(")[0]
            f.write("\n== launch %d: %s\n" % (i, name))
            for k in KEYS:
                if k in d and d[k] != "":
                    f.write("  %-86s %s %s\n" % (k, d[k], u.get(k, "")))
    g1 = [dict(zip(hdr, r)) for r in launches if "k_msm_accumulate<FqOps>" in dict(zip(hdr, r))["Kernel Name"].replace("unnamed>::", "")]
    if g1 and "accumulate" in out_txt:
        top = max(g1, key=lambda d: float(d["gpu__time_duration.sum"].replace(",", "")))
        try:
            commit = subprocess.run(["git", "-C", root, "log", "-1", "--format=%H", "--", "vote-saver-protocol_b200/csrc/msm_impl.cuh",
                                     "vote-saver-protocol_b200/csrc/fp.cuh", "vote-saver-protocol_b200/csrc/ec.cuh"],
                                    capture_output=True, text=True).stdout.strip()
        except Exception:
            commit = ""
        ev = {"kernel": "k_msm_accumulate<FqOps>", "launch": "H_query multiexp of one 256-ballot chunk (the longest G1 accumulation launch)",
              "duration_ms": float(top["gpu__time_duration.sum"].replace(",", "")),
              "dram_bytes_per_launch": to_bytes(top["dram__bytes_read.sum"], u["dram__bytes_read.sum"])
              + to_bytes(top["dram__bytes_write.sum"], u["dram__bytes_write.sum"]),
              "pipe_fmaheavy_pct": float(top["sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"]),
              "registers_per_thread": int(top["launch__registers_per_thread"]),
              "source": os.path.relpath(out_txt, root), "kernel_sources_commit": commit,
              "note": "dram__bytes_read.sum + dram__bytes_write.sum of that launch from the ncu --set full capture in " + os.path.relpath(out_txt, root)}
        with open(os.path.join(root, "profiles", "ncu_accumulate_evidence.json"), "w") as f:
            json.dump(ev, f, indent=1)
        print(json.dumps(ev))


if __name__ == "__main__":
    main(*sys.argv[1:])
