"""Summarise an `ncu --page raw --csv` export (gz) into one line per launch: duration, DRAM traffic, pipe/issue utilisation.

  python tools/ncu_summarise.py gpurun_out/ncu_mixers_raw.csv.gz > profiles/r01_ncu_full_mixers_v5.txt
"""
import csv
import gzip
import re
import sys

COLS = [
    ("gpu__time_duration.sum", "us", 1e-3),
    ("dram__bytes_read.sum", "rdMB", None),
    ("dram__bytes_write.sum", "wrMB", None),
    ("FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 1),
    ("lts__t_sector_hit_rate.pct", "L2hit%", 1),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%", 1),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu%", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1),
    ("smsp__inst_executed.sum", "Minst", 1e-6),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stLong", 1),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stShort", 1),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stBar", 1),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stMath", 1),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stNotSel", 1),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stWait", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
    ("launch__shared_mem_per_block_dynamic", "dsmemKB", None),
]


def to_float(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return float("nan")


def main(path):
    rows = list(csv.reader(l for l in gzip.open(path, "rt") if not l.startswith("==")))
    hdr, units, body = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    print("# " + path + f": {len(body)} launches (ncu --set full --clock-control none; per-launch, cold cache, serialised)")
    print("# " + "  ".join(f"{n:>8s}" for _, n, _ in COLS) + "  kernel")
    for r in body:
        out = []
        for col, name, scale in COLS:
            if col not in idx:
                out.append(f"{'-':>8s}")
                continue
            v, u = to_float(r[idx[col]]), units[idx[col]]
            if scale is None:   # bytes -> MB / KB with the unit ncu printed
                mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Kbyte/block": 1e3, "byte/block": 1}.get(u, 1)
                v = v * mult / (1e3 if name.endswith("KB") else 1e6)
            elif name == "us":
                v = v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
            else:
                v = v * scale
            out.append(f"{v:8.2f}" if name.startswith("st") else f"{v:8.1f}")
        k = re.sub(r"\(.*", "", r[idx["Kernel Name"]])[:70]
        print("  " + "  ".join(out) + "  " + k)


if __name__ == "__main__":
    main(sys.argv[1])
