"""Digest of an `ncu -i X.ncu-rep --page raw --csv` file: one block per kernel name with the numbers
profiles/rNN_summary.md quotes (medians over the captured launches of that kernel).

  python tools/ncu_digest.py raw.csv [name-substring]
"""
import collections
import csv
import statistics
import sys

KEYS = [
    ("duration", "gpu__time_duration.sum"),
    ("dram read", "dram__bytes_read.sum"),
    ("dram write", "dram__bytes_write.sum"),
    ("dram throughput % of peak", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("fp64 pipe active %", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
    ("fp32/alu pipe fma %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("alu pipe %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("issue active %", "sm__issue_active.avg.pct_of_peak_sustained_elapsed"),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("registers/thread", "launch__registers_per_thread"),
    ("dynamic smem/block", "launch__shared_mem_per_block_dynamic"),
    ("occupancy limit blocks (regs)", "launch__occupancy_limit_registers"),
    ("occupancy limit blocks (smem)", "launch__occupancy_limit_shared_mem"),
    ("L2 hit %", "lts__t_sector_hit_rate.pct"),
    ("local load sectors", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"),
    ("local store sectors", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum"),
]
STALL = "smsp__average_warps_issue_stalled_"


def num(s):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    while rows and (not rows[0] or rows[0][0] != "ID"):
        rows.pop(0)
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    col = {h: i for i, h in enumerate(hdr)}
    by = collections.defaultdict(list)
    for r in data:
        name = r[col["Kernel Name"]]
        if want in name:
            by[name.split("(")[0]].append(r)
    for name, rs in by.items():
        print("## %s  (%d launches, grid %s block %s)" % (name, len(rs), rs[0][col["Grid Size"]], rs[0][col["Block Size"]]))
        for label, key in KEYS:
            if key not in col:
                continue
            v = [num(r[col[key]]) for r in rs]
            v = [x for x in v if x is not None]
            if v:
                print("  %-32s %14.3f %s" % (label, statistics.median(v), units[col[key]]))
        st = []
        for h, i in col.items():
            if h.startswith(STALL) and h.endswith("_per_issue_active.ratio"):
                v = [num(r[i]) for r in rs]
                v = [x for x in v if x is not None]
                if v:
                    st.append((statistics.median(v), h[len(STALL):-len("_per_issue_active.ratio")]))
        tot = sum(x for x, _ in st)
        if tot:
            print("  stalls: " + ", ".join("%s %.0f %%" % (n, 100 * x / tot) for x, n in sorted(st, reverse=True)[:7]))


if __name__ == "__main__":
    main()
