"""Summarise `ncu -i X.ncu-rep --page raw --csv` exports: one block of key metrics per profiled launch."""
import csv
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor",
        "sm__inst_executed_pipe_tensor_op_gmma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_umma_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum"]


def main(path, maxn=6):
    rows = list(csv.reader(open(path)))
    hdr = rows[0]
    units = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
    if not stall:
        stall = [h for h in hdr if "warp_issue_stalled" in h and h.endswith(".pct")]
    for r in rows[2:2 + maxn]:
        print("== %s  grid=%s block=%s" % (r[idx["Kernel Name"]][:70], r[idx.get("Grid Size", 0)], r[idx.get("Block Size", 0)]))
        for k in KEYS:
            if k in idx:
                print("   %-75s %s %s" % (k, r[idx[k]], units[idx[k]]))
        st = []
        for h in stall:
            try:
                st.append((float(r[idx[h]].replace(",", "")), h))
            except ValueError:
                pass
        for v, h in sorted(st, reverse=True)[:5]:
            print("   stall %-69s %.2f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 6)
