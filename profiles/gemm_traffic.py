"""Per-launch DRAM traffic of the tcgen05 GEMM from an ncu pass over one eager step (the dram__bytes_* counters are
the ones `--set full` reports; a full-set capture of all 239 launches is > 64 MiB and does not fit gpurun_out/).

    ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
        --clock-control none -k regex:gemm_bf16_kernel --page raw --csv --log-file raw.csv python bench.py --profile-step
    grep -v "^==" raw.csv > raw2.csv; python profiles/gemm_traffic.py raw2.csv profiles/r01_gemm_dram_traffic.json
"""
import csv
import json
import sys


def main(raw, out):
    rows = list(csv.reader(open(raw)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, k):
        v = float(r[ix[k]].replace(",", ""))
        u = units[ix[k]].lower()
        return v * {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)

    rd = wr = t = 0.0
    n = 0
    for r in rows[2:]:
        if "gemm_bf16_kernel" not in r[ix["Kernel Name"]]:
            continue
        rd += val(r, "dram__bytes_read.sum")
        wr += val(r, "dram__bytes_write.sum")
        t += val(r, "gpu__time_duration.sum")
        n += 1
    json.dump({"kernel": "gemm_bf16_kernel (tcgen05)",
               "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, bench.py --profile-step, the %d GEMM launches of one eager "
                         "step (cold caches: ncu flushes L2 between replays)" % n,
               "launches": n, "dram_read_bytes": rd, "dram_write_bytes": wr, "dram_bytes_per_launch": (rd + wr) / max(n, 1),
               "time_us_sum": t}, open(out, "w"), indent=1)
    print(open(out).read())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
