"""Per-kernel DRAM traffic and duration of one eager training step from an ncu pass with
`--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum` (cold caches: ncu flushes L2 between
kernel replays, so these are upper bounds on what a kernel moves inside the running step).

    ncu --profile-from-start off --clock-control none --page raw --csv --log-file raw.csv \
        --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum python bench.py --profile-step
    python profiles/step_traffic.py raw.csv profiles/r02_step_dram_traffic.json profiles/r02_gemm_dram_traffic.json

`bench.py` reads the first file for `roofline_hbm[*].traffic` and the second for `roofline.traffic`."""
import collections
import csv
import json
import re
import sys


def main(raw, out, out_gemm=None):
    rows = [r for r in csv.reader(l for l in open(raw) if not l.startswith("=="))]
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[h], rows[h + 1]
    ix = {k: i for i, k in enumerate(hdr)}
    scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}

    def val(r, k):
        return float(r[ix[k]].replace(",", "")) * scale.get(units[ix[k]].lower(), 1.0)

    agg = collections.OrderedDict()
    for r in rows[h + 2:]:
        if len(r) < len(hdr):
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").strip()
        a = agg.setdefault(name, {"launches": 0, "time_us": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0,
                                  "max_launch_bytes": 0.0, "max_launch_us": 0.0})
        rd, wr, t = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum"), val(r, "gpu__time_duration.sum")
        a["launches"] += 1
        a["time_us"] += t
        a["dram_read_bytes"] += rd
        a["dram_write_bytes"] += wr
        if rd + wr > a["max_launch_bytes"]:
            a["max_launch_bytes"], a["max_launch_us"] = rd + wr, t
    for a in agg.values():
        a["dram_bytes_per_launch"] = (a["dram_read_bytes"] + a["dram_write_bytes"]) / a["launches"]
    json.dump({"source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none "
                         "over bench.py --profile-step (one eager HuBERT Base prune step, 8 x 10 s, bf16)",
               "kernels": agg}, open(out, "w"), indent=1)
    tot = sum(a["time_us"] for a in agg.values())
    print("%d kernels, %d launches, %.1f us serialised" % (len(agg), sum(a["launches"] for a in agg.values()), tot))
    if out_gemm:
        g = [a for n, a in agg.items() if "gemm_bf16_kernel" in n]
        n = sum(a["launches"] for a in g)
        rd, wr = sum(a["dram_read_bytes"] for a in g), sum(a["dram_write_bytes"] for a in g)
        json.dump({"kernel": "gemm_bf16_kernel (tcgen05)",
                   "source": "profiles/step_traffic.py: the %d GEMM launches of one eager step (cold caches)" % n,
                   "launches": n, "dram_read_bytes": rd, "dram_write_bytes": wr,
                   "dram_bytes_per_launch": (rd + wr) / max(n, 1), "time_us_sum": sum(a["time_us"] for a in g)},
                  open(out_gemm, "w"), indent=1)
        print(open(out_gemm).read())


if __name__ == "__main__":
    main(*sys.argv[1:4])
