"""Top stall sites of an `ncu --page source --csv` export (gzip ok): SASS lines ranked by warp-stall samples, with the
dominant stall reason of each and a little context."""
import csv
import gzip
import sys


def main(path, top=25, ctx=0):
    op = gzip.open if path.endswith(".gz") else open
    rows = list(csv.reader(op(path, "rt")))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    body = [r for r in rows[h + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]  # (both captured launches)
    ix = {n: i for i, n in enumerate(hdr)}
    s_i = ix["# Samples"]
    stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    tot = sum(int(r[s_i] or 0) for r in body)
    print("total samples %d over %d SASS lines, %d instructions executed" % (
        tot, len(body), sum(int(r[ix["Instructions Executed"]] or 0) for r in body)))
    agg = {}
    for r in body:
        for n in stalls:
            agg[n] = agg.get(n, 0) + int(r[ix[n]] or 0)
    print("by reason: " + ", ".join("%s %.1f%%" % (k[6:], 100.0 * v / max(tot, 1)) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    order = sorted(range(len(body)), key=lambda i: -int(body[i][s_i] or 0))[:top]
    for i in order:
        r = body[i]
        n = int(r[s_i] or 0)
        why = max(stalls, key=lambda k: int(r[ix[k]] or 0))
        print("%5.1f%% #%-5d %-14s exec %-7s %s" % (100.0 * n / max(tot, 1), i, why[6:], r[ix["Instructions Executed"]], r[ix["Source"]].strip()[:90]))
        for j in range(max(0, i - ctx), i):
            print("          .      %s" % body[j][ix["Source"]].strip()[:90])


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25, int(sys.argv[3]) if len(sys.argv) > 3 else 0)
