"""Reads the SOURCE page of an `ncu --set full --import-source on` report (no GPU needed) and prints, per kernel:
the instructions with the most warp-stall samples (with their top stall reasons), and the instruction mix of the hot
loop (instructions whose execution count lies in a given range).  This is how round 1 found that the prefill attention's
"FMA-pipe" exp2 polynomial still ran FRND.FLOOR / F2I.FLOOR on the XU pipe (profiles/r01_ncu_prefill_gemm.md).

usage: ncu -i report.ncu-rep --page source --csv > src.csv
       python tools/ncu_source_hot.py src.csv [--top 30] [--loop-exec LO HI]"""
import argparse
import collections
import csv


def blocks(path):
    rows = list(csv.reader(open(path)))
    cur = None
    for r in rows:
        if r and r[0] == "Kernel Name":
            if cur:
                yield cur
            cur = {"name": r[1], "hdr": None, "data": []}
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and len(r) == len(cur["hdr"]):
            cur["data"].append(r)
    if cur:
        yield cur


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv"); ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--loop-exec", type=float, nargs=2, default=None, help="execution-count range that identifies the hot loop")
    a = ap.parse_args()
    for b in blocks(a.csv):
        idx = {h: i for i, h in enumerate(b["hdr"])}

        def num(r, k):
            try:
                return float(r[idx[k]].replace(",", ""))
            except (ValueError, KeyError):
                return 0.0
        data = b["data"]
        tot = sum(num(r, "# Samples") for r in data) or 1.0
        stalls = [h for h in b["hdr"] if h.startswith("stall_") and "Not Issued" not in h]
        print(f"== {b['name'][:110]}\n   {len(data)} instructions, {int(tot)} warp-stall samples")
        for r in sorted(data, key=lambda r: -num(r, "# Samples"))[: a.top]:
            st = sorted(((num(r, k), k[6:]) for k in stalls), reverse=True)[:3]
            print(f"   {100 * num(r, '# Samples') / tot:5.1f}%  exec={num(r, 'Instructions Executed'):10.0f}  "
                  f"{r[idx['Source']][:70]:70s} {[(k, int(v)) for v, k in st if v > 0]}")
        if a.loop_exec:
            lo, hi = a.loop_exec
            body = [r for r in data if lo <= num(r, "Instructions Executed") <= hi]
            mix = collections.Counter()
            for r in body:
                t = r[idx["Source"]].split()
                mix[t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")] += 1
            share = 100 * sum(num(r, "# Samples") for r in body) / tot
            print(f"   hot loop: {len(body)} instructions per execution, {share:.1f}% of the samples; mix: {mix.most_common(16)}")


if __name__ == "__main__":
    main()
