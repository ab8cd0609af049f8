"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total time and share per kernel.
usage: python tools/summarize_launches.py gpurun_out/launches.csv [header text] > profiles/rNN_launch_summary.txt"""
import csv, re, sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    rows = [l for l in open(path, newline="") if l.startswith('"')]
    agg = defaultdict(lambda: [0, 0.0])
    for r in csv.DictReader(rows):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = re.sub(r"nvllm::|\(anonymous namespace\)::", "", name)
        ns = float(r["Metric Value"].replace(",", ""))
        if r.get("Metric Unit") in ("us", "usecond"):
            ns *= 1e3
        agg[name][0] += 1; agg[name][1] += ns
    total = sum(v[1] for v in agg.values())
    n = sum(v[0] for v in agg.values())
    if len(sys.argv) > 2:
        print("# " + sys.argv[2])
    print(f"# launches={n}  total={total / 1e3:.1f} us   (per-launch times are cold-cache and serialised: compare SHARES)")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:60]:60s} n={c:4d} total={t / 1e3:9.1f} us  avg={t / c / 1e3:8.2f} us  share={100 * t / total:5.1f}%")


if __name__ == "__main__":
    main()
