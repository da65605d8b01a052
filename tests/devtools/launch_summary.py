"""Kernel shares of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`).

    python tools/launch_summary.py gpurun_out/r1d_launches_bench_steps4.csv profiles/r1d_launch_summary.csv
"""
import collections
import csv
import sys


def main():
    src, out = sys.argv[1], sys.argv[2]
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    head = rows[0]
    ik, iv, iu = head.index("Kernel Name"), head.index("Metric Value"), head.index("Metric Unit")
    tot = collections.Counter()
    cnt = collections.Counter()
    scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3, "second": 1e6, "s": 1e6}
    for r in rows[1:]:
        name = r[ik][:70]
        tot[name] += float(r[iv].replace(",", "")) * scale.get(r[iu], 1.0)
        cnt[name] += 1
    all_us = sum(tot.values())
    with open(out, "w") as f:
        f.write("kernel,launches,total_us,share\n")
        for name, us in tot.most_common():
            f.write(f'"{name}",{cnt[name]},{us:.1f},{us / all_us:.4f}\n')
    print(open(out).read()[:1200])


if __name__ == "__main__":
    main()
