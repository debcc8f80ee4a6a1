"""Per-kernel table of ONE step from an `ncu --metrics gpu__time_duration.sum --csv` launch list of
tools/bench_workloads.py (steps are delimited by the optimizer's adam_kernel launches).

    python tools/launch_table.py gpurun_out/scanvi_list.csv 2 > table.txt    # 2 = Adam launches per step
"""
import collections
import csv
import sys


def main(path, adams_per_step):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    H = rows[h]
    ki, vi = H.index("Kernel Name"), H.index("Metric Value")
    L = [(r[ki].split("(")[0][-64:], float(r[vi].replace(",", "")) / 1000) for r in rows[h + 1:] if len(r) > vi]
    ad = [i for i, l in enumerate(L) if "adam_kernel" in l[0]]
    step = L[ad[-adams_per_step - 1] + 1: ad[-1] + 1]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, t in step:
        agg[n][0] += 1
        agg[n][1] += t
    tot = sum(t for _, t in step)
    print(f"launches in one step: {len(step)}   sum of ncu durations: {tot:.1f} us (cold-cache, serialised)")
    print(f"{'us':>9} {'share':>6} {'count':>5}  kernel")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t:9.1f} {100 * t / tot:5.1f}% {c:5d}  {n}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]))
