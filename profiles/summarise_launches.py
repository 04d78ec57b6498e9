"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`).

    python profiles/summarise_launches.py gpurun_out/r1_launch_list.csv > profiles/r1_launch_summary.md
"""
import csv
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"^void ", "", name)
    m = re.match(r"([A-Za-z0-9_:]+(<[^(]*>)?)\(", name)
    return m.group(1) if m else name[:60]


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] == "gpu__time_duration.sum":
            rows.append((short(r["Kernel Name"]), r["Grid Size"], r["Block Size"], float(r["Metric Value"]) * 1e-3))
    total = sum(r[3] for r in rows)
    by = defaultdict(lambda: [0, 0.0, 0.0])
    for name, grid, block, us in rows:
        e = by[name]
        e[0] += 1
        e[1] += us
        e[2] = max(e[2], us)
    print(f"launches profiled: {len(rows)}, sum of kernel durations {total / 1e3:.2f} ms "
          f"(ncu serialises launches and starts each with cold caches)\n")
    print("| kernel | launches | total us | share | mean us | max us |")
    print("|---|---:|---:|---:|---:|---:|")
    for name, (n, us, mx) in sorted(by.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {n} | {us:.1f} | {us / total:.3f} | {us / n:.2f} | {mx:.1f} |")
    big = defaultdict(lambda: [0, 0.0])
    for name, grid, block, us in rows:
        if us > 30.0:
            big[(name, grid, block)][0] += 1
            big[(name, grid, block)][1] += us
    print("\nLaunches longer than 30 us (finest-level kernels), by launch shape:\n")
    print("| kernel | grid | block | launches | mean us |")
    print("|---|---|---|---:|---:|")
    for (name, grid, block), (n, us) in sorted(big.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {grid} | {block} | {n} | {us / n:.1f} |")


if __name__ == "__main__":
    main(sys.argv[1])
