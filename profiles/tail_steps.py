#!/usr/bin/env python
"""Aggregate the per-command clocks a -DMG_TAIL_TIMING build prints (every CTA of the cluster prints every step):
last program in the log, per (command kind, level index): count, cycles of the slowest CTA (a step that CTA 0 runs
alone shows the others' waiting time at the next cluster barrier: the minimum is taken when the maximum is absurd)."""
import collections, re, sys

NAMES = ["SMOOTH_POINT", "SMOOTH_TILE", "RESIDUAL", "RESTRICT", "RESIDUAL_RESTRICT", "ZERO", "PROLONG", "CGS", "BC"]


def main():
    lines = [l for l in open(sys.argv[1], errors="ignore").read().splitlines() if l.startswith("step ")]
    n = max(int(l.split()[1]) for l in lines[-400:]) + 1
    ctas = 16
    by, meta = collections.defaultdict(list), {}
    for l in lines[-ctas * n:]:
        m = re.match(r"step (\d+) kind (\d+) lv (\d+) colour (-?\d+) star (\d+) cycles (\d+)", l)
        s = tuple(int(x) for x in m.groups())
        by[s[0]].append(s[5])
        meta[s[0]] = s[1:5]
    agg, total = collections.defaultdict(lambda: [0, 0]), 0
    for i in sorted(by):
        v = sorted(by[i])
        c = v[-1] if v[-1] < 60000 or NAMES[meta[i][0]] == "CGS" else v[0]
        total += c
        k = (NAMES[meta[i][0]], meta[i][1])
        agg[k][0] += 1
        agg[k][1] += c
    print(f"{n} commands, {total} cycles = {total / 1.9e3:.1f} us at 1.9 GHz")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[0]:18s} lv {k[1]}  x{v[0]:<3d} {v[1]:8d} cycles  avg {v[1] // v[0]}")


if __name__ == "__main__":
    main()
