"""Key metrics of an `ncu --set full` report exported with `ncu -i X.ncu-rep --page raw --csv`.

    ncu -i gpurun_out/r1_full_stages.ncu-rep --page raw --csv > /tmp/full.csv
    python profiles/summarise_full.py /tmp/full.csv > profiles/r1_full_stages.md
"""
import csv
import re
import sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM % of peak"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__occupancy_limit_registers", "blocks/SM (reg limit)"),
    ("launch__waves_per_multiprocessor", "waves/SM"),
    ("smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "stall long scoreboard %"),
    ("smsp__warp_issue_stalled_barrier_per_warp_active.pct", "stall barrier %"),
    ("smsp__inst_executed.sum", "warp instructions"),
]


def to_bytes(value, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(value.replace(",", "")) * scale.get(unit, 1.0)


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    for d in data:
        name = re.sub(r"^void ", "", d[idx["Kernel Name"]]).split("(")[0]
        print(f"### `{name}`  grid {d[idx['Grid Size']]} block {d[idx['Block Size']]}\n")
        traffic = 0.0
        for key, label in WANT:
            if key not in idx:
                continue
            v, u = d[idx[key]], units[idx[key]]
            if key.startswith("dram__bytes_"):
                b = to_bytes(v, u)
                traffic += b
                print(f"- {label}: {b / 1e6:.1f} MB")
            else:
                print(f"- {label}: {v} {u}")
        print(f"- dram traffic (read+write): {traffic / 1e6:.1f} MB\n")


if __name__ == "__main__":
    main(sys.argv[1])
