#!/usr/bin/env python
"""r1_traffic.json from the `ncu --set full` capture of `bench.py --stages-only --stage-reps 1`:
DRAM read/write bytes of the LAST captured launch of each stage kernel (the first is the warm-up).

    python profiles/make_traffic.py gpurun_out/r1_full_stages.ncu-rep > profiles/r1_traffic.json
"""
import csv, io, json, re, subprocess, sys

STAGE = [("k_star2d<0", "jacobi"), ("k_rbgs2d", "rbgs"), ("k_tile2d", "block_jacobi"), ("k_star2d<1", "residual"),
         ("k_resrestrict2d", "residual_restrict"), ("k_prolong2d", "prolong_add"),
         ("k_rbgs3d", "rbgs3d"), ("k_resrestrict3d", "residual_restrict3d")]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(reports):
    stages = {}
    for rep in reports:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        idx = {h: i for i, h in enumerate(hdr)}
        for d in rows[2:]:
            name = re.sub(r"^void (mg::)?", "", d[idx["Kernel Name"]]).split("(")[0].replace("(int)", "").replace("(bool)", "")
            for prefix, stage in STAGE:
                if name.startswith(prefix):
                    def b(key):
                        return float(d[idx[key]].replace(",", "")) * SCALE.get(units[idx[key]], 1.0)
                    r, w = b("dram__bytes_read.sum"), b("dram__bytes_write.sum")
                    dur = float(d[idx["gpu__time_duration.sum"]].replace(",", ""))
                    dur *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(units[idx["gpu__time_duration.sum"]], 1.0)
                    stages[stage] = {"kernel": name, "grid": d[idx["Grid Size"]], "dram_read_bytes": r, "dram_write_bytes": w,
                                     "traffic_bytes": r + w, "duration_us_under_ncu": round(dur, 3)}
                    break
    print(json.dumps({
        "source": "ncu --set full captures of this round (profiles/r1_capture_final.sh): `bench.py --stages-only --stage-reps 1` "
                  "(4096^2 fp64, last launch of each stage) and the 3D kernels at 257^3",
        "note": "writes still dirty in the 126 MB L2 when the kernel ends are not counted by dram__bytes_write, so traffic can "
                "be below the algorithmic bytes",
        "stages": stages}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1:])
