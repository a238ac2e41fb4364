"""Write profiles/r02_traffic.json from this round's `ncu --set full` captures of the two headline kernels
(batch 8 @ 55x128, bf16x3, tiled): DRAM bytes, SM-side bytes and the duration, per launch, plus the commit
the capture was taken at.  bench.py prints these next to the algorithmic bytes (`roofline.traffic_ncu`).

    python tools/ncu_traffic.py gpurun_out/r02_build_tc_tiled_bf16x3.ncu-rep gpurun_out/r02_lookup_tiled_fwd.ncu-rep
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def metrics(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    get = lambda m: float(vals[hdr.index(m)].replace(",", "")) * UNIT[units[hdr.index(m)]]  # noqa: E731
    return {"name": vals[hdr.index("Kernel Name")], "dram_read": get("dram__bytes_read.sum"),
            "dram_write": get("dram__bytes_write.sum"), "sm_write": get("l1tex__m_l1tex2xbar_write_bytes.sum"),
            "sm_read": get("l1tex__m_xbar2l1tex_read_bytes.sum"), "us": get("gpu__time_duration.sum")}


def main(build_rep, lookup_rep):
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    out = {"config": "batch 8 @ 55x128, D=256, 4 levels, r=4, bf16x3, tiled layout", "commit": commit,
           "source": [os.path.basename(build_rep), os.path.basename(lookup_rep)],
           "kernels": {"build": metrics(build_rep), "lookup": metrics(lookup_rep)}}
    json.dump(out, open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(*sys.argv[1:3])
