"""Summarise an .ncu-rep into the few numbers the roofline discussion uses.
usage: python tools/ncu_summary.py gpurun_out/prof_build.ncu-rep > profiles/r01_build_tc.txt"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__m_l1tex2xbar_write_bytes.sum", "l1tex__m_xbar2l1tex_read_bytes.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_srcunit_tex_op_read.sum", "dram__sectors_read.sum",
    "sm__cycles_elapsed.max",
]


def main(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print(f"## {name[:110]}")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"{w:85s} {vals[i]:>18s} {units[i]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    if len(rows) > 3:
        hdr = rows[1]
        ix = {h: i for i, h in enumerate(hdr)}
        ins = rows[2:]
        stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        tot = {h: sum(int(r[ix[h]] or 0) for r in ins if len(r) > ix[h]) for h in stalls}
        s = sum(tot.values()) or 1
        print("## warp stall samples (all):", ", ".join(f"{k[6:]} {100 * v / s:.0f}%" for k, v in sorted(tot.items(), key=lambda x: -x[1])[:6]))
        ops = set()
        for r in ins:
            tk = r[ix["Source"]].split()
            if tk:
                ops.add((tk[1] if tk[0].startswith("@") and len(tk) > 1 else tk[0]).split(".")[0])
        marks = [m for m in ("UTCHMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "SYNCS", "UTCBAR", "HMMA", "LDGSTS") if m in ops]
        print("## SASS mnemonics present:", ", ".join(marks))


if __name__ == "__main__":
    main(sys.argv[1])
