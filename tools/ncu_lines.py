"""Per-source-line instruction and stall-sample counts of one kernel.
Joins `ncu --page source --csv` (SASS rows, in address order) with `nvdisasm --print-line-info` of the same
kernel extracted from the library, by instruction index.  The library must be the build that was profiled.
usage: python tools/ncu_lines.py <rep> <kernel substring> [lib.so]"""
import csv
import re
import subprocess
import sys
import tempfile
import collections
import os


def main(rep, kern, lib="raft_ros_b200/libraftcorr.so"):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    sass = rows[2:]
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=td, check=True, capture_output=True)
        lines = None
        for f in sorted(os.listdir(td)):
            out = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(td, f)], capture_output=True, text=True).stdout
            if kern not in out:
                continue
            cur, got, inside = None, [], False
            for ln in out.splitlines():
                if ln.startswith("//--------------------- .text."):
                    inside = kern in ln
                    continue
                if not inside:
                    continue
                m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
                if m:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
                    got.append(cur)
            if got:
                lines = got
                break
    if lines is None:
        raise SystemExit("kernel not found in " + lib)
    n = min(len(lines), len(sass))
    if len(lines) != len(sass):
        print(f"# warning: {len(lines)} disassembled vs {len(sass)} profiled instructions", file=sys.stderr)
    inst, samp = collections.Counter(), collections.Counter()
    for i in range(n):
        key = lines[i]
        inst[key] += int(sass[i][ix["Instructions Executed"]] or 0)
        samp[key] += int(sass[i][ix["# Samples"]] or 0)
    ti, ts = sum(inst.values()) or 1, sum(samp.values()) or 1
    print(f"# {kern}: {ti} warp instructions, {ts} samples")
    for key, v in sorted(inst.items(), key=lambda x: -x[1])[:40]:
        print(f"{100 * v / ti:5.1f}% instr {100 * samp[key] / ts:5.1f}% samples  {key}")


if __name__ == "__main__":
    main(*sys.argv[1:])
