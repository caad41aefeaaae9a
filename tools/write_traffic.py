#!/usr/bin/env python3
"""profiles/traffic.json from one `ncu --set full` capture of logmel_kernel (runs here, no GPU): DRAM bytes per launch,
stamped with the hash of the kernel's sources so that bench.py refuses the figure once the kernel changes.

    python tools/write_traffic.py gpurun_out/r02_logmel.ncu-rep
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    to_bytes = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in rows[2:]:
        if "logmel_kernel" not in r[ix["Kernel Name"]]:
            continue
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[ix[m]].replace(",", "")) * to_bytes[units[ix[m]]]
        doc = {"logmel_kernel_bytes_per_launch": int(tot), "source_sha": bench.kernel_source_hash(),
               "capture": f"ncu --set full --clock-control none, one launch of {r[ix['Kernel Name']]} over 1024 x 1 s clips "
                          f"({os.path.basename(rep)}); dram__bytes_read.sum + dram__bytes_write.sum"}
        with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as fh:
            json.dump(doc, fh, indent=1)
        print(doc)
        return
    raise SystemExit("no logmel_kernel launch in the report")


if __name__ == "__main__":
    main()
