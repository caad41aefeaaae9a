#!/usr/bin/env python3
"""Per-source-line dynamic instruction counts and stall samples from an .ncu-rep captured with
--import-source on (kernels built with -lineinfo).

    python tools/ncu_lines.py gpurun_out/x.ncu-rep [--kernel regex] [--per N] [--top 40]
"""
import argparse
import csv
import io
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--kernel", default=None)
    ap.add_argument("--per", type=float, default=1.0)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--smem", action="store_true", help="sort by shared-memory wavefronts")
    a = ap.parse_args()
    cmd = ["ncu", "-i", a.rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--launch-count", "1"]
    if a.kernel:
        cmd += ["--kernel-name", f"regex:{a.kernel}"]
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    path, hdr, agg = None, None, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            path = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
            ie = hdr.index("Instructions Executed")
            ss = hdr.index("# Samples")
            wf = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None
            wx = hdr.index("L1 Wavefronts Shared Excessive") if "L1 Wavefronts Shared Excessive" in hdr else None
        elif hdr and r[0].isdigit():
            def num(v):
                try:
                    return float(v)
                except ValueError:
                    return 0.0
            agg.append((num(r[ie]), int(num(r[ss])), path, int(r[0]), r[1].strip(),
                        num(r[wf]) if wf is not None else 0.0, num(r[wx]) if wx is not None else 0.0))
    tot = sum(x[0] for x in agg)
    tots = sum(x[1] for x in agg)
    print(f"total warp instructions {tot:.0f} ({tot / a.per:.1f} per unit), stall samples {tots}")
    key = (lambda x: x[5]) if a.smem else (lambda x: x[0])
    for n, s, p, ln, src, w, x in sorted(agg, key=key, reverse=True)[:a.top]:
        print(f"{n / a.per:10.1f} {100 * n / tot:5.1f}%  samples {s:5d} {100 * s / max(tots, 1):5.1f}%  "
              f"smem wavefronts {w / a.per:8.1f} (excess {x / a.per:6.1f})  {p}:{ln}  {src[:90]}")


if __name__ == "__main__":
    main()
