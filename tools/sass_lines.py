#!/usr/bin/env python3
"""Static SASS accounting for one kernel of libassl.so (runs without a GPU).

    python tools/sass_lines.py logmel_kernelILi5 [--lines] [--src logmel.cu] [--lib path/to/libassl.so]

Prints the opcode histogram of the kernel and, with --lines, the instruction count per source line
(nvdisasm -g; the library is built with -lineinfo).  Static counts: loop bodies count once.
"""
import argparse
import collections
import os
import re
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kernel", help="substring of the mangled kernel name")
    ap.add_argument("--lib", default=os.path.join(ROOT, "audio-ssl_b200", "csrc", "libassl.so"))
    ap.add_argument("--lines", action="store_true")
    ap.add_argument("--src", default=None, help="only count lines of this source file (basename)")
    ap.add_argument("--range", default=None, help="with --lines: only source lines lo-hi")
    ap.add_argument("--ops", action="store_true", help="with --lines: opcode mix per line")
    args = ap.parse_args()
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", args.lib], cwd=td, check=True, capture_output=True)
        text = None
        for f in sorted(os.listdir(td)):
            out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, f)], capture_output=True, text=True).stdout
            if args.kernel in out:
                text = out
                break
    if text is None:
        raise SystemExit(f"kernel {args.kernel} not found")
    # split into functions
    blocks = re.split(r"\n\s*//-+ \.text\.", text)
    fn = [b for b in blocks if b.split("\n", 1)[0].find(args.kernel) >= 0]
    if not fn:
        raise SystemExit("kernel section not found")
    body = fn[0]
    print("kernel section:", body.split("\n", 1)[0].strip())
    hist = collections.Counter()
    per_line = collections.Counter()
    per_line_ops = collections.defaultdict(collections.Counter)
    cur = ("?", 0)
    n = 0
    for line in body.split("\n"):
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            if op == "NOP":
                continue
            hist[op] += 1
            per_line[cur] += 1
            per_line_ops[cur][op] += 1
            n += 1
    print(f"{n} instructions")
    print("  ".join(f"{k} {v}" for k, v in hist.most_common(40)))
    if args.lines:
        lo, hi = (0, 10 ** 9)
        if args.range:
            lo, hi = (int(x) for x in args.range.split("-"))
        for (f, ln), c in sorted(per_line.items()):
            if args.src and f != args.src:
                continue
            if not (lo <= ln <= hi):
                continue
            extra = ("   " + " ".join(f"{k}:{v}" for k, v in per_line_ops[(f, ln)].most_common(8))) if args.ops else ""
            print(f"{f}:{ln:5d}  {c:5d}{extra}")
        by_file = collections.Counter()
        for (f, ln), c in per_line.items():
            by_file[f] += c
        print("by file:", dict(by_file))


if __name__ == "__main__":
    main()
