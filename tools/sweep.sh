#!/bin/bash
# Throughput sweep over clip lengths and batch sizes on one GPU, one bench.py JSON line each (bench.py itself reports
# BASELINE configs 3-5 in its "configs" block; this is the finer grid).
#   tools/sweep.sh > gpurun_out/sweep_n1.jsonl
set -u
cd "$(dirname "$0")/.."
run() { python bench.py --no-cpu --no-configs --presteady-s 0.3 --sustained-s 0.3 --steps "$3" --warmup 5 --batch "$1" --seconds "$2" 2>/dev/null; }
run 512 0.95 100      # C3: DeLoRes-S shard (4096 / 8 GPUs), length_wave 0.95
run 256 10 50         # C4: DeCAR / DeLoRes-M shard (2048 / 8 GPUs), 10 s clips
run 256 1 100         # C5 grid
run 4096 1 50
run 16384 1 20
run 1024 2 50
run 1024 5 30
run 1024 10 20
run 256 20 20
run 256 30 20
