#!/bin/bash
# Everything profiles/ holds for one round, on one B200: tests, the bench line, the reference arm, the launch list,
# one full ncu capture per kernel (each only after the plain run exited 0), the config sweep, the PCIe ceiling.
#   tools/profile_round.sh      (outputs under gpurun_out/)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 100 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>/dev/null
python bench.py --no-cpu --steps 20 --warmup 3 > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches.csv \
    python bench.py --no-cpu --steps 20 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
for k in logmel_kernel two_views running_stats; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/$k \
      python bench.py --no-cpu --steps 3 --warmup 2 > gpurun_out/ncu_$k.log 2>&1
done
bash tools/sweep.sh > gpurun_out/sweep_n1.jsonl 2> gpurun_out/sweep.err
python tools/pcie_bw.py > gpurun_out/pcie_bw.txt 2>&1
python tools/e2e_probe.py >> gpurun_out/pcie_bw.txt 2>&1
ls -la gpurun_out
