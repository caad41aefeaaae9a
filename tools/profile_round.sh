#!/bin/bash
# Everything profiles/ holds for one round, on one B200: the bench line, the reference arm, the launch list, one full ncu
# capture per kernel (each only after the plain run of the same command exited 0), the PCIe ceiling.
#   tools/profile_round.sh      (outputs under gpurun_out/)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SHORT="--no-cpu --no-configs --steps 3 --warmup 3 --presteady-s 0.05 --sustained-s 0.05"
python bench.py --steps 100 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>/dev/null
python bench.py $SHORT > gpurun_out/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py $SHORT > gpurun_out/ncu_launches.log 2>&1
for k in logmel_kernel two_views running_stats; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 1 -f -o gpurun_out/$k \
      python bench.py $SHORT > gpurun_out/ncu_$k.log 2>&1
done
python tools/pcie_bw.py > gpurun_out/pcie_bw.txt 2>&1
ls -la gpurun_out
