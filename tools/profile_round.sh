#!/bin/bash
# What profiles/ holds for one round, on one B200 -- one kind of run per call (one profiler run per box visit, each only
# after the plain run of the same command exited 0):
#   tools/profile_round.sh bench                       the bench line, the reference arm, the PCIe ceiling
#   tools/profile_round.sh launches                    the launch list (gpu__time_duration per launch)
#   tools/profile_round.sh logmel_kernel | two_views | running_stats      one full ncu capture of that kernel
# Outputs under gpurun_out/; summaries: tools/ncu_summary.py, tools/write_traffic.py.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
what=${1:-bench}
SHORT="--no-cpu --no-configs --steps 3 --warmup 3 --presteady-s 0.05 --sustained-s 0.05"
if [ "$what" = "bench" ]; then
  python bench.py --steps 100 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 1
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>/dev/null
  python tools/pcie_bw.py > gpurun_out/pcie_bw.txt 2>&1
  exit 0
fi
python bench.py $SHORT > gpurun_out/plain.log 2>&1 || exit 1
if [ "$what" = "launches" ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py $SHORT > gpurun_out/ncu_launches.log 2>&1
else
  ncu --set full --clock-control none --import-source on -k regex:$what -s 6 -c 1 -f -o gpurun_out/$what \
      python bench.py $SHORT > gpurun_out/ncu_$what.log 2>&1
fi
ls -la gpurun_out
