#!/bin/bash
# A/B of the deferred log-mel launch (ASSL_FRONT_OVERLAP=0/1), one GPU.
cd "$(dirname "$0")/.."
run() { env "$@" python bench.py --no-cpu --no-configs --steps 300 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['kernel_ms'])"; }
for rep in 1 2 3; do
  echo "overlap=0"; run ASSL_FRONT_OVERLAP=0
  echo "overlap=1"; run ASSL_FRONT_OVERLAP=1
done
