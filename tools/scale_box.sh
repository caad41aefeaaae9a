#!/bin/bash
# The multi-GPU evidence of one round on ONE box with N GPUs: multi-GPU tests, bench at N = 1 and at every N given.
#   tools/scale_box.sh 8        (8-GPU box: tests at 2 and 8 ranks, bench N = 1 and 8)
#   tools/scale_box.sh 2 | 4    (smaller boxes: bench at that N only)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
n=${1:-8}
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus "$1" --steps 100 --warmup 5; }
if [ "$n" = "8" ]; then
  timeout 900 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q > gpurun_out/multigpu_tests.txt 2>&1; tail -3 gpurun_out/multigpu_tests.txt
  python bench.py --steps 100 --warmup 5 --no-cpu > gpurun_out/bench_n1_8gpu_box.json 2>/dev/null
  python tools/pcie_bw.py --all > gpurun_out/pcie_bw_n8.txt 2>&1
fi
tr "$n" > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
for f in gpurun_out/bench_n1_8gpu_box.json gpurun_out/bench_n$n.json; do
  [ -f "$f" ] && python -c "
import json,sys; d=json.load(open('$f')); print('$f', d['n_gpus'], d['value'], d['ms_per_step'], d.get('sharded_equals_unsharded'), d['e2e']['value'], d['e2e_pcm16']['value'], {k: round(v['value']) for k, v in d.get('configs', {}).items()})"
done
