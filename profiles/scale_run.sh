#!/bin/bash
# bench.py at N = 1, 2, 4, 8 (or the list given) on one box, one line per N; used for BASELINE.md §5's scaling rows
for n in ${@:-1 2 4 8}; do
  if [ $n = 1 ]; then timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_n$n.json
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_n$n.json; fi
  python -c "import sys,json; d=json.loads(open('gpurun_out/scale_n$n.json').read()); print(d['n_gpus'], '%.4g'%d['value'], '%.3f ms'%d['ms_per_step'], 'e2e %.4g'%d['e2e']['value'], d['config']['parallelism'], d['roofline']['kernel'][:30])"
done
