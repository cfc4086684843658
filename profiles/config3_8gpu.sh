#!/bin/bash
# BASELINE.json configs[2]: 1M parallel episodes sharded across 8 GPUs (131 072 per GPU), 30-step rollouts
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline --envs-per-gpu 131072 --rollout-steps 30 2>gpurun_out/config3_8gpu.err | tail -1 > gpurun_out/config3_8gpu.json
python -c "import json; d=json.loads(open('gpurun_out/config3_8gpu.json').read()); print(d['n_gpus'], '%.4g'%d['value'], '%.3f ms'%d['ms_per_step'], 'e2e %.4g'%d['e2e']['value'], d['config']['episodes'], d['clocks'], d['roofline']['kernel'])"
