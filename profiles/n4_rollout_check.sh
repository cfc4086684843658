for rs in 0 200; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 2952$((rs/100)) bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu-baseline --rollout-steps $rs 2>/dev/null | tail -1 > gpurun_out/n4_$rs.json
python -c "import sys,json; d=json.loads(open('gpurun_out/n4_$rs.json').read()); n=d['config']['env_steps_per_bench_step']; print('rollout',n,'value %.4g'%d['value'], 'us/step resident %.2f'%(d['ms_per_step']*1e3/n), 'e2e us/step %.2f'%(d['e2e']['ms_per_step']*1e3/n), 'timed200 %.2f'%d['roofline']['us_per_env_step'], d['clocks'])"
done
