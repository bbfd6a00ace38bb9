mkdir -p gpurun_out/m8d
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 scripts/multigpu_check.py > gpurun_out/m8d/check.log 2>&1; echo "check rc $?"; grep -c PASS gpurun_out/m8d/check.log; tail -1 gpurun_out/m8d/check.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/m8d/orr8.json 2> gpurun_out/m8d/orr8.err; echo "bench rc $?"; python -c "
import json; j=json.loads(open('gpurun_out/m8d/orr8.json').read().strip().splitlines()[-1]); print(round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us vs_base', j['vs_baseline'], 'e2e', round(j['e2e']['value']/1e6,1), j['sync_check']['ok'])"
