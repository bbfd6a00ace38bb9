N=$1
mkdir -p gpurun_out/fN
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/fN/orr$N.json 2> gpurun_out/fN/orr$N.err; echo "bench rc $?"; python -c "
import json; j=json.loads(open('gpurun_out/fN/orr$N.json').read().strip().splitlines()[-1]); print($N, round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us vs_base', j['vs_baseline'], 'e2e', round(j['e2e']['value']/1e6,1), j['sync_check']['ok'])"
