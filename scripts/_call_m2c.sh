mkdir -p gpurun_out/m2f
timeout 300 python -m pytest tests/test_gram_f16_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/multigpu_check.py > gpurun_out/m2f/check.log 2>&1; echo "check rc $?"; grep -c PASS gpurun_out/m2f/check.log; grep -v PASS gpurun_out/m2f/check.log | tail -5
for mode in lead tail; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-baseline --tc hubPush=$mode > gpurun_out/m2f/orr2_$mode.json 2> gpurun_out/m2f/orr2_$mode.err; echo "bench $mode rc $?"; python -c "
import json; j=json.loads(open('gpurun_out/m2f/orr2_$mode.json').read().strip().splitlines()[-1]); print('$mode', round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us', j['sync_check'])"
done
