mkdir -p gpurun_out/x8
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py tests/test_sparse_direct.py -m gpu -x -q 2>&1 | tail -4
for c in orr; do
timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/x8/$c.json 2> gpurun_out/x8/$c.err; echo "bench rc $?"; python -c "
import json; j=json.loads(open('gpurun_out/x8/$c.json').read().strip().splitlines()[-1]); print('$c piped', round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us', j['gpu_launches'])"
timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-baseline --no-e2e --no-pipelined-holdout > gpurun_out/x8/${c}_np.json 2> gpurun_out/x8/${c}_np.err; python -c "
import json; j=json.loads(open('gpurun_out/x8/${c}_np.json').read().strip().splitlines()[-1]); print('$c plain', round(j['value']/1e6,1),'M', round(j['ms_per_step']*1e3,2),'us', j['gpu_launches'])"
done
