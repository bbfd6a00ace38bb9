mkdir -p gpurun_out/x4
run() { name=$1; shift; python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e "$@" > gpurun_out/x4/$name.json 2> gpurun_out/x4/$name.err; python -c "
import json,sys
j=json.loads(open('gpurun_out/x4/$name.json').read().strip().splitlines()[-1]); print('$name', round(j['ms_per_step']*1e3,2),'us/step', j['gpu_launches'],'launches', j['windows'],'windows')"; }
run default
run nosolve --hyper solveEvery=1000000
run serial --hyper overlapLoss=false
run serial_nosolve --hyper overlapLoss=false --hyper solveEvery=1000000
run b16k --batch 16384
