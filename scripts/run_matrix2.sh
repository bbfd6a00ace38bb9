#!/bin/bash
out=gpurun_out/matrix2.jsonl; : > $out
run() { n=$1; impl=$2; cfg=$3; shift 3
  if [ "$n" = 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py"; fi
  timeout 120 $cmd --gpus $n --steps 100 --warmup 10 --impl $impl --config $cfg "$@" 2>/dev/null | grep '^{"metric"' >> $out; }
for n in 1 2 4 8; do run $n ours orr; done
run 8 nccl orr --no-e2e
run 8 ours fgm
wc -l $out
