#!/bin/bash
# Bench matrix on an 8-GPU box; every JSON line goes to gpurun_out/matrix.jsonl
out=gpurun_out/matrix.jsonl; : > $out
run() { # n impl config extra...
  n=$1; impl=$2; cfg=$3; shift 3
  if [ "$n" = 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py"; fi
  timeout 150 $cmd --gpus $n --steps 100 --warmup 10 --impl $impl --config $cfg "$@" 2>/dev/null | grep '^{"metric"' >> $out
}
for n in 1 2 4 8; do run $n ours orr; done
run 1 nccl orr; run 8 nccl orr
OMLDM_USE_SYMM_MEM=1 run 8 ours orr --no-e2e
for c in pa fgm kmeans mcpa; do run 8 ours $c; run 8 nccl $c --no-e2e; run 1 ours $c --no-e2e; done
wc -l $out
