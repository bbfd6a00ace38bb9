mkdir -p gpurun_out/m2e gpurun_out/prof
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/multigpu_check.py > gpurun_out/m2e/check.log 2>&1; echo "check rc $?"; tail -3 gpurun_out/m2e/check.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/m2e/orr2.json 2> gpurun_out/m2e/orr2.err; echo "bench rc $?"; cut -c1-220 gpurun_out/m2e/orr2.json
timeout 120 python scripts/prof_gram_push_2gpu.py > gpurun_out/m2e/push_plain.log 2>&1; rc=$?; echo "push plain rc $rc"; tail -2 gpurun_out/m2e/push_plain.log
if [ $rc -eq 0 ]; then
  timeout 300 ncu --set full --clock-control none -k regex:gram_f16 -s 3 -c 1 -f -o /tmp/gram_push python scripts/prof_gram_push_2gpu.py > gpurun_out/m2e/push_ncu.log 2>&1; echo "ncu rc $?"
  ncu -i /tmp/gram_push.ncu-rep --page raw --csv > gpurun_out/prof/gram_push_2gpu_raw.csv 2>/dev/null; wc -l gpurun_out/prof/gram_push_2gpu_raw.csv
fi
