mkdir -p gpurun_out/m2e gpurun_out/prof
timeout 120 python scripts/prof_gram_push_2gpu.py > gpurun_out/m2e/push_plain.log 2>&1; rc=$?; echo "push plain rc $rc"; tail -2 gpurun_out/m2e/push_plain.log
if [ $rc -eq 0 ]; then
  timeout 300 ncu --set full --clock-control none -k regex:gram_f16 -s 3 -c 1 -f -o /tmp/gram_push python scripts/prof_gram_push_2gpu.py > gpurun_out/m2e/push_ncu.log 2>&1; echo "ncu rc $?"
  ncu -i /tmp/gram_push.ncu-rep --page raw --csv > gpurun_out/prof/gram_push_2gpu_raw.csv 2>/dev/null; wc -l gpurun_out/prof/gram_push_2gpu_raw.csv
fi
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-baseline > gpurun_out/m2e/orr2b.json 2> gpurun_out/m2e/orr2b.err; echo "bench rc $?"; wc -l gpurun_out/m2e/orr2b.json; cut -c1-220 gpurun_out/m2e/orr2b.json
