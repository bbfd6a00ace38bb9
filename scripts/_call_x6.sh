mkdir -p gpurun_out/x6
timeout 600 python -m pytest tests/test_gram_f16_gpu.py tests/test_engine_gpu.py tests/test_oracle_parity_gpu.py tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/x6/orr.json 2> gpurun_out/x6/orr.err; echo "bench rc $?"; cut -c1-200 gpurun_out/x6/orr.json
OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so timeout 300 python scripts/gram_in_step_timing.py 2>&1 | grep -E "overlapLoss|CTAs"
