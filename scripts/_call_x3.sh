mkdir -p gpurun_out/x3
python -m pytest tests/test_gram_f16_gpu.py tests/test_engine_gpu.py tests/test_numerics_gate_gpu.py tests/test_oracle_parity_gpu.py -m gpu -x -q -s 2>&1 | grep -E "gram_f16 K|passed|failed|Error|error" | head -20
OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so python -m pytest tests/test_gram_f16_gpu.py -m gpu -q -s -k "speed_report" 2>&1 | grep -E "gram_f16 K|cycles|CTAs|passed|failed" | sed "s/^/timing: /"
python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/x3/orr.json 2> gpurun_out/x3/orr.err; echo "bench rc $?"; cut -c1-200 gpurun_out/x3/orr.json
