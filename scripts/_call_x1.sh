mkdir -p gpurun_out/x1
python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_gram_f16_gpu.py tests/test_wire_f16.py -m gpu -x -q > gpurun_out/x1/pytest.log 2>&1; echo "pytest rc $?"; tail -5 gpurun_out/x1/pytest.log
python bench.py --steps 20 --warmup 5 --no-baseline --no-e2e > gpurun_out/x1/orr.json 2> gpurun_out/x1/orr.err; echo "bench rc $?"; cut -c1-300 gpurun_out/x1/orr.json
for v in 128_3 32_12 64_5 diag; do
  OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_$v.so python -m pytest tests/test_gram_f16_gpu.py -m gpu -q -s -k "speed_report or matches" 2>&1 | grep -E "gram_f16 K|passed|failed" | sed "s/^/var $v: /"
done
python -m pytest tests/test_gram_f16_gpu.py -m gpu -q -s -k "speed_report" 2>&1 | grep -E "gram_f16 K|passed|failed" | sed "s/^/main: /"
OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so python -m pytest tests/test_gram_f16_gpu.py -m gpu -q -s -k "speed_report" 2>&1 | grep -E "gram_f16 K|cycles|CTAs|passed|failed" | sed "s/^/timing: /"
bash scripts/prof_full.sh
tail -3 gpurun_out/prof/full_ncu.log
