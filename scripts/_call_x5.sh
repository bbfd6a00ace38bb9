export OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so
timeout 300 python -m pytest tests/test_gram_f16_gpu.py -m gpu -q -s -k "speed_report" 2>&1 | grep -E "gram_f16 K|cycles|CTAs|passed|failed"
timeout 300 python scripts/gram_in_step_timing.py 2>&1 | tail -20
