OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so timeout 300 python scripts/gram_in_step_timing.py 2>&1 | grep -E "overlapLoss|CTAs|timeline|Error|error"
