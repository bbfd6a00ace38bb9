#!/bin/bash
# One `ncu --set full` row for EVERY kernel of the library: the GPU tests that launch them (protocol kernels with
# emulated ranks in one cooperative launch, learners, preprocessors, tree, NN, GEMM variants, solvers) run under ncu,
# which captures the 2nd invocation of each kernel function whose demangled name contains "omldm::".  Shapes are the
# tests' shapes (small): registers / shared memory / grid / pipe mix are what these rows are for; the rows at bench
# shapes come from scripts/prof_full.sh.  Plain run first (must exit 0), then ONE ncu run; only the CSV travels back.
T="tests/test_sync_gpu.py tests/test_kernels_gpu.py tests/test_nn_gpu.py tests/test_ht_split_gpu.py tests/test_gemm_gpu.py tests/test_sparse_direct.py tests/test_ridge_cg.py tests/test_engine_gpu.py tests/test_wire_f16.py tests/test_gemm_scatter_gpu.py tests/test_gram_f16_gpu.py tests/test_zz_no_alloc.py"
out=gpurun_out/prof; mkdir -p $out
timeout 900 python -m pytest $T -m gpu -x -q > $out/every_plain.log 2>&1 || { echo "plain run failed"; tail -20 $out/every_plain.log; exit 1; }
tail -1 $out/every_plain.log
timeout 1500 ncu --set full --clock-control none --kernel-name-base demangled --kernel-id ::regex:omldm:2 \
    -f -o /tmp/every_kernel python -m pytest $T -m gpu -x -q > $out/every_ncu.log 2>&1
echo "ncu exit $?"; tail -2 $out/every_ncu.log
ncu -i /tmp/every_kernel.ncu-rep --page raw --csv > $out/every_kernel_raw.csv 2>/dev/null
ls -la /tmp/every_kernel.ncu-rep; wc -l $out/every_kernel_raw.csv
