mkdir -p gpurun_out/x2
python -m pytest tests/test_sparse_direct.py tests/test_oracle_parity_gpu.py -m gpu -x -q > gpurun_out/x2/pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/x2/pytest.log
python bench.py --config mcpa --steps 20 --warmup 5 > gpurun_out/x2/mcpa.json 2> gpurun_out/x2/mcpa.err; echo "mcpa rc $?"; cut -c1-260 gpurun_out/x2/mcpa.json
bash scripts/prof_full.sh
tail -2 gpurun_out/prof/full_ncu.log | cut -c1-300
