#!/bin/bash
# First 1-GPU call of the next round (ROADMAP.md §1): GPU tests (reports XPASS for the features whose first hardware
# run is pending), then the bench in its default form and with each opt-in feature, then the timings of the new
# kernels.  Every step has its own timeout; results go to gpurun_out/first/.
#
#   gpurun --timeout 1200 -- 'bash scripts/first_gpu_call.sh'
out=gpurun_out/first
mkdir -p $out
timeout 600 python -m pytest tests -m gpu -q -rxX -p no:cacheprovider > $out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> $out/pytest_gpu.log
tail -25 $out/pytest_gpu.log

bench() { name=$1; shift
  timeout 240 python bench.py --steps 50 --warmup 5 "$@" > $out/$name.json 2> $out/$name.err
  echo "$name exit $? $(grep -o '"value": [0-9.e+]*' $out/$name.json | head -2 | tr '\n' ' ')"; }
bench orr_f32
bench orr_f16 --wire f16
bench orr_cg --hyper solver=cg
bench pa --config pa                        # linear_minibatch_kernel changed after its last hardware run
bench fgm --config fgm                      # poly_scale_linear_kernel changed, FGM kernel changed
bench kmeans --config kmeans
bench mcpa --config mcpa --steps 100
bench mcpa_direct --config mcpa --steps 100 --sparse-direct

timeout 300 python scripts/prof_pending.py > $out/prof_pending.log 2>&1
echo "prof_pending exit $?"
tail -30 $out/prof_pending.log
