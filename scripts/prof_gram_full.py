"""One fp16 Gram kernel call at the bench shape for `ncu --set full` (single GPU, a few launches):
    python scripts/prof_gram_full.py && ncu --set full --clock-control none --import-source on -k regex:gram_f16 -s 3 -c 1 \
        -o gpurun_out/prof/gram_f16 python scripts/prof_gram_full.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R

K, d = 32768, 1000
pool = [R.generate_batch(R.KIND_REGRESSION, d, K, seed=1, batch_index=i, wire="f16s").cuda() for i in range(3)]
C = torch.zeros(1024, 1024, device="cuda")
st = torch.zeros(4, dtype=torch.float64, device="cuda")
wt = torch.zeros(d + 3, device="cuda")
wt[d + 1:] = -1
for i in range(6):
    F.gram_f16(pool[i % 3], d + 3, C)
    F.sq_residual_f16s(pool[i % 3], d, wt, st)
torch.cuda.synchronize()
print("ok", float(C[0, 0]), float(st[0]))
