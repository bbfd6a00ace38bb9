"""Per-launch time and phase cycles of the fp16 Gram kernel with and without the loss formed inside it (tuning build:
scripts/build_variant.sh timing csrc/gram_f16.cu csrc/rows_f16.cu csrc/ingest.cu -- -DOMLDM_GRAM_TIMING -DOMLDM_TIMELINE;
OMLDM_KERNELS_LIB=omldm_b200/lib/libvar_timing.so python scripts/gram_loss_inside_timing.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, "tests")
import torch
from omldm_b200.ops import functional as F
from test_gram_f16_gpu import _rows, gram_phase_report
d = 1000; n = d + 3; ldg = F.pad64(n); K = 32768
pool = [_rows(K, d, seed=s) for s in range(3)]
C = torch.zeros(ldg, ldg, device="cuda")
wt = torch.randn(n, device="cuda") * 0.1; wt[d + 1:] = -1.0
st = torch.zeros(4, dtype=torch.float64, device="cuda")
dots, words = F.gram_loss_workspace(K, "cuda")
desc = F.make_gram_loss(wt, d, st, dots, words)
for use_loss in (False, True):
    for i in range(3):
        F.gram_f16(pool[i], n, C, loss=desc if use_loss else None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(30):
        F.gram_f16(pool[i % 3], n, C, loss=desc if use_loss else None)
    e1.record(); torch.cuda.synchronize()
    print(f"loss inside = {use_loss}: {1000 * e0.elapsed_time(e1) / 30:.1f} us per call")
    for l in gram_phase_report():
        print(l)

import ctypes, numpy as np
from omldm_b200.ops import _lib
lib = _lib.kernels()
if hasattr(lib, "omldm_gram_f16_trace"):
    buf = np.zeros(2 * 5 * 24, dtype=np.int64)
    lib.omldm_gram_f16_trace(buf.ctypes.data_as(ctypes.c_void_p))
    tr = buf.reshape(2, 5, 24)
    t0 = tr[0, 0, 0]
    print("tile 0 slice 0, stage uses 30..41 (cycles relative to the first): E = producer got the free stage, F = MMA warp saw the data, L = loss warp saw the data, A = loss warp released")
    for rk in (0, 1):
        for kb in range(12):
            print(f"  rank {rk} use {30 + kb}: E {tr[rk,0,kb]-t0:7d}  F {tr[rk,2,kb]-t0:7d}  L {tr[rk,3,kb]-t0:7d}  A {tr[rk,4,kb]-t0:7d}")
