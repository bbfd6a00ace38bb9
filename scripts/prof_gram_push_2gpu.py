"""The fp16 Gram kernel with its in-kernel hub push over NVLink, in ONE process that owns two GPUs (not a multi-rank run):
GPU 0 computes the Gram update of a 32 768 x 1003 batch and stores the rows owned by "worker 1" into a slot matrix that
lives on GPU 1.  The kernel waits for nobody, so ncu may replay it:

    python scripts/prof_gram_push_2gpu.py && ncu --set full --clock-control none --import-source on -k regex:gram_f16 \
        -s 3 -c 1 -f -o gpurun_out/prof/gram_push python scripts/prof_gram_push_2gpu.py

The ncu row then carries nvltx / nvlrx bytes of a real NVLink transfer issued from inside the GEMM kernel.  Checks the
pushed rows against the local running sum before printing ok.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R

assert torch.cuda.device_count() >= 2, "needs two GPUs in one process"
torch.cuda.set_device(0)
K, d = 32768, 1000
n, ldg = d + 3, F.pad64(d + 3)
rpo = 512                                              # rows per owner: worker 0 owns rows [0, 512), worker 1 the rest
pool = [R.generate_batch(R.KIND_REGRESSION, d, K, seed=1, batch_index=i, wire="f16s").cuda() for i in range(3)]
acc = torch.zeros(ldg, ldg, device="cuda:0")           # this worker's running sum
slot0 = torch.zeros(ldg, ldg, device="cuda:0")         # this worker's slot at owner 0 (itself)
slot1 = torch.zeros(ldg, ldg, device="cuda:1")         # this worker's slot at owner 1: the other GPU
assert torch.cuda.can_device_access_peer(0, 1)
import ctypes
rt = ctypes.CDLL("libcudart.so.12")                     # kernels on GPU 0 store into GPU 1's memory: explicit peer access
rc = rt.cudaDeviceEnablePeerAccess(ctypes.c_int(1), ctypes.c_uint(0))
assert rc in (0, 704), f"cudaDeviceEnablePeerAccess -> {rc}"      # 704 = already enabled
peers = F.GramPeers()
peers.world, peers.rank, peers.rows_per_owner = 2, 0, rpo
peers.owner_base[0] = slot0.data_ptr()
peers.owner_base[1] = slot1.data_ptr()
scratch = F.gram_f16_scratch("cuda:0")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(6):
    if i == 3:
        torch.cuda.synchronize()
        e0.record()
    F.gram_f16(pool[i % 3], n, acc, scratch, peers)
e1.record()
torch.cuda.synchronize()
torch.cuda.synchronize(1)
got = slot1.to("cuda:0")
i = torch.arange(ldg, device="cuda:0")
upper = ((i[None, :] // 32 * 32 + 31) >= (i[:, None] // 32 * 32)) & (i[:, None] >= rpo) & (i[:, None] < n) & (i[None, :] < n)
err = float((got - acc)[upper].abs().max())
moved = int(upper.sum()) * 4
print(f"ok: pushed rows match the running sum (max abs diff {err:.3g}); {moved / 1e6:.2f} MB per launch over NVLink; "
      f"{1000 * e0.elapsed_time(e1) / 3:.1f} us per launch")
assert err == 0.0
