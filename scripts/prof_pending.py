"""Stand-alone launches of the kernels written after round 1's GPU budget was spent, for the first ncu capture:

    gpurun -- 'python scripts/prof_pending.py > gpurun_out/pending_plain.log 2>&1 && \
      ncu --set full --clock-control none --import-source on \
          -k regex:"ridge_pcg|mcpa_sparse|stage_rows_f16|sparse_rows|scatter" -c 24 \
          -o gpurun_out/prof_pending python scripts/prof_pending.py'
    python scripts/make_profiles.py gpurun_out/prof_pending.ncu-rep pending

Prints CUDA-event timings (after warm-up, larger-than-L2 rotation where it matters)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from omldm_b200.ops import functional as F
from omldm_b200.stream import ring as R

dev = "cuda"


def timed(name, fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {1000 * e0.elapsed_time(e1) / reps:.1f} us", flush=True)


# ---- ORR refresh: Cholesky vs warm-started CG, d = 1000 (ridge system of a stream of 64 k records)
d, ld = 1000, 1024
X = R.generate_batch(R.KIND_REGRESSION, d, 65536, seed=1).to(dev)
W = torch.zeros(ld, ld, device=dev)
W[: d + 2, : d + 2] = X[:, : d + 2].t() @ X[:, : d + 2]
n = d + 1
work, scal, info = F.ridge_cg_workspace(n, dev, 32)
x = torch.zeros(n, device=dev)
F.ridge_cg(W, n, 1.0, x, work, scal, info, 32, 1e-5)
torch.cuda.synchronize()
print("cg cold start: iterations, relres =", info.tolist(), flush=True)
X2 = R.generate_batch(R.KIND_REGRESSION, d, 32768, seed=1, batch_index=5).to(dev)
W[: d + 2, : d + 2] += X2[:, : d + 2].t() @ X2[:, : d + 2]          # the system moves by one refresh interval
x_warm = x.clone()


def cg_warm():
    x.copy_(x_warm)
    F.ridge_cg(W, n, 1.0, x, work, scal, info, 32, 1e-5)


def chol():
    A = W[:n, :n].clone()
    A.diagonal().add_(1.0)
    L, _ = torch.linalg.cholesky_ex(A, check_errors=False)
    torch.cholesky_solve(W[:n, n].clone()[:, None], L)


timed("ridge_cg warm start (incl. 4 KB copy)", cg_warm)
print("   iterations, relres =", info.tolist(), flush=True)
timed("cholesky_ex + cholesky_solve", chol)

# ---- wire format expansion, B = 32768, d = 1000 (67 MB in, 134 MB out)
wires = [R.generate_batch(R.KIND_REGRESSION, d, 32768, seed=2, batch_index=i, wire="f16").to(dev) for i in range(3)]
out = torch.empty(32768, ld, device=dev)
i = [0]


def stage():
    F.stage_rows_f16(wires[i[0] % 3], d, out=out)
    i[0] += 1


timed("stage_rows_f16 32768 x 1000", stage)

# ---- sparse MultiClassPA step, 100k-d model, 128 non-zeros per record
D, nnz, C = 100000, 128, 10
for B in (256, 16384):
    slots = [R.generate_sparse_batch(D, B, nnz, n_aux=C, seed=3, batch_index=j).to(dev) for j in range(4)]
    Wm = torch.zeros(C, F.pad32(D + 2), device=dev)
    st = torch.zeros(4, dtype=torch.float64, device=dev)
    coef = torch.empty(B, 4, device=dev)
    rows = torch.empty(B, F.sparse_row_pitch(nnz), device=dev)
    j = [0]

    def sparse_step():
        idx, val, tgt = R.SparseTopic.unpack(slots[j[0] % 4], B, nnz)
        j[0] += 1
        F.sparse_rows(idx, val, tgt, B, nnz, out=rows)
        prev = st[3].clone()
        F.mcpa_sparse_step(rows, nnz, Wm, D, C, 1, 0.1, coef, st, prev, True)

    timed(f"sparse_rows + mcpa_sparse_step B={B}", sparse_step)
