"""Developer probe: correctness + timing of the tcgen05 GEMMs (prints, never asserts)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.ops import gemm

def timeit(fn, n=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us

torch.manual_seed(0)
for (M, N, K) in [(128, 32, 32), (128, 256, 64), (300, 130, 257), (1024, 1024, 1024)]:
    a = torch.randn(M, (K + 3) // 4 * 4, device="cuda")[:, :K]; b = torch.randn(N, (K + 3) // 4 * 4, device="cuda")[:, :K]
    try:
        out = gemm.gemm_tn(a, b); torch.cuda.synchronize()
        ref = (a.double() @ b.double().t()).float()
        print("tn", M, N, K, "maxerr", (out - ref).abs().max().item(), "refmax", ref.abs().max().item(), flush=True)
    except Exception as e:
        print("tn", M, N, K, "EXC", e, flush=True)
for (M, N, K) in [(128, 32, 32), (128, 256, 64), (434, 434, 777), (1024, 1024, 1024)]:
    a = torch.randn(K, (M + 31) // 32 * 32, device="cuda"); b = torch.randn(K, (N + 31) // 32 * 32, device="cuda")
    try:
        out = gemm.gemm_outer(a, b, m=M, n=N); torch.cuda.synchronize()
        ref = (a[:, :M].double().t() @ b[:, :N].double()).float()
        print("outer", M, N, K, "maxerr", (out - ref).abs().max().item(), "refmax", ref.abs().max().item(), flush=True)
    except Exception as e:
        print("outer", M, N, K, "EXC", e, flush=True)
# timing: Gram 1024x1024 += Z^T Z
d = 1024
g = torch.zeros(d, d, device="cuda")
torch.backends.cuda.matmul.allow_tf32 = True
for B in (4096, 8192, 16384, 65536):
    z = torch.randn(B, d, device="cuda")
    for ks, up in ((1, False), (4, False), (7, True), (4, True), (8, True), (14, True)):
        t = timeit(lambda: gemm.gemm_outer(z, z, g, accumulate=True, ksplit=ks, upper_only=up))
        fl = 2.0 * B * d * d
        print(f"gram B={B} ksplit={ks} upper={up}: {t:.1f} us  {fl / t / 1e6:.1f} TFLOP/s(full-equiv)  {B / t:.2f} Mrec/s", flush=True)
    t = timeit(lambda: torch.addmm(g, z.t(), z, out=g))
    print(f"cublas tf32 addmm B={B}: {t:.1f} us {2.0 * B * d * d / t / 1e6:.1f} TFLOP/s", flush=True)
