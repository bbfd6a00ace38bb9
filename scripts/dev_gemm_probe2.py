import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.ops import gemm
torch.manual_seed(0)
for mode in ("3d", "2d"):
    os.environ["OMLDM_MN_LOAD"] = mode
    for (M, N, K) in [(128, 32, 32), (128, 256, 64), (434, 434, 777)]:
        a = torch.randn(K, (M + 31) // 32 * 32, device="cuda"); b = torch.randn(K, (N + 31) // 32 * 32, device="cuda")
        out = gemm.gemm_outer(a, b, m=M, n=N); torch.cuda.synchronize()
        ref = (a[:, :M].double().t() @ b[:, :N].double()).float()
        print(mode, "outer", M, N, K, "maxerr", (out - ref).abs().max().item(), "refmax", ref.abs().max().item(),
              "outabsmax", out.abs().max().item(), "nnz", (out != 0).sum().item(), flush=True)
        if M == 128 and N == 32:
            print(out[:4, :6]); print(ref[:4, :6])
            # structured probe: a = one-hot rows to see which (k,m) element lands where
            a2 = torch.zeros_like(a); b2 = torch.zeros_like(b)
            a2[3, 5] = 1.0; b2[3, 7] = 2.0
            o2 = gemm.gemm_outer(a2, b2, m=M, n=N); torch.cuda.synchronize()
            print("onehot nz:", o2.nonzero().tolist(), o2[o2 != 0].tolist())
            a2.zero_(); b2.zero_(); a2[9, 40] = 1.0; b2[9, 1] = 3.0
            o2 = gemm.gemm_outer(a2, b2, m=M, n=N); torch.cuda.synchronize()
            print("onehot2 nz:", o2.nonzero().tolist(), o2[o2 != 0].tolist())
