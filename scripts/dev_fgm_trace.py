import sys, os, math, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.ops import _lib
from omldm_b200.parallel.comm import PAD_FGM, PAD_PHI, PipelineComm
_i, _f, _u, _sz = ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_size_t
k, n, T = 4, 256, 1.0
comm = PipelineComm(n, "cuda", 0, 1, None, emulate=k)
e = torch.zeros(k, comm.n, device="cuda")
state = torch.zeros(8 * k, device="cuda"); state[1::8] = -1.0
fl = comm.pads[0].view(torch.float32)
fl[PAD_FGM + 2] = math.sqrt(T) / 2
fl[PAD_PHI:PAD_PHI + 16] = -math.sqrt(T)
for step in range(1, 45):
    for r in range(k): comm.bufs[r].add_(0.001 * (r + 1))
    _lib.check(_lib.kernels().omldm_sync_fgm(comm.struct_ptr(), _lib.ptr(e), _sz(comm.n), _f(1.0 / k), _i(0), _f(T),
               _f(0.01), _u(step), _lib.ptr(state), _i(0), _lib.stream_ptr()), "fgm")
    torch.cuda.synchronize()
    p0 = comm.pads[0]
    st = state.view(k, 8)
    print(step, "counter", int(p0[PAD_FGM]), "sub", int(p0[PAD_FGM+1]), "theta", round(float(fl[PAD_FGM+2]),4),
          "round", [int(comm.pads[r][PAD_FGM+3]) for r in range(k)], "phi", [round(float(x),3) for x in fl[PAD_PHI:PAD_PHI+k]],
          "zeta", [round(float(x),3) for x in st[:,2]], "sent", st[:,3].tolist(), "rounds", st[:,4].tolist(), flush=True)
