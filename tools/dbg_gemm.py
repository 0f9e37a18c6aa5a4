import ctypes, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_exp_b200 import _lib as L
from sevennet_exp_b200.engine import tc_image_for
lib = L.load(); st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
import os as _os
shape = eval(_os.environ.get("DBG_SHAPE", "(198356, 1, 960, 64, 3)"))
for (n, R, K, N, epi) in [shape]:
    A = torch.randn(n, R * K, device='cuda'); B = torch.randn(K, N, device='cuda')
    C = torch.zeros(n, R * N, device='cuda'); C2 = torch.zeros(n, R * N, device='cuda'); P = torch.randn(n, R * N, device='cuda')
    img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
    g = L.GemmArgs(A.data_ptr(), R * K, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, epi, 0, 1.679, C2.data_ptr(), P.data_ptr())
    for _ in range(2):
        lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(img.data_ptr()), st)
    torch.cuda.synchronize()
