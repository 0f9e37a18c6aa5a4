"""Micro-benchmark of the two dense-contraction arms (tcgen05 3xTF32 vs SIMT fp32) at the path's shapes."""
import ctypes
import sys
import os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_exp_b200 import _lib as L
from sevennet_exp_b200.engine import tc_image_for

lib = L.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
shapes = [(198356, 1, 64, 960, 0), (198356, 1, 960, 64, 3), (198356, 1, 64, 64, 2), (198356, 1, 8, 64, 2),
          (4096, 1, 224, 224, 0), (4096, 3, 384, 64, 0), (4096, 5, 352, 32, 0), (4096, 1, 128, 128, 0)]
only = [a for a in sys.argv[1:] if a in ('tc', 'simt')]
if 'cfg3' in sys.argv:        # SevenNet-l3i5 inner layer at 32,768 atoms: si2 forward, then si2 input-gradient
    shapes = [(32768, 1, 256, 256, 0), (32768, 3, 480, 64, 0), (32768, 5, 544, 32, 0), (32768, 7, 480, 32, 0),
              (32768, 1, 256, 256, 0), (32768, 3, 64, 480, 0), (32768, 5, 32, 544, 0), (32768, 7, 32, 480, 0)]
if 'one' in sys.argv:
    shapes = [(32768, 5, 32, 544, 0)]
flush = torch.empty(64 * 1024 * 1024, device='cuda')
for (n, R, K, N, epi) in shapes:
    A = torch.randn(n, R * K, device='cuda')
    B = torch.randn(K, N, device='cuda') / K ** 0.5
    C = torch.zeros(n, R * N, device='cuda')
    P = torch.randn(n, R * N, device='cuda')
    C2 = torch.zeros(n, R * N, device='cuda')
    img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
    g = L.GemmArgs(A.data_ptr(), R * K, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, epi, 0, 1.679,
                   C2.data_ptr(), P.data_ptr())
    res = {}
    for name in ('tc', 'simt'):
        if only and name not in only:
            continue
        ts = []
        for it in range(6):
            flush.fill_(0.0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            if name == 'tc':
                lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(img.data_ptr()), st)
            else:
                lib.call('sevenn_gemm_f32', ctypes.byref(g), st)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        res[name] = min(ts[1:])
    byt = 4 * n * R * (K + N * (2 if epi in (1, 2, 3) else 1))
    print(f'M={n * R:7d} K={K:4d} N={N:4d} epi={epi} ' + ' '.join(
        f'{k}: {v * 1e3:8.1f} us ({byt / v / 1e6:7.0f} GB/s)' for k, v in res.items()), flush=True)
