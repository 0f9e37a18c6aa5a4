import ctypes, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_exp_b200 import _lib as L
lib = L.load(); st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for (M, R, K, N) in [(148310, 1, 64, 960), (148310, 1, 64, 64), (148310, 1, 8, 64), (5307, 1, 224, 224), (5307, 3, 384, 64), (5307, 5, 352, 32), (5307, 1, 128, 128), (5307, 5, 32, 32)]:
    A = torch.randn(M, R * K, device='cuda'); B = torch.randn(M, R * N, device='cuda')
    ns = int(lib.cdll.sevenn_gemm_tn_splits(M * R))
    part = torch.empty(ns * K * N, device='cuda'); out = torch.zeros(K, N, device='cuda')
    def run():
        lib.call('sevenn_gemm_tn', ctypes.c_void_p(A.data_ptr()), R * K, K, ctypes.c_void_p(B.data_ptr()), R * N, N, M * R, R, K, N,
                 None, 0, 1.0, 0, ctypes.c_void_p(part.data_ptr()), ctypes.c_void_p(out.data_ptr()), st)
    for _ in range(3): run()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): run()
    b.record(); torch.cuda.synchronize()
    t = a.elapsed_time(b) / 10
    ref = torch.einsum('mrk,mrn->kn', A.view(M, R, K).double(), B.view(M, R, N).double())
    print(f'M={M*R} K={K} N={N}: {t*1e3:.1f} us, {2*M*R*K*N/t/1e9:.1f} TFLOP/s, splits {ns}, err {(out.double()-ref).abs().max().item():.2e}')
