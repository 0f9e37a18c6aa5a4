"""Micro-benchmark of the per-atom stages of one SevenNet-0 inner layer (node-linear GEMMs through the tcgen05 arm, gate
forward / backward) at a given atom count: time, bytes touched once, GB/s.  usage: python tools/bench_node.py [atoms]"""
import ctypes
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_exp_b200 import _lib as L
from sevennet_exp_b200.engine import tc_image_for

lib = L.load()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 262144
flush = torch.empty(64 * 1024 * 1024, device='cuda')


def timed(fn, reps=6):
    ts = []
    for _ in range(reps):
        flush.fill_(0.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts[1:])


# (name, R, K, N, epi): the per-l blocks of si1 / sc / si2 / si2^T of a SevenNet-0 inner layer
shapes = [('si1 l0', 1, 128, 128, 0), ('si1 l1', 3, 64, 64, 0), ('si1 l2', 5, 32, 32, 0), ('sc  l0', 1, 128, 224, 0),
          ('si2 l0', 1, 224, 224, 1), ('si2 l1', 3, 384, 64, 1), ('si2 l2', 5, 352, 32, 1),
          ('si2T l0', 1, 224, 224, 0), ('si2T l1', 3, 64, 384, 0), ('si2T l2', 5, 32, 352, 0),
          ('si1T l1 acc', 3, 64, 64, 1), ('si1T l2 acc', 5, 32, 32, 1)]
sel = os.environ.get('NODE_ONLY')
tot = 0.0
for (name, R, K, N, epi) in shapes:
    if sel and sel not in name:
        continue
    A = torch.randn(n, R * K, device='cuda')
    B = torch.randn(K, N, device='cuda') / K ** 0.5
    C = torch.zeros(n, R * N, device='cuda')
    img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
    g = L.GemmArgs(A.data_ptr(), R * K, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, epi, 0, 1.679,
                   C.data_ptr(), C.data_ptr())
    t = timed(lambda: lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(img.data_ptr()), st))
    byt = 4 * n * R * (K + N * (2 if epi else 1))
    tot += t
    ref = (A.view(n, R, K)[:4096].double() @ B.double()).float().reshape(-1, R * N)
    err = float((C[:4096] / (6 if epi else 1) - ref).abs().max()) if not epi else -1.0
    print(f'{name:12s} M={n * R:8d} K={K:4d} N={N:4d} epi={epi} {t * 1e3:8.1f} us {byt / t / 1e6:7.0f} GB/s err {err:.1e}', flush=True)
    del A, B, C
print(f'gemm total {tot:.3f} ms')

# gate of the SevenNet-0 inner layer: [128 scalars | 96 gates | 64x1e | 32x2e] -> [128 | 64x1e | 32x2e]
c = 1.6791767923989418
pieces = [(0, 0, 128, 0, 0, 0, 0, 128, 0, c), (1, 1, 64, 224, 64, 128, 128, 64, 0, c), (1, 2, 32, 416, 32, 192, 320, 32, 0, c)]
pa = (L.GatePiece * 3)(*[L.GatePiece(*p) for p in pieces])
din, dout = 576, 480
pre = torch.randn(n, din, device='cuda'); out = torch.zeros(n, dout, device='cuda')
go = torch.randn(n, dout, device='cuda'); dpre = torch.zeros(n, din, device='cuda')
t = timed(lambda: lib.call('sevenn_gate_fwd', ctypes.c_void_p(pre.data_ptr()), din, ctypes.c_void_p(out.data_ptr()), dout, n, pa, 3, st))
print(f'gate_fwd {t * 1e3:8.1f} us {4 * n * (din + dout) / t / 1e6:7.0f} GB/s')
silu = torch.nn.functional.silu
ref = torch.cat([c * silu(pre[:, :128]), (c * silu(pre[:, 128:192]))[:, None, :].expand(n, 3, 64).reshape(n, 192) * pre[:, 224:416],
                 (c * silu(pre[:, 192:224]))[:, None, :].expand(n, 5, 32).reshape(n, 160) * pre[:, 416:576]], 1)
print('gate_fwd max err', float((out - ref).abs().max()))
t = timed(lambda: lib.call('sevenn_gate_bwd', ctypes.c_void_p(pre.data_ptr()), din, ctypes.c_void_p(go.data_ptr()), dout,
                           ctypes.c_void_p(dpre.data_ptr()), n, pa, 3, st))
print(f'gate_bwd {t * 1e3:8.1f} us {4 * n * (2 * din + dout) / t / 1e6:7.0f} GB/s')
pr = pre[:65536].clone().requires_grad_(True)
o = torch.cat([c * silu(pr[:, :128]), (c * silu(pr[:, 128:192]))[:, None, :].expand(-1, 3, 64).reshape(-1, 192) * pr[:, 224:416],
               (c * silu(pr[:, 192:224]))[:, None, :].expand(-1, 5, 32).reshape(-1, 160) * pr[:, 416:576]], 1)
o.backward(go[:65536])
print('gate_bwd max err', float((dpre[:65536] - pr.grad).abs().max()))
