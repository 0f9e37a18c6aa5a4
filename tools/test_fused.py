"""GPU check of the fused radial-MLP-last-layer + convolution forward against (GEMM, unfused conv) on random data."""
import ctypes, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sevennet_exp_b200 import _lib as L, presets, spec as sp
from sevennet_exp_b200.engine import tc_weight_image
from sevennet_exp_b200.neighbor import NeighborList
from sevennet_exp_b200.codegen import fused_codegen as fc
from tests.structures import fluorite_hfo2

lib = L.load()
dev = torch.device('cuda')
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
reps = int(os.environ.get('REPS', '4'))
layer = int(os.environ.get('LAYER', '1'))
ms = sp.build_spec(presets.SEVENNET_0)
conv = ms.layers[layer].conv
sig = conv.signature()
num, pos, cell = fluorite_hfo2((reps, reps, reps), a=4.8)
r = NeighborList().build(pos, cell, [True] * 3, 5.0)
rowptr, col, E, n = r['rowptr'], r['col'], r['n_edges'], len(pos)
deg = (rowptr[1:] - rowptr[:-1]).max().item()
RT = max(1, min(8, 128 // deg))
print('atoms', n, 'edges', E, 'max deg', deg, 'RT', RT, 'sig', sig, 'wn', conv.wn)
torch.manual_seed(0)
h2 = torch.randn(E, 64, device=dev)
W2 = (torch.randn(64, conv.wn, device=dev) / 8.0)
x = torch.randn(n, conv.lx.dim, device=dev)
S = (conv.lmax_sh + 1) ** 2
Y = torch.randn(E, S, device=dev)
w = (h2.double() @ W2.double()).float().contiguous()
p = lambda t: ctypes.c_void_p(t.data_ptr())
fwd, _ = lib.conv(sig)
ref = torch.zeros(n, conv.lmid.dim, device=dev)
assert fwd(p(x), conv.lx.dim, p(Y), p(w), p(rowptr), p(col), n, 0.5, p(ref), conv.lmid.dim, st) == 0
NCH = fc.n_chunks(conv)
Wp = np.zeros((64, NCH * 64), np.float32); Wp[:, :conv.wn] = W2.cpu().numpy()
img = torch.from_numpy(tc_weight_image(Wp, 64, NCH, 2)).to(dev)
out = torch.zeros(n, conv.lmid.dim, device=dev)
wdbg = torch.zeros(E, NCH * 64, device=dev) if os.environ.get('WDBG', '1') == '1' else None
f = lib.fused_fwd(sig)
assert f is not None
rc = f(p(h2), p(img), p(x), conv.lx.dim, p(Y), p(rowptr), p(col), n, RT, 0.5, p(out), conv.lmid.dim,
       p(wdbg) if wdbg is not None else None, st)
torch.cuda.synchronize()
print('rc', rc)
if wdbg is not None:
    dw = (wdbg[:, :conv.wn] - w).abs().max().item()
    print('max |w_fused - w|', dw, 'scale', w.abs().max().item())
d = (out - ref).abs().max().item()
print('max |out_fused - out_ref|', d, 'scale', ref.abs().max().item())
ok = d < 2e-4 * max(1.0, ref.abs().max().item())
# timing
wdbg = None
for name, fn in (('fused', lambda: f(p(h2), p(img), p(x), conv.lx.dim, p(Y), p(rowptr), p(col), n, RT, 0.5, p(out), conv.lmid.dim, None, st)),
                 ('unfused conv only', lambda: fwd(p(x), conv.lx.dim, p(Y), p(w), p(rowptr), p(col), n, 0.5, p(ref), conv.lmid.dim, st))):
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    print(f'{name}: {min(ts) * 1e3:.1f} us')
print('PASS' if ok else 'FAIL')
