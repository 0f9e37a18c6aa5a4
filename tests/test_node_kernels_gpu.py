"""Kernel-level GPU checks of the node stages rewritten late in round 2: tcgen05 GEMM plans (resident image, streamed
image, TMEM-A kernel, grouped launches) at sizes where every CTA walks many row tiles, the grouped exact-fp32 GEMM, and
the 16-byte gate kernels against the scalar ones.  Reference math: e3nn o3.Linear / nn.Gate as used by
sevenn/nn/linear.py:14-100 and sevenn/nn/equivariant_gate.py:13-61 (restated in torch fp64 here)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
CST = 1.6791767923989418


@pytest.fixture(scope='module')
def lib():
    from sevennet_exp_b200 import _lib as L
    return L.load()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ref(A, R, K, B):
    n = A.shape[0]
    return torch.einsum('nrk,kc->nrc', A[:, :R * K].reshape(n, R, K).double(), B.double()).reshape(n, -1)


# (nodes, R, K, N, epi): SevenNet-0 / l3i5 node-linear blocks; 40k+ rows so that each of the 148 CTAs takes several row
# tiles with both MMA-issuing warps and all accumulator buffers in rotation; ragged M (not a multiple of 128)
NODE_SHAPES = [(40001, 1, 128, 128, 0),    # image resident beside two A stages
               (40001, 1, 128, 224, 0),    # one column tile, image streamed by the loader warp
               (30011, 1, 224, 224, 1),    # streamed, accumulate epilogue, K not a multiple of 32... (224 = 7 x 32)
               (20003, 3, 64, 64, 0), (20003, 5, 32, 32, 1), (20003, 3, 64, 384, 0), (20003, 5, 32, 352, 0),
               (20003, 3, 384, 64, 1),     # TMEM-A kernel
               (20003, 5, 352, 32, 1), (9001, 7, 480, 32, 0), (9001, 1, 256, 256, 0), (9001, 1, 36, 20, 0)]


@pytest.mark.parametrize('shape', NODE_SHAPES)
def test_gemm_tc_node_shapes_many_tiles(lib, shape):
    from sevennet_exp_b200 import _lib as L
    from sevennet_exp_b200.engine import tc_image_for
    n, R, K, N, epi = shape
    torch.manual_seed(n + K)
    ld = R * K + 4
    A = torch.randn(n, ld, device='cuda')
    B = torch.randn(K, N, device='cuda') / K ** 0.5
    C = torch.randn(n, R * N + 4, device='cuda')
    C0 = C.clone()
    img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
    g = L.GemmArgs(A.data_ptr(), ld, K, B.data_ptr(), N, C.data_ptr(), R * N + 4, N, n * R, R, N, K, epi, 0, 1.0, None, None)
    lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(img.data_ptr()), _stream())
    torch.cuda.synchronize()
    ref = _ref(A, R, K, B) + (C0[:, :R * N].double() if epi else 0.0)
    err = (C[:, :R * N].double() - ref).abs().max().item()
    assert err < 2e-5 * max(1.0, K / 128), (shape, err)
    assert torch.equal(C[:, R * N:], C0[:, R * N:])          # padding columns of the output rows untouched


def test_gemm_tc_grouped_equals_single_launches(lib):
    """Grouped launch (small members share CTAs, large ones run alone) == one launch per problem, bit for bit."""
    from sevennet_exp_b200 import _lib as L
    from sevennet_exp_b200.engine import tc_image_for
    torch.manual_seed(5)
    for n in (700, 60001):
        probs = []
        for (R, K, N) in [(1, 224, 224), (3, 384, 64), (5, 352, 32), (1, 128, 128)]:
            A = torch.randn(n, R * K, device='cuda')
            B = torch.randn(K, N, device='cuda') / K ** 0.5
            img = torch.from_numpy(tc_image_for(lib, B.cpu().numpy())).cuda()
            probs.append((R, K, N, A, B, img))
        outs = []
        for grouped in (False, True):
            Cs = [torch.zeros(n, R * N, device='cuda') for (R, K, N, *_r) in probs]
            gs = [L.GemmArgs(A.data_ptr(), R * K, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, 0, 0, 1.0, None, None)
                  for (R, K, N, A, B, img), C in zip(probs, Cs)]
            if grouped:
                arr = (L.GemmArgs * len(gs))(*gs)
                imgs = (ctypes.c_void_p * len(gs))(*[p[5].data_ptr() for p in probs])
                lib.call('sevenn_gemm_tc_grouped', len(gs), arr, imgs, _stream())
            else:
                for g, p in zip(gs, probs):
                    lib.call('sevenn_gemm_tc', ctypes.byref(g), ctypes.c_void_p(p[5].data_ptr()), _stream())
            torch.cuda.synchronize()
            outs.append(Cs)
        for (R, K, N, A, B, img), a, b in zip(probs, *outs):
            assert (a.double() - _ref(A, R, K, B)).abs().max() < 2e-5 * max(1.0, K / 128)
            # the TMEM-A kernel (alone) and the shared-memory-A kernel (in a group) order the k-steps differently
            assert (a - b).abs().max() < 4e-5 * max(1.0, K / 128)


def test_gemm_f32_grouped_is_bit_identical(lib):
    from sevennet_exp_b200 import _lib as L
    torch.manual_seed(6)
    for n in (64, 5000):          # one shared launch / beyond SEVENN_GEMM_GROUP_MAX_TILES: library falls back to single launches
        probs = []
        for (R, K, N, epi) in [(1, 128, 224, 0), (3, 64, 64, 0), (5, 32, 32, 1), (1, 224, 224, 1), (3, 384, 64, 0), (7, 20, 12, 0)]:
            A = torch.randn(n, R * K + 3, device='cuda')
            B = torch.randn(K, N, device='cuda')
            probs.append((R, K, N, epi, A, B))
        outs = []
        for grouped in (False, True):
            torch.manual_seed(7)
            Cs = [torch.randn(n, R * N, device='cuda') for (R, K, N, *_r) in probs]
            gs = [L.GemmArgs(A.data_ptr(), R * K + 3, K, B.data_ptr(), N, C.data_ptr(), R * N, N, n * R, R, N, K, epi, 0, 1.0, None, None)
                  for (R, K, N, epi, A, B), C in zip(probs, Cs)]
            if grouped:
                lib.call('sevenn_gemm_f32_grouped', len(gs), (L.GemmArgs * len(gs))(*gs), _stream())
            else:
                for g in gs:
                    lib.call('sevenn_gemm_f32', ctypes.byref(g), _stream())
            torch.cuda.synchronize()
            outs.append(Cs)
        for a, b in zip(*outs):
            assert torch.equal(a, b)
    too_many = (L.GemmArgs * 9)()
    assert lib.cdll.sevenn_gemm_f32_grouped(9, too_many, None) != 0


def _gate_pieces(L, shift):
    """SevenNet-0 inner gate: [128 scalars | 64 + 32 gates | 64x1e | 32x2e] -> [128 | 64x1e | 32x2e]; `shift` moves every
    offset off the 16-byte grid (the library must then take the scalar kernels)."""
    s = shift
    pieces = [(0, 0, 128, s, 0, 0, s, 128, 0, CST), (1, 1, 64, 224 + s, 64, 128 + s, 128 + s, 64, 0, CST),
              (1, 2, 32, 416 + s, 32, 192 + s, 320 + s, 32, 0, CST)]
    return (L.GatePiece * 3)(*[L.GatePiece(*p) for p in pieces])


@pytest.mark.parametrize('n', [1, 37, 4096, 70001])
def test_gate_float4_kernels_equal_scalar_kernels_and_torch(lib, n):
    from sevennet_exp_b200 import _lib as L
    torch.manual_seed(n)
    silu = torch.nn.functional.silu
    res = {}
    for shift in (0, 1):
        if shift == 0:
            din, dout = 576, 480
            pre, go = torch.randn(n, din, device='cuda'), torch.randn(n, dout, device='cuda')
            res['pre'], res['go'] = pre, go
        else:                                   # same values one column in, rows padded by four floats
            din, dout = 580, 484
            pre, go = torch.zeros(n, din, device='cuda'), torch.zeros(n, dout, device='cuda')
            pre[:, 1:577] = res['pre']
            go[:, 1:481] = res['go']
        out = torch.zeros(n, dout, device='cuda')
        dpre = torch.zeros(n, din, device='cuda')
        pa = _gate_pieces(L, shift)
        lib.call('sevenn_gate_fwd', ctypes.c_void_p(pre.data_ptr()), din, ctypes.c_void_p(out.data_ptr()), dout, n, pa, 3, _stream())
        lib.call('sevenn_gate_bwd', ctypes.c_void_p(pre.data_ptr()), din, ctypes.c_void_p(go.data_ptr()), dout,
                 ctypes.c_void_p(dpre.data_ptr()), n, pa, 3, _stream())
        torch.cuda.synchronize()
        res[shift] = (out[:, shift:shift + 480].clone(), dpre[:, shift:shift + 576].clone())
    # float4 kernels (shift 0) against the scalar kernels (shift 1): forward identical, backward up to the order of the
    # 2l+1 products in the gate-channel sum
    assert torch.equal(res[0][0], res[1][0])
    assert (res[0][1] - res[1][1]).abs().max() < 2e-5
    p = res['pre'].double().requires_grad_(True)
    o = torch.cat([CST * silu(p[:, :128]),
                   (CST * silu(p[:, 128:192]))[:, None, :].expand(n, 3, 64).reshape(n, 192) * p[:, 224:416],
                   (CST * silu(p[:, 192:224]))[:, None, :].expand(n, 5, 32).reshape(n, 160) * p[:, 416:576]], 1)
    o.backward(res['go'].double())
    assert (res[0][0].double() - o.detach()).abs().max() < 1e-5
    assert (res[0][1].double() - p.grad).abs().max() < 2e-5
