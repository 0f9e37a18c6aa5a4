"""Differentiable-to-second-order autograd Functions over the C ABI: the training path.

The reference trains through ``torch.autograd.grad(E, edge_vec, create_graph=True)`` (sevenn/nn/force_output.py:173-178)
followed by ``loss.backward()`` (sevenn/train/trainer.py:156-170): the loss contains forces and stress, so the backward
pass itself is differentiated.  Every op of the path is multilinear (o3.Linear, the uvu tensor-product convolution) or
elementwise (activations, gate, edge embedding), so each Function below has a backward written in terms of Functions of
the same closed family -- autograd can walk it again, and parameter gradients come out of the same hand-written kernels:

    BlockMM   y = x W (or x W^T) per irrep block      <->  BlockTN  G = a^T b (weight gradients, csrc/train.cu gemm_tn)
    ConvT     gather + uvu TP + segment sum (trilinear in h, Y, w)  <->  ConvBwd (dh, dY, dw), itself trilinear
    ActMul    m * cst * f^(k)(x)                       (radial MLP; k -> k + 1 under differentiation)
    GateFn -> GateBwd -> gate_bwd2 kernel
    EdgeFn -> EdgeBwd -> edge_train kernel (directional derivatives, Bessel-coefficient gradients)

First-order inference keeps using nn/functional.py (fused epilogues, no parameter gradients).
"""
import ctypes

import torch

from ._backend import kernels, ptr


def _c(t):
    """contiguous fp32 (differentiable torch ops: call it OUTSIDE Function.forward)"""
    t = t if t.dtype == torch.float32 else t.float()
    return t if t.is_contiguous() else t.contiguous()


def _chk(t):
    """Function inputs arrive contiguous fp32 (the wrappers / backward formulas see to it), so that
    ctx.save_for_backward keeps the ORIGINAL tensors and the second-order graph stays connected."""
    assert t.is_contiguous() and t.dtype == torch.float32, (t.dtype, t.is_contiguous())
    return t


class _DenseLin:
    """What Kernels._linear_fwd / _linear_bwd need, built from dense block tensors that live in the autograd graph."""

    def __init__(self, spec, Ws, by_type_S=0):
        self.spec, self.by_type = spec, by_type_S
        self.W = [_c(w.detach()) for w in Ws]
        self.WT = [_c(w.detach().transpose(-1, -2)) for w in Ws]
        for t in self.W + self.WT:
            t._sevenn_dynamic = True          # no host-built tcgen05 image for per-step weights (engine._gemm_args)
        self.const, self.has_const = [None] * len(Ws), False
        cover = [0] * len(spec.lin.groups)
        for b in spec.blocks:
            cover[b.g_in] += 1
        self.bwd_needs_zero = any(c != 1 for c in cover)
        self.bwd_overlap = any(c > 1 for c in cover)
        self.fwd_full = sum(b.N * (2 * b.l + 1) for b in spec.blocks) == spec.lout.dim


class BlockMM(torch.autograd.Function):
    """transposed=False: y[lout layout] = x[lin layout] @ W per block;  True: y[lin layout] = x[lout layout] @ W^T."""

    @staticmethod
    def forward(ctx, x, spec, transposed, types, by_S, *Ws):
        k = kernels(x.device)
        xd = _chk(x)
        n = xd.shape[0]
        lin = _DenseLin(spec, Ws, by_S)
        if not transposed:
            y = k._empty(n, spec.lout.dim) if lin.fwd_full else k._zeros(n, spec.lout.dim)
            if n:
                k._linear_fwd(lin, xd, n, y, types=types)
        else:
            y = k._zeros(n, spec.lin.dim) if lin.bwd_needs_zero else k._empty(n, spec.lin.dim)
            if n:
                k._linear_bwd(lin, xd, n, y, types=types)
        ctx.save_for_backward(x, *Ws)              # the inputs themselves: the saved tensors stay in the graph
        ctx.meta = (spec, transposed, types, by_S)
        return y

    @staticmethod
    def backward(ctx, dy):
        xd, *Ws = ctx.saved_tensors
        spec, transposed, types, by_S = ctx.meta
        dy = _c(dy)
        dx = BlockMM.apply(dy, spec, not transposed, types, by_S, *Ws) if ctx.needs_input_grad[0] else None
        dWs = [None] * len(Ws)
        if any(ctx.needs_input_grad[5:]):
            a, b = (xd, dy) if not transposed else (dy, xd)
            dWs = list(BlockTN.apply(a, b, spec, types, by_S))
        return (dx, None, None, None, None, *dWs)


class BlockTN(torch.autograd.Function):
    """G_b = sum_rows a_block^T b_block  (a in the lin layout, b in the lout layout): the weight gradients of BlockMM."""

    @staticmethod
    def forward(ctx, a, b, spec, types, by_S):
        k = kernels(a.device)
        _chk(a), _chk(b)
        n = a.shape[0]
        outs = []
        for blk in spec.blocks:
            R = 2 * blk.l + 1
            M = n * R
            a_off = spec.lin.group_off[blk.g_in]
            b_off = spec.lout.group_off[blk.g_out] + blk.u_off_out
            n_sp = int(k.lib.cdll.sevenn_gemm_tn_splits(M))
            part = k._empty(max(n_sp * blk.K * blk.N, 1))
            G = k._zeros(*((by_S, blk.K, blk.N) if by_S else (blk.K, blk.N)))
            for s in range(max(by_S, 1)):
                k.lib.call('sevenn_gemm_tn', ctypes.c_void_p(a.data_ptr() + 4 * a_off), spec.lin.dim, blk.K,
                           ctypes.c_void_p(b.data_ptr() + 4 * b_off), spec.lout.dim, spec.lout.groups[blk.g_out][0],
                           M, R, blk.K, blk.N, ptr(types) if by_S else None, s, 1.0, 0, ptr(part),
                           ctypes.c_void_p(G.data_ptr() + 4 * s * blk.K * blk.N), k._stream())
                k.n_launch += 2
            outs.append(G)
        ctx.save_for_backward(a, b)
        ctx.meta = (spec, types, by_S)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *Cs):
        a, b = ctx.saved_tensors
        spec, types, by_S = ctx.meta
        k = kernels(a.device)
        Cs = [_c(c) if c is not None else k._zeros(*((by_S, blk.K, blk.N) if by_S else (blk.K, blk.N)))
              for c, blk in zip(Cs, spec.blocks)]
        da = BlockMM.apply(b, spec, True, types, by_S, *Cs) if ctx.needs_input_grad[0] else None
        db = BlockMM.apply(a, spec, False, types, by_S, *Cs) if ctx.needs_input_grad[1] else None
        return da, db, None, None, None


class ActMul(torch.autograd.Function):
    """out = m * cst * f^(order)(x) (m may be None)."""

    @staticmethod
    def forward(ctx, x, m, act, cst, order):
        k = kernels(x.device)
        xd = _chk(x)
        md = _chk(m) if m is not None else None
        out = torch.empty_like(xd)
        if order > 2:
            raise NotImplementedError('third derivative of the activation: not needed by first-order force losses')
        k.lib.call('sevenn_act_mul', ptr(xd), ptr(md), xd.numel(), act, cst, order, ptr(out), k._stream())
        k.n_launch += 1
        ctx.save_for_backward(xd, md) if md is not None else ctx.save_for_backward(xd)
        ctx.meta = (act, cst, order, md is not None)
        return out

    @staticmethod
    def backward(ctx, c):
        act, cst, order, has_m = ctx.meta
        c = _c(c)
        xd = ctx.saved_tensors[0]
        md = ctx.saved_tensors[1] if has_m else None
        dm = ActMul.apply(xd, c, act, cst, order) if (has_m and ctx.needs_input_grad[1]) else None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ActMul.apply(xd, _c(c * md) if has_m else c, act, cst, order + 1)
        return dx, dm, None, None, None


class ConvT(torch.autograd.Function):
    """agg = segment_sum_over_edges(uvu(h[col], Y, w)) * inv_denom : trilinear in (h, Y, w)."""

    @staticmethod
    def forward(ctx, h, Y, w, graph, meta):
        k = kernels(h.device)
        hd, Yd, wd = _chk(h), _chk(Y), _chk(w)
        n_dst = graph['n_dst']
        agg = k._zeros(n_dst, meta['dmid'])
        fwd, _ = k.lib.conv(meta['sig'])
        rc = fwd(ptr(hd), meta['dx'], ptr(Yd), ptr(wd), ptr(graph['rowptr']), ptr(graph['col']), n_dst,
                 meta['inv_denom'], ptr(agg), meta['dmid'], k._stream())
        if rc:
            raise RuntimeError(f'conv_fwd {meta["sig"]} failed with status {rc}')
        k.n_launch += 1
        ctx.save_for_backward(hd, Yd, wd)
        ctx.graph, ctx.meta = graph, meta
        return agg

    @staticmethod
    def backward(ctx, g):
        hd, Yd, wd = ctx.saved_tensors
        dh, dY, dw = ConvBwd.apply(hd, Yd, wd, _c(g), ctx.graph, ctx.meta)
        return dh, dY, dw, None, None


class ConvBwd(torch.autograd.Function):
    """(dh, dY, dw) = d<g, ConvT(h, Y, w)>/d(h, Y, w): one launch of the generated reverse kernel + the source-side
    reduction.  Each output is trilinear in the remaining inputs and g, so its own backward is three forward and three
    reverse launches with swapped arguments."""

    @staticmethod
    def forward(ctx, h, Y, w, g, graph, meta):
        k = kernels(h.device)
        hd, Yd, wd, gd = _chk(h), _chk(Y), _chk(w), _chk(g)
        E, S, nn_ = wd.shape[0], Yd.shape[1], hd.shape[0]
        dW, dxe = k._empty(max(E, 1), meta['wn']), k._zeros(max(E, 1), meta['dx'])
        dY = k._zeros(meta['units'], max(E, 1), S)
        _, bwd = k.lib.conv(meta['sig'])
        rc = bwd(ptr(hd), meta['dx'], ptr(Yd), ptr(wd), ptr(graph['rowptr']), ptr(graph['col']), graph['n_dst'],
                 meta['inv_denom'], ptr(gd), meta['dmid'], ptr(dW), ptr(dxe), meta['dx'], ptr(dY), max(E, 1) * S,
                 k._stream())
        if rc:
            raise RuntimeError(f'conv_bwd {meta["sig"]} failed with status {rc}')
        dh = k._empty(nn_, meta['dx'])
        k.lib.call('sevenn_segment_gather_sum', ptr(dxe), meta['dx'], ptr(graph['cptr']), ptr(graph['cedge']), nn_,
                   ptr(dh), 0, k._stream())
        k.n_launch += 2
        ctx.save_for_backward(hd, Yd, wd, gd)
        ctx.graph, ctx.meta = graph, meta
        return dh, _c(dY.sum(0)[:E]), _c(dW[:E])

    @staticmethod
    def backward(ctx, ch, cY, cw):
        hd, Yd, wd, gd = ctx.saved_tensors
        graph, meta = ctx.graph, ctx.meta
        need_h, need_Y, need_w, need_g = ctx.needs_input_grad[:4]
        dg = dh = dY = dw = None

        def add(a, b):
            return b if a is None else (a if b is None else a + b)
        ch, cY, cw = (None if t is None else _c(t) for t in (ch, cY, cw))
        if ch is not None:
            if need_g:
                dg = add(dg, ConvT.apply(ch, Yd, wd, graph, meta))
            if need_Y or need_w:
                _, y1, w1 = ConvBwd.apply(ch, Yd, wd, gd, graph, meta)
                dY, dw = add(dY, y1 if need_Y else None), add(dw, w1 if need_w else None)
        if cY is not None:
            if need_g:
                dg = add(dg, ConvT.apply(hd, cY, wd, graph, meta))
            if need_h or need_w:
                h2, _, w2 = ConvBwd.apply(hd, cY, wd, gd, graph, meta)
                dh, dw = add(dh, h2 if need_h else None), add(dw, w2 if need_w else None)
        if cw is not None:
            if need_g:
                dg = add(dg, ConvT.apply(hd, Yd, cw, graph, meta))
            if need_h or need_Y:
                h3, y3, _ = ConvBwd.apply(hd, Yd, cw, gd, graph, meta)
                dh, dY = add(dh, h3 if need_h else None), add(dY, y3 if need_Y else None)
        return dh, dY, dw, dg, None, None


class GateFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pre, gate):
        k = kernels(pre.device)
        pd = _chk(pre)
        n = pd.shape[0]
        out = k._empty(n, gate['dout'])
        if n:
            k.lib.call('sevenn_gate_fwd', ptr(pd), gate['din'], ptr(out), gate['dout'], n, gate['pieces'],
                       gate['n_pieces'], k._stream())
            k.n_launch += 1
        ctx.save_for_backward(pd)
        ctx.gate = gate
        return out

    @staticmethod
    def backward(ctx, dout):
        return GateBwd.apply(ctx.saved_tensors[0], _c(dout), ctx.gate), None


class GateBwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pre, dout, gate):
        k = kernels(pre.device)
        pd, dd = _chk(pre), _chk(dout)
        n = pd.shape[0]
        dpre = k._empty(n, gate['din'])
        if n:
            k.lib.call('sevenn_gate_bwd', ptr(pd), gate['din'], ptr(dd), gate['dout'], ptr(dpre), n, gate['pieces'],
                       gate['n_pieces'], k._stream())
            k.n_launch += 1
        ctx.save_for_backward(pd, dd)
        ctx.gate = gate
        return dpre

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, c):
        pd, dd = ctx.saved_tensors
        gate = ctx.gate
        k = kernels(pd.device)
        n = pd.shape[0]
        d_dout, d_pre = k._zeros(n, gate['dout']), k._zeros(n, gate['din'])
        if n:
            k.lib.call('sevenn_gate_bwd2', ptr(pd), gate['din'], ptr(dd), gate['dout'], ptr(_c(c.float())), n,
                       gate['pieces'], gate['n_pieces'], ptr(d_dout), ptr(d_pre), k._stream())
            k.n_launch += 1
        return d_pre, d_dout, None


def _coeff_sum(k, G, E, nb):
    """deterministic fp64 column sums of the per-edge coefficient-gradient contributions"""
    part, out = k._empty(296 * 8, dtype=torch.float64), k._empty(nb, dtype=torch.float64)
    k.lib.call('sevenn_colsum_f64', ptr(G), E, nb, ptr(part), ptr(out), k._stream())
    k.n_launch += 2
    return out.float()


class EdgeFn(torch.autograd.Function):
    """(edge_vec, coeffs) -> (Y, emb); coeffs = trainable Bessel frequencies (edge_embedding.py:96-100)."""

    @staticmethod
    def forward(ctx, edge_vec, coeffs, args):
        k = kernels(edge_vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = args
        assert nb <= 8, 'the coefficient-gradient reduction handles up to 8 radial basis functions'
        vec, cf = _chk(edge_vec), _chk(coeffs)
        E, S = vec.shape[0], (lmax + 1) ** 2
        Y, emb = k._empty(max(E, 1), S), k._empty(max(E, 1), nb)
        k.lib.call('sevenn_edge_fwd', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf), ptr(Y),
                   ptr(emb), k._stream())
        k.n_launch += 1
        ctx.save_for_backward(vec, cf)
        ctx.args = args
        return _c(Y[:E]), _c(emb[:E])

    @staticmethod
    def backward(ctx, dY, demb):
        vec, cf = ctx.saved_tensors
        args = ctx.args
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = args
        E, S = vec.shape[0], (lmax + 1) ** 2
        dY = _c(dY) if dY is not None else k._zeros(max(E, 1), S)
        demb = _c(demb) if demb is not None else k._zeros(max(E, 1), nb)
        dvec = EdgeBwd.apply(vec, cf, dY, demb, args) if ctx.needs_input_grad[0] else None
        dcf = EdgeCoeffGrad.apply(vec, cf, demb, args) if ctx.needs_input_grad[1] else None
        return dvec, dcf, None


class EdgeCoeffGrad(torch.autograd.Function):
    """dL/d coeff_k = sum_e demb[e,k] d emb_k/d coeff_k (linear in demb; its derivative w.r.t. edge_vec is not needed:
    positions are not trained)."""

    @staticmethod
    def forward(ctx, vec, cf, demb, args):
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = args
        E = vec.shape[0]
        G = k._zeros(max(E, 1), nb)
        dd = _chk(demb)
        k.lib.call('sevenn_edge_train', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf), ptr(dd),
                   None, None, None, ptr(G), k._stream())
        k.n_launch += 1
        ctx.save_for_backward(vec, cf)
        ctx.args = args
        return _coeff_sum(k, G, E, nb)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, c):
        # d/d demb[e,k] = c_k d emb_k/d coeff_k
        vec, cf = ctx.saved_tensors
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = ctx.args
        E = vec.shape[0]
        G = k._zeros(max(E, 1), nb)
        k.lib.call('sevenn_edge_train', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf), None, None,
                   None, None, ptr(G), k._stream())
        k.n_launch += 1
        return None, None, G[:E] * c.reshape(1, nb), None


class EdgeBwd(torch.autograd.Function):
    """dvec = J_Y^T dY + J_emb^T demb (csrc/edge.cu reverse kernel).  Linear in (dY, demb): its backward returns the
    directional derivatives J_Y c, J_emb c and the mixed Bessel-coefficient term (csrc/train.cu edge_train)."""

    @staticmethod
    def forward(ctx, vec, cf, dY, demb, args):
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = args
        E, S = vec.shape[0], (lmax + 1) ** 2
        dYd, dd = _chk(dY), _chk(demb)
        f = k._empty(max(E, 1), 3)
        k.lib.call('sevenn_edge_bwd', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf), ptr(dYd), 1,
                   max(E, 1) * S, ptr(dd), ptr(f), k._stream())
        k.n_launch += 1
        ctx.save_for_backward(vec, cf, dd)
        ctx.args = args
        return _c(f[:E])

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, c):
        vec, cf, dd = ctx.saved_tensors
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = ctx.args
        E, S = vec.shape[0], (lmax + 1) ** 2
        JY, Jemb, G = k._zeros(max(E, 1), S), k._zeros(max(E, 1), nb), k._zeros(max(E, 1), nb)
        k.lib.call('sevenn_edge_train', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf), ptr(dd),
                   ptr(_c(c.float())), ptr(JY), ptr(Jemb), ptr(G), k._stream())
        k.n_launch += 1
        dcf = _coeff_sum(k, G, E, nb) if ctx.needs_input_grad[1] else None
        # no gradient w.r.t. edge_vec itself (second derivatives of the harmonics): positions are inputs, not parameters
        return None, dcf, JY[:E], Jemb[:E], None
