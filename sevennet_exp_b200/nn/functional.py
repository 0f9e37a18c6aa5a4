"""torch.autograd.Function shims over the C ABI (include/sevenn_b200.h).

Each Function pairs a forward launcher with the hand-written reverse kernel of the same op, so that the
reference's own ``torch.autograd.grad(E, edge_vec)`` in ForceStressOutputFromEdge
(sevenn/nn/force_output.py:173-178) walks *our* kernels.  First order only: gradients flow to activations
(x, edge_vec), not to parameters -- this is the inference path (E/F/S); training is SURVEY 8f-1."""
import torch

from ..neighbor import csc_build
from ._backend import kernels, ptr


def _c(t):
    return t if t.is_contiguous() else t.contiguous()


class EdgeEmbedFn(torch.autograd.Function):
    """(edge_vec) -> (Y [E,S], emb [E,nb]);  backward = analytic d/d(edge_vec) (csrc/edge.cu)."""

    @staticmethod
    def forward(ctx, edge_vec, coeffs, args):
        k = kernels(edge_vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = args
        vec = _c(edge_vec.detach().float())
        E, S = vec.shape[0], (lmax + 1) ** 2
        Y, emb = k._empty(max(E, 1), S), k._empty(max(E, 1), nb)
        cf = _c(coeffs.detach().float())
        k.lib.call('sevenn_edge_fwd', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf),
                   ptr(Y), ptr(emb), k._stream())
        k.n_launch += 1
        ctx.save_for_backward(vec, cf)
        ctx.args = args
        return Y[:E], emb[:E]

    @staticmethod
    def backward(ctx, dY, demb):
        vec, cf = ctx.saved_tensors
        k = kernels(vec.device)
        lmax, normalize, nb, ckind, rc, p, r_on = ctx.args
        E, S = vec.shape[0], (lmax + 1) ** 2
        dY = _c(dY) if dY is not None else k._zeros(max(E, 1), S)
        demb = _c(demb) if demb is not None else k._zeros(max(E, 1), nb)
        f = k._empty(max(E, 1), 3)
        k.lib.call('sevenn_edge_bwd', ptr(vec), E, lmax, int(normalize), nb, ckind, rc, p, r_on, ptr(cf),
                   ptr(dY), 1, max(E, 1) * S, ptr(demb), ptr(f), k._stream())
        k.n_launch += 1
        return f[:E], None, None


class LinearFn(torch.autograd.Function):
    """Block-wise o3.Linear on device-layout rows; backward = the transposed blocks."""

    @staticmethod
    def forward(ctx, x, lin, types):
        k = kernels(x.device)
        x = _c(x.detach())
        n = x.shape[0]
        y = k._empty(n, lin.spec.lout.dim) if lin.fwd_full else k._zeros(n, lin.spec.lout.dim)
        if n:
            k._linear_fwd(lin, x, n, y, types=types)
        ctx.lin, ctx.types = lin, types
        return y

    @staticmethod
    def backward(ctx, dy):
        lin = ctx.lin
        k = kernels(dy.device)
        dy = _c(dy)
        n = dy.shape[0]
        dx = k._zeros(n, lin.spec.lin.dim) if lin.bwd_needs_zero else k._empty(n, lin.spec.lin.dim)
        if n:
            k._linear_bwd(lin, dy, n, dx, types=ctx.types)
        return dx, None, None


class RadialMLPFn(torch.autograd.Function):
    """e3nn FullyConnectedNet (sevenn/nn/convolution.py:92-101): emb [E,nb] -> per-edge weights [E,wn]."""

    @staticmethod
    def forward(ctx, emb, lay):
        k = kernels(emb.device)
        emb = _c(emb.detach())
        E = emb.shape[0]
        if E == 0:
            ctx.lay, ctx.pres, ctx.nb = lay, None, emb.shape[1]
            return k._empty(0, lay['mlp'][-1].shape[1])
        pres, _, w = k._mlp_fwd(lay, emb, E)
        ctx.lay, ctx.pres, ctx.nb = lay, pres, emb.shape[1]
        return w

    @staticmethod
    def backward(ctx, dW):
        k = kernels(dW.device)
        E = dW.shape[0]
        demb = k._zeros(max(E, 1), ctx.nb)
        if E:
            k._mlp_bwd(ctx.lay, _c(dW), ctx.pres, demb, E)
        return demb[:E], None


class ConvFn(torch.autograd.Function):
    """Fused gather -> uvu tensor product -> segment sum -> /denominator (generated kernels).
    Edges must be grouped by destination (CSR); ``graph`` = dict(rowptr, col, cptr, cedge, n_dst)."""

    @staticmethod
    def forward(ctx, h, Y, w, graph, meta):
        k = kernels(h.device)
        h, Y, w = _c(h.detach()), _c(Y.detach()), _c(w.detach())
        n_dst = graph['n_dst']
        agg = k._zeros(n_dst, meta['dmid'])
        fwd, _ = k.lib.conv(meta['sig'])
        rc = fwd(ptr(h), meta['dx'], ptr(Y), ptr(w), ptr(graph['rowptr']), ptr(graph['col']), n_dst,
                 meta['inv_denom'], ptr(agg), meta['dmid'], k._stream())
        if rc:
            raise RuntimeError(f'conv_fwd {meta["sig"]} failed with status {rc}')
        k.n_launch += 1
        ctx.save_for_backward(h, Y, w)
        ctx.graph, ctx.meta = graph, meta
        return agg

    @staticmethod
    def backward(ctx, dagg):
        h, Y, w = ctx.saved_tensors
        graph, meta = ctx.graph, ctx.meta
        k = kernels(h.device)
        E, S, nn_ = w.shape[0], Y.shape[1], h.shape[0]
        dW, dxe = k._empty(max(E, 1), meta['wn']), k._zeros(max(E, 1), meta['dx'])
        dY = k._zeros(meta['units'], max(E, 1), S)
        _, bwd = k.lib.conv(meta['sig'])
        rc = bwd(ptr(h), meta['dx'], ptr(Y), ptr(w), ptr(graph['rowptr']), ptr(graph['col']), graph['n_dst'],
                 meta['inv_denom'], ptr(_c(dagg)), meta['dmid'], ptr(dW), ptr(dxe), meta['dx'], ptr(dY),
                 max(E, 1) * S, k._stream())
        if rc:
            raise RuntimeError(f'conv_bwd {meta["sig"]} failed with status {rc}')
        dh = k._empty(nn_, meta['dx'])
        k.lib.call('sevenn_segment_gather_sum', ptr(dxe), meta['dx'], ptr(graph['cptr']), ptr(graph['cedge']),
                   nn_, ptr(dh), 0, k._stream())
        k.n_launch += 2
        return dh, dY.sum(0)[:E], dW[:E], None, None


class GateFn(torch.autograd.Function):
    """e3nn Gate (sevenn/nn/equivariant_gate.py:49-61) on device-layout rows."""

    @staticmethod
    def forward(ctx, pre, gate):
        k = kernels(pre.device)
        pre = _c(pre.detach())
        n = pre.shape[0]
        out = k._empty(n, gate['dout'])
        if n:
            k.lib.call('sevenn_gate_fwd', ptr(pre), gate['din'], ptr(out), gate['dout'], n, gate['pieces'],
                       gate['n_pieces'], k._stream())
            k.n_launch += 1
        ctx.save_for_backward(pre)
        ctx.gate = gate
        return out

    @staticmethod
    def backward(ctx, dout):
        (pre,), gate = ctx.saved_tensors, ctx.gate
        k = kernels(pre.device)
        n = pre.shape[0]
        dpre = k._empty(n, gate['din'])
        if n:
            k.lib.call('sevenn_gate_bwd', ptr(pre), gate['din'], ptr(_c(dout)), gate['dout'], ptr(dpre), n,
                       gate['pieces'], gate['n_pieces'], k._stream())
            k.n_launch += 1
        return dpre, None


def build_graph(edge_index, n_nodes, n_dst, device):
    """CSR by destination (edge_index[0]) + CSC by source for an arbitrary edge order.
    Returns (graph dict, perm or None): sorted edge k is input edge perm[k]."""
    k = kernels(device)
    dst, src = edge_index[0], edge_index[1]
    E = dst.shape[0]
    perm = None
    if E > 1 and bool((dst[1:] < dst[:-1]).any()):
        perm = torch.sort(dst, stable=True)[1]
        dst, src = dst[perm], src[perm]
    cnt = torch.bincount(dst, minlength=n_dst)[:n_dst]
    rowptr = torch.zeros(n_dst + 1, dtype=torch.int32, device=device)
    rowptr[1:] = torch.cumsum(cnt, 0).to(torch.int32)
    col = _c(src.to(torch.int32))
    cptr, cedge = csc_build(k.lib, col, n_nodes, torch.device(device), k._stream())
    return dict(rowptr=rowptr, col=col, cptr=cptr, cedge=cedge, n_dst=n_dst, n_nodes=n_nodes), perm
