"""Python binding of the C++ whole-model host (csrc/model.cu, `sevenn_model_*` in include/sevenn_b200.h).

``pack_engine`` serialises an :class:`engine.Engine` (tabulated-radial mode: block weights with the e3nn normalisation
folded in, their pre-swizzled tcgen05 images, radial tables, gate pieces, conv signatures) into the flat
(int32 program, fp32 data) pair that ``sevenn_model_create`` reads; ``NativeModel`` owns the handle, the workspace
and the result buffer and runs one E/F/S evaluation per call: no torch allocation, no Python per kernel, and no host
synchronisation before the result copy.  What it replaces in the reference: the per-step body of
``SevenNetCalculator.calculate`` (sevenn/sevennet_calculator.py:163-202) and of ``PairE3GNN::compute``
(sevenn/pair_e3gnn/pair_e3gnn.cpp:71-274)."""
import ctypes
import os
from typing import Optional

import numpy as np
import torch

from . import _lib
from .engine import Engine, tc_image_for

MAGIC, VERSION = 0x37454E42, 1
SENT_LIN, SENT_END = 0x4C494E, 0x454E44


class _Writer:
    def __init__(self):
        self.ints, self.chunks, self.n = [], [], 0

    def i(self, *vals):
        self.ints.extend(int(v) for v in vals)

    def f(self, v):
        self.ints.append(int(np.float32(v).view(np.int32)))

    def arr(self, a) -> int:
        """append a float array, 256-byte aligned; returns its offset in floats"""
        a = np.ascontiguousarray(np.asarray(a, dtype=np.float32)).reshape(-1)
        pad = (-self.n) % 64
        if pad:
            self.chunks.append(np.zeros(pad, dtype=np.float32))
            self.n += pad
        off = self.n
        self.chunks.append(a)
        self.n += a.size
        return off


def _host(t):
    return t.detach().cpu().numpy()


def _pack_lin(w: _Writer, lin, lib):
    sp = lin.spec
    w.i(lin.by_type, sp.lin.dim, sp.lout.dim, int(lin.fwd_full), int(lin.bwd_needs_zero), int(lin.bwd_overlap),
        len(sp.blocks))
    for b, W, WT, c in zip(sp.blocks, lin.W, lin.WT, lin.const):
        Wh, WTh = _host(W), _host(WT)
        w_off, wt_off = w.arr(Wh), w.arr(WTh)
        wi = wti = -1
        if not lin.by_type:
            if b.K % 4 == 0:
                wi = w.arr(tc_image_for(lib, Wh))
            if b.N % 4 == 0:
                wti = w.arr(tc_image_for(lib, WTh))
        cs = w.arr(_host(c)) if c is not None else -1
        w.i(b.l, sp.lin.group_off[b.g_in], sp.lout.group_off[b.g_out] + b.u_off_out, b.K, b.N, b.K,
            sp.lout.groups[b.g_out][0], w_off, wt_off, wi, wti, cs)
    w.i(SENT_LIN)


def supported(engine: Engine) -> Optional[str]:
    """None when the C++ host can run this engine, else the reason (the caller then keeps the Python launcher)."""
    if engine.radial != 'table':
        return "radial='mlp' (the C++ host runs the tabulated-radial kernels)"
    if engine.readout_fcn is not None:
        return 'readout_as_fcn'
    return None


def pack_engine(engine: Engine):
    """-> (prog int32 array, data float32 array) for sevenn_model_create."""
    why = supported(engine)
    if why:
        raise NotImplementedError(f'C++ host: {why}')
    sp, lib = engine.spec, engine.lib
    w = _Writer()
    w.i(MAGIC, VERSION, sp.lmax_sh, int(sp.normalize_sph), engine.tab_n, sp.n_species, engine.d0,
        int(sp.rescale_species_wise), engine.tc_min_rows)
    w.f(engine.tab_inv_h)
    w.f(sp.cutoff)
    w.i(w.arr(_host(engine.embed_table)), w.arr(_host(engine.scale)), w.arr(_host(engine.shift)))
    w.i(len(engine.layers))
    for lay in engine.layers:
        L = lay['spec']
        unused = any(not any(p.gx == gi for p in L.conv.paths) for gi in range(len(L.conv.lx.groups)))
        w.i(L.conv.lx.dim, L.conv.lmid.dim, L.gate.lin.dim, L.gate.lout.dim, lay['units'], int(unused))
        w.f(lay['inv_denom'])
        w.i(w.arr(_host(lay['tab'])))
        sig = lay['sig'].encode()
        w.i(len(sig), *sig)
        w.i(len(L.gate.pieces))
        for p in L.gate.pieces:
            w.i(p.kind, p.l, p.mul, p.in_off, p.in_stride, p.gate_off, p.out_off, p.out_stride, p.act)
            w.f(p.cst)
        for k in ('sc', 'si1', 'si2'):
            _pack_lin(w, lay[k], lib)
    _pack_lin(w, engine.read1, lib)
    _pack_lin(w, engine.read2, lib)
    w.i(SENT_END)
    return np.asarray(w.ints, dtype=np.int32), np.concatenate(w.chunks) if w.chunks else np.zeros(1, np.float32)


def save_model(path: str, engine: Engine):
    """model.bin for torch-free C++ hosts (csrc/md_host.cpp): int64 magic, n_prog, n_data | int32 program | fp32 data."""
    prog, data = pack_engine(engine)
    with open(path, 'wb') as f:
        np.asarray([MAGIC, prog.size, data.size], dtype=np.int64).tofile(f)
        prog.tofile(f)
        data.tofile(f)
    return path


class Outputs(ctypes.Structure):
    _fields_ = [('energy', ctypes.c_void_p), ('atomic_energy', ctypes.c_void_p), ('forces', ctypes.c_void_p),
                ('virial6', ctypes.c_void_p), ('virial_atoms', ctypes.c_void_p), ('status', ctypes.c_void_p)]


class GraphStruct(ctypes.Structure):
    _fields_ = [('n_local', ctypes.c_longlong), ('n_nodes', ctypes.c_longlong), ('max_edges', ctypes.c_longlong),
                ('n_edges_dev', ctypes.c_void_p), ('types', ctypes.c_void_p), ('edge_vec', ctypes.c_void_p),
                ('rowptr', ctypes.c_void_p), ('col', ctypes.c_void_p), ('cptr', ctypes.c_void_p),
                ('cedge', ctypes.c_void_p)]


HALO_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p)


class HaloHooks(ctypes.Structure):
    _fields_ = [('user', ctypes.c_void_p), ('forward', HALO_FN), ('backward', HALO_FN),
                ('n_interior', ctypes.c_longlong), ('side_stream', ctypes.c_void_p)]


class NativeModel:
    """One model on one GPU behind the C ABI.  ``evaluate`` returns device tensors that live in this object's result
    buffer (valid until the next call); ``evaluate_host`` adds the one device->host copy of E / virial / F / e_i."""

    HEADER = 64        # bytes: energy f64, virial f64 x 6, status i32 x 2

    def __init__(self, engine: Engine, slack: float = 1.05):
        self.lib, self.dev = engine.lib, engine.dev
        self.cutoff = float(engine.spec.cutoff)
        prog, data = pack_engine(engine)
        self._prog, self._data = prog, data
        h = ctypes.c_void_p()
        with torch.cuda.device(self.dev):
            self.lib.call('sevenn_model_create', prog.ctypes.data_as(ctypes.c_void_p), prog.size,
                          data.ctypes.data_as(ctypes.c_void_p), data.size, ctypes.byref(h))
        self.handle = h
        self.slack = slack
        self._ws = None
        self._res = None
        self._cap = {}             # (n_local, n_nodes) -> edge capacity
        self._hooks_keep = None
        self._side = None
        self._graphs, self._graph_seen = {}, None
        self._replayed = [0, 0]      # library calls / tcgen05 launches executed through graph replays
        self._stage = {}

    def __del__(self):
        try:
            if getattr(self, 'handle', None) is not None and self.handle.value:
                self.lib.cdll.sevenn_model_destroy(self.handle)
                self.handle = ctypes.c_void_p()
        except Exception:  # noqa: BLE001
            pass

    @property
    def n_launch(self):
        return int(self.lib.cdll.sevenn_model_launches(self.handle)) + self._replayed[0]

    @property
    def n_tc_launch(self):
        return int(self.lib.cdll.sevenn_model_tc_launches(self.handle)) + self._replayed[1]

    def set_tc_min_rows(self, rows: int):
        self.lib.call('sevenn_model_set_tc_min_rows', self.handle, int(rows))

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)

    def _workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=self.dev)
        return self._ws

    def _result(self, n_local, n_nodes):
        need = self.HEADER + 4 * (3 * n_nodes + n_local + 6 * n_nodes)
        if self._res is None or self._res.numel() < need:
            self._res = torch.empty(int(need * 1.25) + 256, dtype=torch.uint8, device=self.dev)
        return self._res

    def _views(self, res, n_local, n_nodes):
        hdr = res[:self.HEADER]
        o = self.HEADER
        F = res[o:o + 12 * n_nodes].view(torch.float32).view(n_nodes, 3)
        o += 12 * n_nodes
        ae = res[o:o + 4 * n_local].view(torch.float32)
        o += 4 * n_local
        vir = res[o:o + 24 * n_nodes].view(torch.float32).view(n_nodes, 6)
        return hdr, F, ae, vir

    def count_edges(self, pos_dev, cell, pbc, n_rows, pos_host=None) -> int:
        n = int(pos_dev.shape[0])
        ws = self._workspace(max(256 * n + (1 << 20), 1 << 20))
        out = ctypes.c_longlong()
        cell_a = np.ascontiguousarray(np.asarray(cell, dtype=np.float64).reshape(9))
        pbc_a = np.asarray([int(bool(b)) for b in np.asarray(pbc).reshape(3)], dtype=np.int32)
        self.lib.call('sevenn_model_count_edges', self.handle, int(n_rows), n, ctypes.c_void_p(pos_dev.data_ptr()),
                      pos_host.ctypes.data_as(ctypes.c_void_p) if pos_host is not None else None,
                      cell_a.ctypes.data_as(ctypes.c_void_p), pbc_a.ctypes.data_as(ctypes.c_void_p),
                      ctypes.c_void_p(ws.data_ptr()), ws.numel(), self._stream(), ctypes.byref(out))
        return int(out.value)

    def evaluate(self, types_dev, pos_dev, cell, pbc, pos_host=None, n_local=None, halo=None, want_forces=True):
        """types_dev int32 [n_nodes], pos_dev float64 [n_nodes, 3] (device).  Returns a dict of device tensors:
        header (energy f64 | virial f64 x 6 | status), forces [n_nodes, 3], atomic_energy [n_local],
        virial_per_atom [n_nodes, 6] -- all views of the result buffer."""
        n_nodes = int(pos_dev.shape[0])
        n_local = n_nodes if n_local is None else int(n_local)
        key = (n_local, n_nodes)
        if pos_host is not None:
            pos_host = np.ascontiguousarray(np.asarray(pos_host, dtype=np.float64))
        if key not in self._cap:
            e = self.count_edges(pos_dev, cell, pbc, n_local, pos_host)
            self._cap[key] = int(e * self.slack) + 1024
        cap = self._cap[key]
        nbytes = int(self.lib.cdll.sevenn_model_workspace_bytes(self.handle, n_local, n_nodes, cap, 1))
        if nbytes < 0:
            raise _lib.SevennError('sevenn_model_workspace_bytes failed')
        ws = self._workspace(nbytes)
        res = self._result(n_local, n_nodes)
        hdr, F, ae, vir = self._views(res, n_local, n_nodes)
        out = Outputs(hdr.data_ptr(), ae.data_ptr(), F.data_ptr() if want_forces else None, hdr.data_ptr() + 8,
                      vir.data_ptr(), hdr.data_ptr() + 56)
        cell_a = np.ascontiguousarray(np.asarray(cell, dtype=np.float64).reshape(9))
        pbc_a = np.asarray([int(bool(b)) for b in np.asarray(pbc).reshape(3)], dtype=np.int32)
        hooks = None
        if halo is not None:
            hooks = self._make_hooks(halo, n_nodes)
        self.lib.call('sevenn_model_evaluate_positions', self.handle, n_local, n_nodes,
                      ctypes.c_void_p(pos_dev.data_ptr()),
                      pos_host.ctypes.data_as(ctypes.c_void_p) if pos_host is not None else None,
                      cell_a.ctypes.data_as(ctypes.c_void_p), pbc_a.ctypes.data_as(ctypes.c_void_p),
                      ctypes.c_void_p(types_dev.data_ptr()), cap, ctypes.byref(hooks) if hooks is not None else None,
                      ctypes.c_void_p(ws.data_ptr()), ws.numel(), self._stream(), ctypes.byref(out))
        return {'header': hdr, 'forces': F, 'atomic_energy': ae, 'virial_per_atom': vir, 'key': key}

    def _make_hooks(self, halo, n_nodes):
        dev = self.dev

        def wrap(fn):
            def cb(user, layer, ptr, dim, stream):
                try:
                    t = _tensor_from_ptr(ptr, n_nodes * dim, dev).view(n_nodes, dim)
                    # the exchange runs on the stream the host names (the main one, or the side stream of the overlap)
                    if stream:
                        with torch.cuda.stream(torch.cuda.ExternalStream(int(stream), device=dev)):
                            fn(layer, t)
                    else:                      # NULL = the legacy default stream
                        with torch.cuda.stream(torch.cuda.default_stream(dev)):
                            fn(layer, t)
                    return 0
                except Exception as e:  # noqa: BLE001
                    import sys
                    print(f'[halo hook] {type(e).__name__}: {e}', file=sys.stderr, flush=True)
                    return 10003
            return HALO_FN(cb)
        # SURVEY 8e "Overlap": when the decomposition put the owned atoms without ghost neighbours first, the forward
        # exchange of a layer runs on a side stream under the convolution of those rows (SEVENN_HALO_OVERLAP=0: in line)
        n_int = int(getattr(getattr(halo, 'dec', None), 'n_interior', 0))
        side = None
        if n_int > 0 and os.environ.get('SEVENN_HALO_OVERLAP', '1') != '0':
            if self._side is None:
                self._side = torch.cuda.Stream(device=dev)
            side = self._side.cuda_stream
        hooks = HaloHooks(None, wrap(halo.forward), wrap(halo.backward), n_int if side else 0, side)
        self._hooks_keep = hooks
        return hooks

    # ---- CUDA-graph replay of a repeated step.  A step is launches + async memsets only (no allocation, no host
    # synchronisation), and every kernel argument is a function of (sizes, capacity, cell, pbc, buffer addresses): when
    # those repeat -- MD at constant cell with positions refreshed in place -- the second identical call captures the
    # step and later calls replay it (one driver call instead of ~100 launches).
    def evaluate_graphed(self, types_dev, pos_dev, cell, pbc, want_forces=True):
        periodic = all(bool(b) for b in np.asarray(pbc).reshape(3))
        n = int(pos_dev.shape[0])
        if not periodic or (n, n) not in self._cap or not want_forces:
            self._graph_seen = None
            return self.evaluate(types_dev, pos_dev, cell, pbc, want_forces=want_forces)
        cell_b = np.ascontiguousarray(np.asarray(cell, dtype=np.float64).reshape(9)).tobytes()
        key = (n, self._cap[(n, n)], cell_b, pos_dev.data_ptr(), types_dev.data_ptr(),
               self._ws.data_ptr() if self._ws is not None else 0, self._res.data_ptr() if self._res is not None else 0,
               torch.cuda.current_stream(self.dev).cuda_stream)
        hit = self._graphs.get(key)
        if hit is not None:
            hit[0].replay()
            self._replayed[0] += hit[2][0]
            self._replayed[1] += hit[2][1]
            return hit[1]
        if self._graph_seen != key:            # first sighting: run eagerly, capture if it comes again
            self._graph_seen = key
            out = self.evaluate(types_dev, pos_dev, cell, pbc)
            if self._ws.data_ptr() != key[5] or self._res.data_ptr() != key[6]:
                self._graph_seen = key[:5] + (self._ws.data_ptr(), self._res.data_ptr()) + key[7:]
            return out
        graph = torch.cuda.CUDAGraph()
        c0 = (self.n_launch, self.n_tc_launch)
        with torch.cuda.graph(graph):
            out = self.evaluate(types_dev, pos_dev, cell, pbc)
        per = (self.n_launch - c0[0], self.n_tc_launch - c0[1])       # library calls captured in one step
        graph.replay()
        if len(self._graphs) >= 8:
            self._graphs.pop(next(iter(self._graphs)))
        self._graphs[key] = (graph, out, per)
        return out

    @staticmethod
    def parse_header(hdr_host: np.ndarray):
        """bytes of the header on the host -> (energy, virial6, overflow flag, edges found)"""
        e = float(hdr_host[:8].view(np.float64)[0])
        v = hdr_host[8:56].view(np.float64).copy()
        st = hdr_host[56:64].view(np.int32)
        return e, v, bool(st[0] & 1), int(st[1])

    def _staging(self, n):
        st = self._stage.get(n)
        if st is None:
            if len(self._stage) >= 8:
                self._stage.pop(next(iter(self._stage)))
            st = {'pos_pin': torch.empty(n, 3, dtype=torch.float64).pin_memory(),
                  'types_pin': torch.empty(n, dtype=torch.int32).pin_memory(),
                  'pos_dev': torch.empty(n, 3, dtype=torch.float64, device=self.dev),
                  'types_dev': torch.empty(n, dtype=torch.int32, device=self.dev),
                  'res_pin': torch.empty(self.HEADER + 16 * n, dtype=torch.uint8).pin_memory()}
            self._stage[n] = st
        return st

    def evaluate_host(self, types, pos, cell, pbc):
        """Host arrays in (types int32 [n], pos float64 [n, 3]), E / virial / forces / atomic energies out on the host:
        pinned staging buffers, one host->device copy of positions + types, the step (replayed from a CUDA graph when the
        shapes and the cell repeat), one device->host copy; grows the edge capacity and repeats the step when the
        neighbour list did not fit."""
        n = int(len(types))
        st = self._staging(n)
        st['pos_pin'].numpy()[...] = pos
        st['types_pin'].numpy()[...] = types
        stream = torch.cuda.current_stream(self.dev)
        periodic = all(bool(b) for b in np.asarray(pbc).reshape(3))
        for _ in range(3):
            st['pos_dev'].copy_(st['pos_pin'], non_blocking=True)
            st['types_dev'].copy_(st['types_pin'], non_blocking=True)
            if periodic:
                out = self.evaluate_graphed(st['types_dev'], st['pos_dev'], cell, pbc)
            else:
                out = self.evaluate(st['types_dev'], st['pos_dev'], cell, pbc, pos_host=st['pos_pin'].numpy())
            st['res_pin'].copy_(self._res[:self.HEADER + 16 * n], non_blocking=True)
            stream.synchronize()
            host = st['res_pin'].numpy()
            e, v, overflow, found = self.parse_header(host[:self.HEADER])
            if not overflow:
                F = host[self.HEADER:self.HEADER + 12 * n].view(np.float32).reshape(n, 3)
                ae = host[self.HEADER + 12 * n:self.HEADER + 16 * n].view(np.float32)
                return {'energy': e, 'virial6': v, 'forces': F, 'atomic_energy': ae, 'n_edges': found}
            self._cap[out['key']] = int(found * self.slack) + 1024
            self._graphs.clear()
            self._graph_seen = None
        raise _lib.SevennError('neighbour list keeps overflowing the edge capacity')


class _PtrHolder:
    pass


def _tensor_from_ptr(ptr, numel, dev):
    """float32 tensor view of workspace memory handed to a halo hook (no ownership)."""
    iface = {'shape': (int(numel),), 'typestr': '<f4', 'data': (int(ptr), False), 'version': 2}
    h = _PtrHolder()
    h.__cuda_array_interface__ = iface
    return torch.as_tensor(h, device=dev)
