"""Spatial domain decomposition with per-layer ghost-atom feature exchange (one process per GPU).

Mirrors LAMMPS ``pair_style e3gnn/parallel`` (sevenn/pair_e3gnn/pair_e3gnn_parallel.cpp:191-525) and the
SevenNet patch of ``CommBrick`` (sevenn/pair_e3gnn/comm_brick.cpp:1057-1123):

  * the periodic cell is cut into a brick grid (px, py, pz) in fractional coordinates; a rank owns the
    atoms inside its brick and keeps ghost copies of every atom image within the cutoff of the brick;
  * node order is [owned | ghosts grouped by owner rank], so received rows land contiguously behind the
    owned rows and the convolution reads x[0 : n_local + n_ghost] without a copy
    (reference: torch.cat(x, x_ghost), sevenn/nn/convolution.py:123-124);
  * before every convolution t >= 1 the owners send the rows their peers need (forward), and in the
    reverse pass ghost gradients travel back and are added at the owner (pair_e3gnn_parallel.cpp:872-897);
    layer 0 needs no exchange because ghost embeddings follow from ghost types (model_build.py:319-357);
  * instead of CommBrick's six sequential relay swaps each rank talks to each peer directly (NVSwitch gives
    every pair full bandwidth): one grouped isend/irecv per exchange via torch.distributed (NCCL on GPUs,
    gloo in the CPU tests).  Pack = gather kernel into one send buffer, unpack-add = scatter-add kernel.
"""
import ctypes
import os
import sys
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from .engine import Graph


def brick_grid(world: int):
    """LAMMPS-style processor grid for a roughly cubic cell: 1, 2x1x1, 2x2x1, 2x2x2, ..."""
    g, d = [1, 1, 1], 0
    w = world
    while w > 1:
        f = 2 if w % 2 == 0 else w
        g[d % 3] *= f
        w //= f
        d += 1
    return tuple(g)


class Decomposition:
    """Host-side (numpy) split of a global periodic structure; every rank runs it on the same input."""

    def __init__(self, numbers, pos, cell, cutoff, grid, rank):
        numbers, pos = np.asarray(numbers), np.asarray(pos, dtype=np.float64)
        cell = np.asarray(cell, dtype=np.float64).reshape(3, 3)
        inv = np.linalg.inv(cell)
        frac = pos @ inv
        wrap = np.floor(frac)
        frac = frac - wrap
        frac[frac >= 1.0] -= 1.0
        posw = frac @ cell
        grid = np.asarray(grid)
        coord = np.minimum((frac * grid).astype(np.int64), grid - 1)
        owner = (coord[:, 0] * grid[1] + coord[:, 1]) * grid[2] + coord[:, 2]
        me = np.array([rank // (grid[1] * grid[2]), (rank // grid[2]) % grid[1], rank % grid[2]])
        lo, hi = me / grid, (me + 1) / grid
        h = 1.0 / np.linalg.norm(inv, axis=0)            # plane spacings
        halo = cutoff / h + 1e-9                          # fractional halo width per direction
        if np.any(halo * grid > 1.0):
            raise ValueError('sub-domain thinner than the cutoff (comm_brick.cpp:1074 aborts likewise)')
        self.local_ids = np.nonzero(owner == rank)[0]
        # owned atoms farther than the cutoff from every face of the brick have no ghost neighbour: they come first, so a
        # host can convolve rows [0, n_interior) while the ghost rows are still in flight (SURVEY 8e "Overlap")
        fl = frac[self.local_ids]
        interior = np.all((fl >= lo + halo) & (fl < hi - halo), axis=1)
        self.local_ids = np.concatenate([self.local_ids[interior], self.local_ids[~interior]])
        self.n_interior = int(interior.sum())
        # ghost candidates: every atom, every image shift in {-1,0,1}^3 whose position falls in the
        # brick grown by the halo, except the owned atoms themselves at shift 0
        g_ids, g_shift = [], []
        for sx in (-1, 0, 1):
            for sy in (-1, 0, 1):
                for sz in (-1, 0, 1):
                    s = np.array([sx, sy, sz], dtype=np.float64)
                    f = frac + s
                    m = np.all((f >= lo - halo) & (f < hi + halo), axis=1)
                    if sx == 0 and sy == 0 and sz == 0:
                        m &= owner != rank
                    idx = np.nonzero(m)[0]
                    g_ids.append(idx)
                    g_shift.append(np.tile(s, (len(idx), 1)))
        g_ids = np.concatenate(g_ids)
        g_shift = np.concatenate(g_shift)
        order = np.lexsort((g_shift[:, 2], g_shift[:, 1], g_shift[:, 0], g_ids, owner[g_ids]))
        self.ghost_ids, self.ghost_shift = g_ids[order], g_shift[order]
        self.ghost_owner = owner[self.ghost_ids]
        self.n_local, self.n_ghost = len(self.local_ids), len(self.ghost_ids)
        self.numbers = np.concatenate([numbers[self.local_ids], numbers[self.ghost_ids]])
        self.pos = np.concatenate([posw[self.local_ids], posw[self.ghost_ids] + self.ghost_shift @ cell])
        self.cell, self.owner, self.rank, self.grid = cell, owner, rank, tuple(int(x) for x in grid)
        self.volume = abs(float(np.linalg.det(cell)))
        self.n_global = len(numbers)


class Halo:
    """Exchange plan + the forward / backward / force-fold collectives of one rank."""

    def __init__(self, dec: Decomposition, lib, device, group=None):
        self.dec, self.lib, self.dev, self.group = dec, lib, torch.device(device), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dec.rank
        # what I need from each owner: global ids, in my ghost order (ghosts are grouped by owner)
        need = [dec.ghost_ids[dec.ghost_owner == o] for o in range(self.world)]
        self.recv_count = [len(n) for n in need]
        if self.world > 1:
            gathered: List[Optional[list]] = [None] * self.world
            dist.all_gather_object(gathered, [n.tolist() for n in need], group=group)
        else:
            gathered = [[n.tolist() for n in need]]
        # what each peer needs from me -> my local row indices
        g2l = -np.ones(dec.n_global, dtype=np.int64)
        g2l[dec.local_ids] = np.arange(dec.n_local)
        send = [g2l[np.asarray(gathered[p][self.rank], dtype=np.int64)] for p in range(self.world)]
        assert all((s >= 0).all() for s in send)
        self.send_count = [len(s) for s in send]
        self.send_idx = torch.tensor(np.concatenate(send) if sum(self.send_count) else np.zeros(0, np.int64),
                                     dtype=torch.int32, device=self.dev)
        self.send_off = np.concatenate([[0], np.cumsum(self.send_count)]).astype(np.int64)
        self.recv_off = np.concatenate([[0], np.cumsum(self.recv_count)]).astype(np.int64)
        self.bytes_fwd = 0
        self.peer = None
        mode = os.environ.get('SEVENN_HALO', 'auto')
        if mode in ('p2p', 'auto') and self.world > 1 and self.dev.type == 'cuda':
            try:
                self._setup_peer(group)
            except Exception as e:                      # no peer access between these GPUs: NCCL carries the rows
                if mode == 'p2p':
                    raise
                print(f'[halo rank {self.rank}] peer-memory halo unavailable ({e}); using NCCL', file=sys.stderr, flush=True)
                self.peer = None
            # every rank must use the same transport: one failed rendezvous sends all of them to NCCL
            ok = torch.tensor([1 if self.peer is not None else 0], dtype=torch.int32, device=self.dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                self.peer = None

    # ---- peer-memory halo (NVLink / NVSwitch): every rank owns a symmetric staging area of two slots; a sender
    # packs its rows straight into the receiver's slot (one kernel for all peers), a device-side barrier over the
    # signal pads orders the exchange, the receiver copies / scatter-adds out of its own slot.  Slots alternate per
    # exchange: a peer can only write slot s for exchange k + 2 after it passed the barrier of exchange k + 1, which
    # this rank enters (in stream order) after it has consumed slot s of exchange k.
    def _setup_peer(self, group):
        import torch.distributed._symmetric_memory as symm_mem
        if self.world > 16:
            raise RuntimeError('more than SEVENN_MAX_PEERS ranks')
        grp = group if group is not None else dist.group.WORLD
        rows = max(int(self.send_off[-1]), int(self.recv_off[-1]), 1)
        cap = torch.tensor([rows], dtype=torch.int64, device=self.dev)
        dist.all_reduce(cap, op=dist.ReduceOp.MAX, group=group)
        self._slot_rows = int(cap.item())
        self._slot_dim = int(os.environ.get('SEVENN_HALO_MAX_DIM', '1024'))
        self._slot_floats = self._slot_rows * self._slot_dim
        tables: List[Optional[list]] = [None] * self.world
        dist.all_gather_object(tables, (self.send_off.tolist(), self.recv_off.tolist()), group=group)
        # where my rows start inside peer p's slot: forward = p's receive offset for me, reverse = p's send offset
        ll = ctypes.c_longlong * self.world
        self._fwd_dst_row = ll(*[int(tables[p][1][self.rank]) for p in range(self.world)])
        self._bwd_dst_row = ll(*[int(tables[p][0][self.rank]) for p in range(self.world)])
        buf = symm_mem.empty(2 * self._slot_floats, dtype=torch.float32, device=self.dev)
        hdl = symm_mem.rendezvous(buf, grp)
        self._stage = buf
        self._peer_ptrs = [int(q) for q in hdl.buffer_ptrs]
        self._n_exchange = 0
        llp1, pp = ctypes.c_longlong * (self.world + 1), ctypes.c_void_p * self.world
        self._seg_fwd = llp1(*[int(v) for v in self.send_off])
        self._seg_bwd = llp1(*[int(v) for v in self.recv_off])
        self._slot_base = [pp(*[q + 4 * slot * self._slot_floats for q in self._peer_ptrs]) for slot in (0, 1)]
        self.peer = hdl
        hdl.barrier(channel=0, timeout_ms=20000)

    def _peer_push(self, src, D, idx, first_row, n, dst_row, seg):
        if D > self._slot_dim:
            raise ValueError(f'halo rows of {D} floats exceed SEVENN_HALO_MAX_DIM={self._slot_dim}')
        slot = self._n_exchange & 1
        self._n_exchange += 1
        self.lib.call('sevenn_push_rows_peers', ctypes.c_void_p(src.data_ptr()), D,
                      ctypes.c_void_p(idx.data_ptr()) if idx is not None else None, first_row, n, self.world,
                      self._slot_base[slot], dst_row, seg, self._stream())
        self.peer.barrier(channel=0, timeout_ms=20000)
        return self._stage[slot * self._slot_floats:]

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream) if self.dev.type == 'cuda' else None

    def _exchange(self, send_buf, send_counts, send_off, recv_buf, recv_counts, recv_off):
        """rows send_buf[send_off[p] : +send_counts[p]] -> peer p ; recv_buf[recv_off[p] : ...] <- peer p."""
        me = self.rank
        if self.world > 1 and send_buf.is_cuda:
            # NCCL: ONE all-to-all-v per exchange (rows are already grouped by peer on both sides), instead of a
            # batch of 2 (world - 1) point-to-point operations: the host cost per exchange drops several-fold, and at
            # ~5 ms per step with nine exchanges the host, not the link, is what the small-brick case waits for
            ns, nr = int(send_off[-1]), int(recv_off[-1])
            dist.all_to_all_single(recv_buf[:nr], send_buf[:ns], [int(c) for c in recv_counts],
                                   [int(c) for c in send_counts], group=self.group)
            return
        if send_counts[me]:
            recv_buf[recv_off[me]:recv_off[me] + recv_counts[me]].copy_(
                send_buf[send_off[me]:send_off[me] + send_counts[me]])
        if self.world == 1:
            return
        ops = []
        for p in range(self.world):
            if p == me:
                continue
            if recv_counts[p]:
                ops.append(dist.P2POp(dist.irecv, recv_buf[recv_off[p]:recv_off[p] + recv_counts[p]], p, self.group))
            if send_counts[p]:
                ops.append(dist.P2POp(dist.isend, send_buf[send_off[p]:send_off[p] + send_counts[p]], p, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    def forward(self, t, h):
        """Fill ghost rows h[n_local:] with the owners' rows (after self_interaction_1 of layer t)."""
        D = h.shape[1]
        nl = self.dec.n_local
        ns = int(self.send_off[-1])
        if self.peer is not None:
            nr = int(self.recv_off[-1])
            mine = self._peer_push(h, D, self.send_idx, 0, ns, self._fwd_dst_row, self._seg_fwd)
            if nr:
                h[nl:nl + nr].copy_(mine[:nr * D].view(nr, D))
            self.bytes_fwd += ns * D * 4
            return
        buf = torch.empty(max(ns, 1), D, dtype=h.dtype, device=self.dev)
        if ns:
            self.lib.call('sevenn_gather_rows', ctypes.c_void_p(h.data_ptr()), D,
                          ctypes.c_void_p(self.send_idx.data_ptr()), ns, ctypes.c_void_p(buf.data_ptr()),
                          self._stream())
        self._exchange(buf, self.send_count, self.send_off, h[nl:], self.recv_count, self.recv_off)
        self.bytes_fwd += ns * D * 4

    def backward(self, t, dh):
        """Send ghost gradient rows dh[n_local:] to their owners and add them into dh[:n_local]."""
        D = dh.shape[1]
        nl = self.dec.n_local
        ns = int(self.send_off[-1])
        if self.peer is not None:
            # ghost rows are grouped by owner: segment p = dh[nl + recv_off[p] : nl + recv_off[p + 1]]
            mine = self._peer_push(dh, D, None, nl, int(self.recv_off[-1]), self._bwd_dst_row, self._seg_bwd)
            if ns:
                self.lib.call('sevenn_scatter_add_rows', ctypes.c_void_p(mine.data_ptr()), D,
                              ctypes.c_void_p(self.send_idx.data_ptr()), ns, ctypes.c_void_p(dh.data_ptr()),
                              self._stream())
            return
        buf = torch.empty(max(ns, 1), D, dtype=dh.dtype, device=self.dev)
        ghost = dh[nl:].contiguous()
        self._exchange(ghost, self.recv_count, self.recv_off, buf, self.send_count, self.send_off)
        if ns:
            self.lib.call('sevenn_scatter_add_rows', ctypes.c_void_p(buf.data_ptr()), D,
                          ctypes.c_void_p(self.send_idx.data_ptr()), ns, ctypes.c_void_p(dh.data_ptr()),
                          self._stream())

    def fold_forces(self, F):
        """Forces accumulated on ghost copies belong to their owners (LAMMPS newton reverse comm,
        pair_e3gnn_parallel.cpp:472-477,669-671)."""
        self.backward(-1, F)
        return F[:self.dec.n_local]

    def all_reduce_(self, t):
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return t


class _NativeStepInfo:
    """What callers read off the graph of a step (edge count) when the C++ host built it on the device."""

    def __init__(self, native, header):
        self._native, self._header = native, header

    @property
    def n_edges(self):
        return self._native.parse_header(self._header.cpu().numpy())[3]


def evaluate_decomposed(calc, dec: Decomposition, halo: Halo, types_dev=None, pos_dev=None, host='auto'):
    """One E/F/S evaluation of this rank's sub-domain; returns (E_total, forces of owned atoms, virial6, graph info)
    with energy and virial already summed over ranks.  host='auto': the C++ whole-model host (one C call per step, the
    ghost exchanges plugged in through `sevenn_halo_hooks`) when the calculator has one, else the per-launch Python
    driver; host='python' forces the latter (stage profiling, emulation tests)."""
    dev = calc.device
    if types_dev is None:
        types_dev = torch.from_numpy(calc._lut[dec.numbers].astype(np.int32)).to(dev)
    # owned + ghost atoms as an open (non-periodic) point set: rows are needed for owned atoms only,
    # and those come first, so the CSR is simply truncated
    # rows for the owned atoms only (they come first); ghosts are neighbours, never centres
    nl = dec.n_local
    native = getattr(calc, 'native', None)
    if native is not None and host != 'python':
        if pos_dev is None:
            pos_dev = torch.from_numpy(np.ascontiguousarray(dec.pos, dtype=np.float64)).to(dev)
        out = native.evaluate(types_dev, pos_dev, np.eye(3), [False] * 3, pos_host=dec.pos, n_local=nl,
                              halo=halo if halo.world > 1 else None)
        F = halo.fold_forces(out['forces'])
        tot = out['header'][:56].view(torch.float64).clone()          # energy, virial6 (fp64 sums of this rank)
        halo.all_reduce_(tot)
        return tot[0], F, tot[1:], _NativeStepInfo(native, out['header'])
    r = calc.nl.build(dec.pos, np.eye(3), [False] * 3, calc.cutoff, pos_dev=pos_dev, n_rows=nl)
    g = Graph(types_dev, r['edge_vec'], r['rowptr'], r['col'], r['cptr'], r['cedge'], nl, dec.volume)
    out = calc.engine.evaluate(g, halo=halo)
    F = halo.fold_forces(out['forces'])
    tot = torch.cat([out['energy'].reshape(1), out['virial6']])
    halo.all_reduce_(tot)
    return tot[0], F, tot[1:], g
