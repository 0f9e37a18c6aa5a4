"""sevenn_push_rows_peers (halo pack + peer write in one kernel) through the emulation build: rows must land in the
right peer buffer at the right row, for the indexed (forward) and the contiguous-block (reverse) form."""
import ctypes

import numpy as np

from tests.emu_util import build_full_emu


def test_push_rows_peers_emu():
    lib = build_full_emu([])
    rng = np.random.default_rng(0)
    D, n_src, world = 5, 40, 3
    src = rng.normal(size=(n_src, D)).astype(np.float32)
    counts = [4, 0, 7]                                   # rows for peer 0, 1, 2 (peer 1 gets nothing)
    seg = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    idx = rng.integers(0, n_src, size=int(seg[-1])).astype(np.int32)
    dst_row = [3, 0, 1]
    bufs = [np.full((16, D), -7.0, dtype=np.float32) for _ in range(world)]
    PP, LL, LL1 = ctypes.c_void_p * world, ctypes.c_longlong * world, ctypes.c_longlong * (world + 1)
    base = PP(*[b.ctypes.data for b in bufs])
    lib.call('sevenn_push_rows_peers', src.ctypes.data_as(ctypes.c_void_p), D,
             idx.ctypes.data_as(ctypes.c_void_p), 0, int(seg[-1]), world, base, LL(*dst_row),
             LL1(*[int(v) for v in seg]), None)
    for p in range(world):
        want = np.full((16, D), -7.0, dtype=np.float32)
        want[dst_row[p]:dst_row[p] + counts[p]] = src[idx[seg[p]:seg[p + 1]]]
        assert np.array_equal(bufs[p], want), p
    # reverse form: contiguous block src[first_row + i]
    bufs2 = [np.zeros((16, D), dtype=np.float32) for _ in range(world)]
    base2 = PP(*[b.ctypes.data for b in bufs2])
    first = 20
    lib.call('sevenn_push_rows_peers', src.ctypes.data_as(ctypes.c_void_p), D, None, first, int(seg[-1]), world,
             base2, LL(*dst_row), LL1(*[int(v) for v in seg]), None)
    for p in range(world):
        assert np.array_equal(bufs2[p][dst_row[p]:dst_row[p] + counts[p]], src[first + seg[p]:first + seg[p + 1]])
    # inconsistent segment table is refused
    bad = LL1(*[0, 4, 4, 10])
    assert lib.cdll.sevenn_push_rows_peers(src.ctypes.data_as(ctypes.c_void_p), D, None, 0, 11, world, base2,
                                          LL(*dst_row), bad, None) != 0
