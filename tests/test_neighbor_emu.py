"""Cell-list neighbour list (emulation build) vs the brute-force edge-set oracle: bit-exact sorted edge set."""
import numpy as np
import pytest

from oracle import sevenn_ref as S
from sevennet_exp_b200.neighbor import NeighborList
from tests.emu_util import build_full_emu
from tests.structures import diamond_si, random_cell


def _check(pos, cell, pbc, rc):
    nl = NeighborList(lib=build_full_emu([]), device='cpu')
    r = nl.build(pos, cell, pbc, rc)
    i, j, D, Sh = S.neighbor_list_bruteforce(pos, cell, pbc, rc)
    assert r['n_edges'] == len(i)
    assert np.array_equal(r['center'].numpy(), i)
    assert np.array_equal(r['col'].numpy(), j)
    assert np.array_equal(r['shift'].numpy(), Sh.astype(np.int64))
    assert np.array_equal(r['edge_vec'].numpy(), D.astype(np.float32))
    rp = r['rowptr'].numpy()
    assert rp[-1] == len(i) and np.all(np.diff(rp) == np.bincount(i, minlength=len(pos)))
    # CSC: edges grouped by neighbour, ascending edge id inside a segment
    cp, ce = r['cptr'].numpy(), r['cedge'].numpy()
    assert np.array_equal(np.sort(ce), np.arange(len(i)))
    for k in range(len(pos)):
        seg = ce[cp[k]:cp[k + 1]]
        assert np.all(j[seg] == k) and np.all(np.diff(seg) > 0)


def test_nl_triclinic_random():
    num, pos, cell = random_cell(60, [1, 2], density=0.05, seed=1)
    _check(pos, cell, [True] * 3, 4.0)


def test_nl_unwrapped_positions():
    num, pos, cell = random_cell(30, [1], density=0.04, seed=2)
    pos = pos + np.array([2.3, -1.7, 0.4]) @ cell * 1.0      # whole-cell translations, some atoms outside
    pos[::3] += cell[0] * 2
    _check(pos, cell, [True] * 3, 4.5)


def test_nl_small_cell_multiple_images():
    cell = np.array([[0, 2.815, 2.815], [2.815, 0, 2.815], [2.815, 2.815, 0]])
    pos = np.array([[0, 0, 0], [2.815, 2.815, 2.815]])
    _check(pos, cell, [True] * 3, 4.0)                          # NaCl primitive: 36 edges (test_data.py:48)
    _check(np.zeros((1, 3)), np.eye(3) * 3.35, [True] * 3, 4.0)
    _check(np.zeros((1, 3)), np.eye(3) * 1.7, [True] * 3, 5.0)


def test_nl_silicon_bulk():
    num, pos, cell = diamond_si((2, 2, 2))
    _check(pos, cell, [True] * 3, 5.0)


@pytest.mark.parametrize('pbc', [[False] * 3, [True, True, False], [False, True, False]])
def test_nl_mixed_pbc(pbc):
    num, pos, cell = random_cell(40, [1], density=0.05, seed=5, triclinic=False)
    _check(pos, cell, pbc, 4.0)


def test_nl_molecule_and_isolated():
    h2o = np.array([[0, 0, 0.119262], [0, 0.763239, -0.477047], [0, -0.763239, -0.477047]])
    _check(h2o, np.zeros((3, 3)), [False] * 3, 4.0)            # 6 edges (test_data.py:48)
    _check(np.zeros((1, 3)), np.zeros((3, 3)), [False] * 3, 4.0)   # 0 edges
