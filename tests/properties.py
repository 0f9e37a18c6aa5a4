"""Size-independent properties of an E / F / S evaluation (used at BASELINE's full sizes, where the oracle is too
slow): rigid rotation and translation, zero net force, energy <-> force and energy <-> stress consistency by central
finite differences.  ``compute(numbers, pos, cell) -> dict`` with the calculator's result keys."""
import numpy as np

KBAR = 1602.1766208


def random_rotation(seed):
    q, r = np.linalg.qr(np.random.default_rng(seed).normal(size=(3, 3)))
    q = q * np.sign(np.diag(r))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def stress_matrix(s6):
    """(xx, yy, zz, xy, yz, zx) -> symmetric 3x3."""
    xx, yy, zz, xy, yz, zx = s6
    return np.array([[xx, xy, zx], [xy, yy, yz], [zx, yz, zz]])


def check_properties(compute, numbers, pos, cell, tol_e=5e-6, tol_f=1e-4, tol_s_kbar=1e-2, fd_rel=2e-2, seed=0):
    n = len(numbers)
    base = compute(numbers, pos, cell)
    F, S0 = np.asarray(base['forces']), stress_matrix(base['inferred_stress'])
    fmax = max(1.0, np.abs(F).max())
    out = {}
    # rigid rotation of positions and cell (row vectors: x -> x R^T)
    R = random_rotation(seed)
    rot = compute(numbers, pos @ R.T, cell @ R.T)
    out['rot_dE_per_atom'] = abs(rot['energy'] - base['energy']) / n
    out['rot_dF'] = np.abs(np.asarray(rot['forces']) - F @ R.T).max()
    out['rot_dS_kbar'] = np.abs(stress_matrix(rot['inferred_stress']) - R @ S0 @ R.T).max() * KBAR
    assert rot['n_edges'] == base['n_edges']
    assert out['rot_dE_per_atom'] < tol_e, out
    assert out['rot_dF'] < tol_f * fmax, out
    assert out['rot_dS_kbar'] < tol_s_kbar * max(1.0, np.abs(S0).max() * KBAR * 1e-2), out
    # translation by an arbitrary vector (atoms leave the cell: the neighbour list wraps them)
    tr = compute(numbers, pos + np.array([1.234, -5.678, 9.1011]), cell)
    out['tr_dE_per_atom'] = abs(tr['energy'] - base['energy']) / n
    out['tr_dF'] = np.abs(np.asarray(tr['forces']) - F).max()
    assert tr['n_edges'] == base['n_edges']
    assert out['tr_dE_per_atom'] < tol_e and out['tr_dF'] < tol_f * fmax, out
    # periodic system: forces sum to zero
    out['net_force'] = np.abs(F.sum(0)).max()
    assert out['net_force'] < 2e-5 * n * fmax ** 0 + 2e-3, out
    # dE along the force direction: E(x + eps d) - E(x - eps d) = -2 eps sum F.d
    d = F / np.abs(F).max()
    eps = 2e-3
    ep, em = compute(numbers, pos + eps * d, cell)['energy'], compute(numbers, pos - eps * d, cell)['energy']
    pred = -2 * eps * float((F * d).sum())
    out['fd_force'] = (ep - em, pred)
    assert abs((ep - em) - pred) < fd_rel * abs(pred) + 5e-4, out
    # isotropic strain: dE/d(eps) = -V tr(inferred_stress): the model's `inferred_stress` is minus the ASE stress
    # (force_output.py:194-222; the calculator flips it, sevennet_calculator.py:196-201)
    h = 1e-3
    V = abs(np.linalg.det(cell))
    es = [compute(numbers, pos * (1 + s), cell * (1 + s))['energy'] for s in (h, -h)]
    pred = -2 * h * V * float(np.trace(S0))
    out['fd_stress'] = (es[0] - es[1], pred)
    assert abs((es[0] - es[1]) - pred) < fd_rel * abs(pred) + 2e-6 * n + 5e-4, out
    return out
