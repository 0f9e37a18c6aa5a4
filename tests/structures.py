"""Seeded synthetic structures."""
import numpy as np


def random_cell(n, species, density=0.05, seed=0, rmin=1.6, triclinic=True):
    rng = np.random.default_rng(seed)
    L = (n / density) ** (1 / 3)
    cell = np.eye(3) * L
    if triclinic:
        cell = cell + rng.normal(0, 0.08 * L, (3, 3)) * (1 - np.eye(3))
    pos = []
    while len(pos) < n:
        f = rng.random(3)
        p = f @ cell
        ok = True
        for q in pos:
            d = (p - q) @ np.linalg.inv(cell)
            d -= np.round(d)
            if np.linalg.norm(d @ cell) < rmin:
                ok = False
                break
        if ok:
            pos.append(p)
    numbers = rng.choice(species, n)
    return numbers, np.array(pos), cell


def diamond_si(reps, a=5.431, sigma=0.05, seed=0):
    base = np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0],
                     [.25, .25, .25], [.25, .75, .75], [.75, .25, .75], [.75, .75, .25]])
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(*[np.arange(r) for r in reps], indexing='ij'), -1).reshape(-1, 3)
    frac = (g[:, None, :] + base[None]).reshape(-1, 3)
    pos = frac * a + rng.normal(0, sigma, frac.shape)
    cell = np.diag(np.array(reps) * a)
    return np.full(len(pos), 14), pos, cell


def fluorite_hfo2(reps, a=4.85, sigma=0.05, seed=0, n_remove=0):
    hf = np.array([[0, 0, 0], [0, .5, .5], [.5, 0, .5], [.5, .5, 0]])
    ox = np.array([[x, y, z] for x in (.25, .75) for y in (.25, .75) for z in (.25, .75)])
    base = np.concatenate([hf, ox])
    z = np.array([72] * 4 + [8] * 8)
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(*[np.arange(r) for r in reps], indexing='ij'), -1).reshape(-1, 3)
    frac = (g[:, None, :] + base[None]).reshape(-1, 3)
    numbers = np.tile(z, len(g))
    if n_remove:
        keep = np.ones(len(frac), bool)
        keep[rng.choice(len(frac), n_remove, replace=False)] = False
        frac, numbers = frac[keep], numbers[keep]
    pos = frac * a + rng.normal(0, sigma, frac.shape)
    cell = np.diag(np.array(reps) * a)
    return numbers, pos, cell


def jittered_lips(n_side=32, density=0.048, jitter=0.4, seed=1):
    """BASELINE config 3 stand-in: n_side^3 atoms, Li:P:S = 3:1:4, disordered positions at `density` atoms/A^3 with
    a hard core (simple-cubic sites of spacing density^(-1/3) = 2.75 A, each coordinate displaced uniformly by up to
    +-`jitter` A, so no pair is closer than spacing - 2 jitter = 1.95 A).  O(n), unlike random_cell()."""
    rng = np.random.default_rng(seed)
    a = density ** (-1.0 / 3.0)
    g = np.stack(np.meshgrid(*[np.arange(n_side)] * 3, indexing='ij'), -1).reshape(-1, 3).astype(np.float64)
    pos = g * a + rng.uniform(-jitter, jitter, g.shape)
    numbers = rng.choice(np.array([3, 15, 16]), len(pos), p=[3 / 8, 1 / 8, 4 / 8])
    return numbers, pos, np.eye(3) * (n_side * a)
