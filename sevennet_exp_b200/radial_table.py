"""Tabulation of the radial weight function of one convolution layer.

``w(r) = FullyConnectedNet(BesselBasis(r) * cutoff(r))`` (sevenn/nn/convolution.py:92-101,121 with
sevenn/nn/edge_embedding.py:81-131 as its input) depends on the edge length only.  With the model weights
fixed (inference) it is evaluated once, in fp64, on the nodes ``r_i = i * h`` of a uniform grid over
``[0, rc]`` together with its exact derivative; the convolution kernels interpolate it per edge with cubic
Hermite polynomials (codegen/conv_codegen.py "Tabulated radial weights").  Interpolation error for the
SevenNet-0 shape at ``h = rc / 1000``: |dw| < 3e-9 (below the fp32 rounding of the table itself),
|d(dw/dr)| < 2e-6 on values of order 4 -- see tests/test_radial_table.py.

The XPLOR envelope has a jump in its second derivative at ``cutoff_on``; the grid is chosen so that this
point is a node (the Hermite pieces are then smooth inside every interval).

Weight preprocessing only: nothing here runs per evaluation.
"""
import math
from typing import Sequence, Tuple

import numpy as np

SMOOTH_ACTS = ('silu', 'tanh', 'sigmoid', 'ssp')


def _act(name: str, x: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """activation and its derivative (fp64)."""
    if name == 'silu':
        s = 1.0 / (1.0 + np.exp(-x))
        return x * s, s * (1.0 + x * (1.0 - s))
    if name == 'tanh':
        t = np.tanh(x)
        return t, 1.0 - t * t
    if name == 'sigmoid':
        s = 1.0 / (1.0 + np.exp(-x))
        return s, s * (1.0 - s)
    if name == 'ssp':
        return np.logaddexp(0.0, x) - math.log(2.0), 1.0 / (1.0 + np.exp(-x))
    raise ValueError(f'activation {name!r} is not smooth: the radial function cannot be tabulated')


def cutoff_and_derivative(r: np.ndarray, kind: str, rc: float, p: float, r_on: float):
    """Envelope value and d/dr: sevenn/nn/edge_embedding.py:134-181 (PolynomialCutoff, XPLORCutoff)."""
    if kind == 'poly_cut':
        x = r / rc
        c0, c1, c2 = (p + 1.0) * (p + 2.0) / 2.0, p * (p + 2.0), p * (p + 1.0) / 2.0
        f = 1.0 - c0 * x ** p + c1 * x ** (p + 1) - c2 * x ** (p + 2)
        df = (-c0 * p * x ** (p - 1) + c1 * (p + 1) * x ** p - c2 * (p + 2) * x ** (p + 1)) / rc
        return f, df
    den = (rc * rc - r_on * r_on) ** 3
    a, b = rc * rc - r * r, rc * rc + 2.0 * r * r - 3.0 * r_on * r_on
    f = np.where(r < r_on, 1.0, a * a * b / den)
    df = np.where(r < r_on, 0.0, (-4.0 * r * a * b + 4.0 * r * a * a) / den)
    return f, df


def radial_function(r: np.ndarray, coeffs: np.ndarray, kind: str, rc: float, p: float, r_on: float,
                    weights: Sequence[np.ndarray], act: str, act_cst: float):
    """(w, dw/dr) at the lengths ``r`` ([n]) -> two [n, wn] fp64 arrays.  ``weights`` already carry the
    1/sqrt(fan_in) factors (as in Engine.layers[...]['mlp'])."""
    r = np.asarray(r, dtype=np.float64)[:, None]
    c = np.asarray(coeffs, dtype=np.float64)[None, :]
    small = r < 1e-6
    rs = np.where(small, 1.0, r)
    cr = c * rs
    # sin(c r) / r and its derivative; the r -> 0 limits are c and 0
    b = np.where(small, c, np.sin(cr) / rs)
    db = np.where(small, 0.0, (c * np.cos(cr) - np.sin(cr) / rs) / rs)
    fc, dfc = cutoff_and_derivative(r, kind, rc, p, r_on)
    pref = 2.0 / rc
    cur, dcur = pref * b * fc, pref * (db * fc + b * dfc)
    for k, W in enumerate(weights):
        W = np.asarray(W, dtype=np.float64)
        pre, dpre = cur @ W, dcur @ W
        if k < len(weights) - 1:
            a, da = _act(act, pre)
            cur, dcur = act_cst * a, act_cst * da * dpre
        else:
            cur, dcur = pre, dpre
    return cur, dcur


def choose_intervals(rc: float, kind: str, r_on: float, n_min: int = 1000, n_max: int = 2048,
                     extra_nodes: Sequence[float] = ()) -> int:
    """Number of grid intervals over [0, rc]: the smallest n >= n_min for which ``cutoff_on`` and every radius in
    ``extra_nodes`` (per-layer cutoffs of multi-cutoff models) is (numerically) a node."""
    special = [float(x) for x in extra_nodes if 0.0 < float(x) < rc]
    if kind != 'poly_cut' and 0.0 < r_on < rc:
        special.append(float(r_on))
    if not special:
        return n_min
    best, best_err = n_min, 1.0
    for n in range(n_min, n_max + 1):
        err = max(abs(x * n / rc - round(x * n / rc)) for x in special)
        if err < 1e-9:
            return n
        if err < best_err:
            best, best_err = n, err
    return best


def build_table(coeffs, kind: str, rc: float, p: float, r_on: float, weights: Sequence[np.ndarray], act: str,
                act_cst: float, n_int: int, grid_rc: float = None, layout: str = 'hermite') -> np.ndarray:
    """fp32 [n_int, wn, 4]: (f_i, h f'_i, f_{i+1}, h f'_{i+1}) per interval and column.  ``grid_rc`` (>= rc, default
    rc) is the extent of the grid: a layer whose own cutoff ``rc`` is shorter than the neighbour-list cutoff sees no
    edge beyond ``rc`` (MultiCutoffEdgeEmbedding keeps r <= rc), i.e. its radial function is identically zero
    there; value and slope vanish at ``rc`` for both envelopes, so the zero continuation is C1."""
    grid_rc = rc if grid_rc is None else grid_rc
    h = grid_rc / n_int
    r = np.arange(n_int + 1, dtype=np.float64) * h
    inside = r <= rc * (1 + 1e-12)
    f, df = radial_function(np.where(inside, r, rc), coeffs, kind, rc, p, r_on, weights, act, act_cst)
    f, df = f * inside[:, None], df * inside[:, None]
    g = h * df
    if layout == 'poly':
        # the same cubic in the monomial basis of t = (r - r_i) / h, converted in fp64: w = c0 + t (c1 + t (c2 + t c3)),
        # h dw/dr = c1 + 2 t c2 + 3 t^2 c3 -- 3 (+2) FMAs per column in the kernels instead of 4 (+4) with the Hermite basis
        f0, g0, f1, g1 = f[:-1], g[:-1], f[1:], g[1:]
        tab = np.stack([f0, g0, 3.0 * (f1 - f0) - 2.0 * g0 - g1, 2.0 * (f0 - f1) + g0 + g1], axis=-1)
    else:
        tab = np.stack([f[:-1], g[:-1], f[1:], g[1:]], axis=-1)
    return np.ascontiguousarray(tab.astype(np.float32))


def hermite_eval(tab: np.ndarray, r: np.ndarray, rc: float, layout: str = 'hermite') -> Tuple[np.ndarray, np.ndarray]:
    """Host restatement of the kernels' interpolation (tests): (w, dw/dr) at lengths ``r`` in fp64."""
    n_int = tab.shape[0]
    inv_h = n_int / rc
    s = np.asarray(r, dtype=np.float64) * inv_h
    i = np.clip(s.astype(np.int64), 0, n_int - 1)
    t = (s - i)[:, None]
    T = tab[i].astype(np.float64)
    if layout == 'poly':
        w = T[..., 0] + t * (T[..., 1] + t * (T[..., 2] + t * T[..., 3]))
        return w, (T[..., 1] + 2 * t * T[..., 2] + 3 * t * t * T[..., 3]) * inv_h
    u = 1.0 - t
    h00, h10, h01, h11 = (1 + 2 * t) * u * u, t * u * u, t * t * (3 - 2 * t), -t * t * u
    w = T[..., 0] * h00 + T[..., 1] * h10 + T[..., 2] * h01 + T[..., 3] * h11
    d00, d10, d11 = -6 * t * u, u * (1 - 3 * t), t * (3 * t - 2)
    dw = (d00 * (T[..., 0] - T[..., 2]) + d10 * T[..., 1] + d11 * T[..., 3]) * inv_h
    return w, dw
