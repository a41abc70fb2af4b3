"""Host-side SO(3) constant tables for the EF_layers path (float64, numpy only).

Everything here is a pure function of (lmax, mmax); nothing is read from disk.
The tables replace three constant sources of the reference:

* ``model/Jd.pt`` (the J matrices used by ``wigner_D``, EF_layers.py:2198-2217):
  recomputed by :func:`j_matrix` as the real-SH representation of the improper
  axis swap x <-> -y (polar axis is y), which is what the shipped file holds.
* ``e3nn.o3.ToS2Grid / FromS2Grid`` (EF_layers.py:562-587, e3nn 0.5.1, not
  vendored): restated by :func:`s2_grid_matrices`.
* ``CoefficientMappingModule`` index tables (EF_layers.py:1421-1549):
  :func:`coefficient_tables`.

Real spherical harmonics convention (e3nn): coefficient index i = l*l + l + m,
m = -l..l; m<0 carries sin(|m| alpha), m>0 carries cos(m alpha); polar axis y,
alpha = atan2(x, z), beta = acos(y); no Condon-Shortley phase.
"""
from __future__ import annotations

import functools
import math

import numpy as np


# ----------------------------------------------------------------------------
# real spherical harmonics
# ----------------------------------------------------------------------------
def legendre_table(lmax: int, z: np.ndarray, y: np.ndarray) -> np.ndarray:
    """Normalised associated Legendre values, shape [len(z), (lmax+1)^2].

    ``out[:, l*l+l+m] = sqrt((2l+1)/(4pi) (l-|m|)!/(l+|m|)!) P_l^{|m|}(z)`` with
    ``y = sqrt(1-z^2)`` and no Condon-Shortley phase (same value for +m and -m).
    """
    z = np.asarray(z, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    P = np.zeros((lmax + 1, lmax + 1) + z.shape, dtype=np.float64)  # P[l, m]
    for m in range(lmax + 1):
        dfact = 1.0
        for k in range(1, 2 * m, 2):
            dfact *= k
        P[m, m] = dfact * y ** m
        if m + 1 <= lmax:
            P[m + 1, m] = (2 * m + 1) * z * P[m, m]
        for l in range(m + 2, lmax + 1):
            P[l, m] = ((2 * l - 1) * z * P[l - 1, m] - (l + m - 1) * P[l - 2, m]) / (l - m)
    out = np.zeros(z.shape + ((lmax + 1) ** 2,), dtype=np.float64)
    for l in range(lmax + 1):
        for m in range(-l, l + 1):
            am = abs(m)
            norm = math.sqrt((2 * l + 1) / (4 * math.pi) * math.factorial(l - am) / math.factorial(l + am))
            out[..., l * l + l + m] = norm * P[l, am]
    return out


def alpha_table(lmax: int, alpha: np.ndarray) -> np.ndarray:
    """Azimuthal factors, shape [len(alpha), 2*lmax+1], column index lmax+m."""
    alpha = np.asarray(alpha, dtype=np.float64)
    out = np.ones(alpha.shape + (2 * lmax + 1,), dtype=np.float64)
    for m in range(1, lmax + 1):
        out[..., lmax - m] = math.sqrt(2.0) * np.sin(m * alpha)
        out[..., lmax + m] = math.sqrt(2.0) * np.cos(m * alpha)
    return out


def real_sph_harm(lmax: int, xyz: np.ndarray) -> np.ndarray:
    """Integral-normalised real SH of unit vectors, shape [n, (lmax+1)^2]."""
    xyz = np.asarray(xyz, dtype=np.float64)
    xyz = xyz / np.linalg.norm(xyz, axis=-1, keepdims=True)
    beta = np.arccos(np.clip(xyz[..., 1], -1.0, 1.0))
    alpha = np.arctan2(xyz[..., 0], xyz[..., 2])
    leg = legendre_table(lmax, np.cos(beta), np.abs(np.sin(beta)))
    sha = alpha_table(lmax, alpha)
    out = np.empty_like(leg)
    for l in range(lmax + 1):
        for m in range(-l, l + 1):
            out[..., l * l + l + m] = leg[..., l * l + l + m] * sha[..., lmax + m]
    return out


# ----------------------------------------------------------------------------
# J matrices (replacement for model/Jd.pt)
# ----------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def j_matrix(l: int) -> np.ndarray:
    """(2l+1)x(2l+1) float64 J_l with D^l(a,b,c) = Z(a) J Z(b) J Z(c).

    J_l represents the orthogonal map (x, y, z) -> (-y, -x, z) on degree-l real
    SH: Y_l(S r) = J_l Y_l(r).  Solved exactly from 4(2l+1) sample directions.
    """
    if l == 0:
        return np.ones((1, 1))
    S = np.array([[0.0, -1.0, 0.0], [-1.0, 0.0, 0.0], [0.0, 0.0, 1.0]])
    rng = np.random.RandomState(1234 + l)
    pts = rng.normal(size=(8 * (2 * l + 1), 3))
    pts /= np.linalg.norm(pts, axis=1, keepdims=True)
    sl = slice(l * l, (l + 1) ** 2)
    Y = real_sph_harm(l, pts)[:, sl]
    YS = real_sph_harm(l, pts @ S.T)[:, sl]
    # YS^T = J Y^T  ->  J = YS^T Y (Y^T Y)^-1
    J, *_ = np.linalg.lstsq(Y, YS, rcond=None)
    J = J.T
    J[np.abs(J) < 1e-13] = 0.0
    return J


# ----------------------------------------------------------------------------
# coefficient index tables (CoefficientMappingModule, EF_layers.py:1421-1549)
# ----------------------------------------------------------------------------
@functools.lru_cache(maxsize=None)
def j_structural_nonzero(l: int) -> np.ndarray:
    """Boolean [2l+1, 2l+1]: entries of ``J_l`` that can be non-zero.  With the class of order m defined as
    (s, q) = (m < 0, parity of l + |m|), J_l couples (1,1) and (0,0) to themselves and (1,0) with (0,1); every other
    entry is exactly zero.  csrc/wigner.cu (``wig_jnz``) skips the zero products at compile time."""
    m = np.arange(-l, l + 1)
    s, q = (m < 0).astype(int), (l + np.abs(m)) % 2
    same = (s[:, None] == s[None, :]) & (q[:, None] == q[None, :])
    swap = (s[:, None] != s[None, :]) & (q[:, None] != q[None, :])
    return np.where((s == q)[:, None], same, swap)


def coefficient_tables(lmax: int, mmax: int) -> dict:
    """Index tables for one resolution.

    keep      : l-primary indices (into (lmax+1)^2) of coefficients with |m|<=mmax
                (= ``coefficient_idx(lmax, mmax)``), length M.
    l_of/m_of : degree / signed order of each kept coefficient (l-primary order).
    to_m_perm : permutation p with ``x_m[j] = x_l[p[j]]`` (the one-hot ``to_m``
                matrix as an index vector), m-primary order = m=0 block, then for
                m=1..mmax: +m block ("real"), -m block ("imag").
    m_size    : number of degrees per order, [lmax+1-m].
    rescale   : per-degree rotate-back rescale sqrt((2l+1)/(2mmax+1)) for l>mmax.
    """
    keep, l_of, m_of = [], [], []
    for l in range(lmax + 1):
        mm = min(mmax, l)
        for m in range(-mm, mm + 1):
            keep.append(l * l + l + m)
            l_of.append(l)
            m_of.append(m)
    keep = np.array(keep, dtype=np.int64)
    l_of = np.array(l_of, dtype=np.int64)
    m_of = np.array(m_of, dtype=np.int64)
    perm = []
    m_size = []
    for m in range(mmax + 1):
        idx_r = [j for j in range(len(keep)) if m_of[j] == m]
        perm.extend(idx_r)
        m_size.append(len(idx_r))
        if m > 0:
            perm.extend([j for j in range(len(keep)) if m_of[j] == -m])
    rescale = np.ones(lmax + 1, dtype=np.float64)
    for l in range(lmax + 1):
        if l > mmax:
            rescale[l] = math.sqrt((2 * l + 1) / (2 * mmax + 1))
    return dict(keep=keep, l_of=l_of, m_of=m_of, to_m_perm=np.array(perm, dtype=np.int64),
                m_size=np.array(m_size, dtype=np.int64), rescale=rescale)


# ----------------------------------------------------------------------------
# S2 grid matrices (SO3_Grid, EF_layers.py:539-601)
# ----------------------------------------------------------------------------
def _quadrature_weights(b: int) -> np.ndarray:
    k = np.arange(b, dtype=np.float64)
    w = np.array([
        (2.0 / b) * math.sin(math.pi * (2.0 * j + 1.0) / (4.0 * b))
        * np.sum(np.sin((2 * j + 1) * (2 * k + 1) * math.pi / (4.0 * b)) / (2 * k + 1))
        for j in range(2 * b)
    ])
    return w / (2.0 * (2 * b) ** 2)


@functools.lru_cache(maxsize=None)
def s2_grid_matrices(lmax: int, mmax: int, normalization: str = "component", resolution=None):
    """(to_grid_mat, from_grid_mat), float64, each [res_beta, res_alpha, M].

    Restates ``SO3_Grid.__init__`` on top of e3nn 0.5.1's ToS2Grid/FromS2Grid:
    grid beta_i=(i+1/2)pi/res_beta, alpha_j=2 pi j/res_alpha; 'component'
    normalisation; rows with l>mmax rescaled by sqrt((2l+1)/(2mmax+1)) when
    lmax != mmax; columns restricted to |m|<=mmax.
    """
    res_b = 2 * (lmax + 1)
    res_a = 2 * (mmax + 1) + 1 if lmax == mmax else 2 * mmax + 1
    if resolution is not None:
        res_b = res_a = int(resolution)
    assert res_b % 2 == 0
    betas = (np.arange(res_b) + 0.5) / res_b * math.pi
    alphas = np.arange(res_a) / res_a * 2.0 * math.pi
    leg = legendre_table(lmax, np.cos(betas), np.abs(np.sin(betas)))  # [b, i]
    sha = alpha_table(lmax, alphas)  # [a, lmax+m]
    L = (lmax + 1) ** 2
    if normalization == "component":
        n_to = [math.sqrt(4 * math.pi) / math.sqrt(2 * l + 1) / math.sqrt(lmax + 1) for l in range(lmax + 1)]
        n_from = [math.sqrt(4 * math.pi) * math.sqrt(2 * l + 1) * math.sqrt(lmax + 1) for l in range(lmax + 1)]
    elif normalization == "norm":
        n_to = [math.sqrt(4 * math.pi) / math.sqrt(lmax + 1)] * (lmax + 1)
        n_from = [math.sqrt(4 * math.pi) * math.sqrt(lmax + 1)] * (lmax + 1)
    elif normalization == "integral":
        n_to = [1.0] * (lmax + 1)
        n_from = [4 * math.pi] * (lmax + 1)
    else:
        raise ValueError(normalization)
    qw = _quadrature_weights(res_b // 2) * res_b ** 2 / res_a  # [b]
    to_mat = np.zeros((res_b, res_a, L))
    from_mat = np.zeros((res_b, res_a, L))
    for l in range(lmax + 1):
        scale = math.sqrt((2 * l + 1) / (2 * mmax + 1)) if (lmax != mmax and l > mmax) else 1.0
        for m in range(-l, l + 1):
            i = l * l + l + m
            to_mat[:, :, i] = scale * n_to[l] * leg[:, i][:, None] * sha[:, lmax + m][None, :]
            from_mat[:, :, i] = scale * n_from[l] * (qw * leg[:, i])[:, None] * sha[:, lmax + m][None, :]
    keep = coefficient_tables(lmax, mmax)["keep"]
    return to_mat[:, :, keep].copy(), from_mat[:, :, keep].copy()
