"""Constant tables: J matrices vs the shipped Jd.pt, S2 grid self-checks, Wigner-D equivariance."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import ef_oracle as eo
from oracle import ref_loader as rl
from singa_b200 import so3_constants as sc
from tests.parity_util import load_golden, oracle_tables


def test_j_matrices_are_symmetric_involutions():
    for l in range(9):
        J = sc.j_matrix(l)
        assert np.abs(J - J.T).max() < 1e-12
        assert np.abs(J @ J - np.eye(2 * l + 1)).max() < 1e-12


@pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference (build container only)")
def test_j_matrix_zero_pattern_used_by_the_wigner_kernel():
    """csrc/wigner.cu skips products with structurally zero J entries: the pattern must cover every non-zero."""
    for l in range(0, 12):
        J, nz = sc.j_matrix(l), sc.j_structural_nonzero(l)
        assert float(np.abs(J[~nz]).max() if (~nz).any() else 0.0) < 1e-13, l
        assert nz.sum() <= (2 * l + 1) ** 2 // 4 + l + 1


def test_j_matrices_match_shipped_jd():
    jd = torch.load(os.path.join(rl.REFERENCE_ROOT, "model", "Jd.pt"), weights_only=False)
    for l in range(len(jd)):
        assert np.abs(sc.j_matrix(l) - jd[l].numpy()).max() < 1e-12


def test_grid_matrices_match_golden_buffers():
    """Product-side grid constants equal the buffers the reference built (through the e3nn shim) bit-closely."""
    g = load_golden("example_3wi2.pt")["grid"]
    for (l, m), kt, kf in (((6, 2), "to62", "from62"), ((6, 6), "to66", "from66")):
        to, fr = sc.s2_grid_matrices(l, m)
        assert np.abs(to - g[kt].numpy()).max() < 2e-7
        assert np.abs(fr - g[kf].numpy()).max() < 2e-7


def test_grid_roundtrip_is_identity():
    to, fr = sc.s2_grid_matrices(6, 6)
    assert to.shape == (14, 15, 49)
    eye = np.einsum("bai,baj->ij", fr, to)
    assert np.abs(eye - np.eye(49)).max() < 1e-12
    to2, fr2 = sc.s2_grid_matrices(6, 2)
    assert to2.shape == (14, 5, 29)


def test_component_normalisation_gives_unit_variance():
    to, _ = sc.s2_grid_matrices(6, 6)
    assert np.abs((to ** 2).sum(-1) - 1.0).max() < 1e-12


def test_wigner_equivariance_of_real_sh():
    """Y(R r) = D(R) Y(r) with D from the oracle's wigner_dense and the frame convention of the path."""
    t = oracle_tables()
    rs = np.random.RandomState(0)
    vec = torch.from_numpy(rs.normal(size=(8, 3))).double()
    rot = eo.edge_rot_mat(vec, torch.from_numpy(rs.uniform(size=(8, 3))).double())
    D = eo.wigner_dense(rot, t)
    pts = rs.normal(size=(5, 3))
    pts /= np.linalg.norm(pts, axis=1, keepdims=True)
    for e in range(8):
        R = rot[e].numpy()
        Y = sc.real_sph_harm(6, pts)
        YR = sc.real_sph_harm(6, pts @ R.T)
        assert np.abs(YR - Y @ D[e].numpy().T).max() < 1e-10
        # the edge direction is mapped to the polar axis: only m = 0 survives
        ye = sc.real_sph_harm(6, vec[e:e + 1].numpy())[0]
        rotated = D[e].numpy() @ ye
        for l in range(7):
            blk = rotated[l * l:(l + 1) ** 2].copy()
            blk[l] = 0.0
            assert np.abs(blk).max() < 1e-10


def test_coefficient_tables():
    t = sc.coefficient_tables(6, 2)
    assert len(t["keep"]) == 29 and list(t["m_size"]) == [7, 6, 5]
    assert sorted(t["to_m_perm"].tolist()) == list(range(29))
    assert abs(t["rescale"][6] - math.sqrt(13 / 5)) < 1e-15 and t["rescale"][2] == 1.0


def test_real_spherical_harmonics_match_scipy():
    """Independent anchor for the e3nn-convention real harmonics (e3nn itself is not vendored: the grid constants are
    otherwise pinned only through the reference's buffers built on a shim).  scipy's complex Y_l^m (Condon-Shortley
    phase, polar angle from the y axis, azimuth atan2(x, z)) -> real form without the phase: m > 0 carries
    sqrt(2) (-1)^m Re Y_l^m, m < 0 carries sqrt(2) (-1)^m Im Y_l^|m|."""
    sp = pytest.importorskip("scipy.special")
    if not hasattr(sp, "sph_harm_y"):
        pytest.skip("scipy without sph_harm_y")
    rs = np.random.RandomState(0)
    xyz = rs.normal(size=(200, 3))
    xyz /= np.linalg.norm(xyz, axis=1, keepdims=True)
    beta, alpha = np.arccos(xyz[:, 1]), np.arctan2(xyz[:, 0], xyz[:, 2])
    ours = sc.real_sph_harm(8, xyz)
    for l in range(9):
        for m in range(-l, l + 1):
            Y = sp.sph_harm_y(l, abs(m), beta, alpha)
            ref = Y.real if m == 0 else np.sqrt(2.0) * (-1) ** m * (Y.real if m > 0 else Y.imag)
            assert np.abs(ref - ours[:, l * l + l + m]).max() < 1e-12, (l, m)
