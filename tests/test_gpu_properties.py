"""Size-independent properties of the CUDA path at BASELINE.json config-2 shapes (full 400-500 atom complexes,
8 A radius graphs, all four passes), where the CPU oracle would take minutes:
  * SO(3) equivariance: rotating every atom position by Q maps each output block of degree l by the Wigner matrix
    D^l(Q) (l = 0 rows invariant); the random gauge of the edge frames must not matter;
  * batch independence: a complex gives the same embedding alone and inside a batch of 32 (disjoint union);
  * gradient linearity: doubling the upstream gradient doubles every parameter gradient (bit for bit up to
    flush-to-zero of subnormals).
Tolerances relative to ||ref||_inf: equivariance 1e-2 (fp32 mode; the S2 grid is only approximately equivariant, see
the test) / 5e-2 (bf16 mode), l=0 invariance 1e-3 / 5e-2, batch independence 1e-6, linearity 1e-6."""
import numpy as np
import pytest
import torch

from tests.parity_util import LIGAND, PROTEIN, build_model, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rotation(seed):
    q, r = np.linalg.qr(np.random.RandomState(seed).normal(size=(3, 3)))
    q = q * np.sign(np.diag(r))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return torch.from_numpy(q.astype(np.float32))


def _run(model, g):
    from singa_b200 import synthetic as syn
    out = model(syn.GraphView({k: v.to(DEV) for k, v in g.items()}))
    return {k: out[k].embedding for k in (PROTEIN, LIGAND)}


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-2), ("bf16", 5e-2)])
def test_rotating_the_complex_rotates_the_embedding(precision, tol):
    """The S2-grid activations sample the sphere on a finite grid, so the architecture itself is equivariant only to
    ~3e-3 (measured in fp32 mode; the reference shares this); l = 0 rows are invariant to rounding."""
    from singa_b200 import EF_layers as ef
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    try:
        model = build_model(DEV, 4, precision)
        g = syn.make_complex(21)  # U{400..500} protein atoms, 8 A graph: ~28 k edges
        Q = _rotation(3)
        g2 = dict(g)
        g2["pos_p"] = g["pos_p"] @ Q.t()
        g2["pos_l"] = g["pos_l"] @ Q.t()
        with torch.no_grad():
            torch.manual_seed(1)
            a = _run(model, g)
            torch.manual_seed(2)      # different random gauge of the edge frames as well
            b = _run(model, g2)
            rot = ef.SO3_Rotation(6, device=DEV)
            D = rot.RotationToWignerDMatrix(Q.to(DEV).view(1, 3, 3), 0, 6)[0]   # [49, 49] block diagonal
    finally:
        ops.set_precision("fp32")
    for k in a:
        tol0 = 1e-3 if precision == "fp32" else tol
        assert rel_err(b[k][:, 0], a[k][:, 0]) < tol0, ("l=0 invariance", k, rel_err(b[k][:, 0], a[k][:, 0]))
        want = torch.einsum("ij,njc->nic", D, a[k])
        alt = torch.einsum("ji,njc->nic", D, a[k])
        e = min(rel_err(b[k], want), rel_err(b[k], alt))
        assert e < tol, ("equivariance", k, rel_err(b[k], want), rel_err(b[k], alt))
        # the higher degrees are really populated (the check above is not vacuous)
        assert float(a[k][:, 1:].abs().max()) > 1e-3 * float(a[k].abs().max())


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-6), ("bf16", 1e-6)])
def test_complex_alone_equals_complex_in_batch_of_32(precision, tol):
    """Disjoint union batching (GAN.py:48-55): no information crosses complexes; the kernels are deterministic and
    row-independent, so the values agree to rounding of the summation order only."""
    import bench
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    try:
        model = build_model(DEV, 4, precision)
        batch, graphs = bench.make_batch(0, 32, None)
        from singa_b200 import EF_layers as ef
        torch.manual_seed(7)
        frames_b, per = {}, [dict() for _ in graphs]
        for key, (sp, dp, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                      lp=("pos_l", "pos_p", "ei_lp")).items():
            e = batch[ei]
            frames_b[key] = ef.init_edge_rot_mat((batch[sp][e[0]] - batch[dp][e[1]]).to(DEV), device=DEV)
            off = 0
            for i, gr in enumerate(graphs):  # collate() concatenates the edge lists complex by complex
                n = gr[ei].shape[1]
                per[i][key] = frames_b[key][off:off + n]
                off += n
        with torch.no_grad():
            big = model(syn.GraphView({k: v.to(DEV) for k, v in batch.items()}), frames=frames_b)
            ptr_p, ptr_l = batch["ptr_p"].tolist(), batch["ptr_l"].tolist()
            for i in (0, 13, 31):
                one = model(syn.GraphView({k: v.to(DEV) for k, v in graphs[i].items()}),
                            frames=per[i])
                ep = rel_err(big[PROTEIN].embedding[ptr_p[i]:ptr_p[i + 1]], one[PROTEIN].embedding)
                el = rel_err(big[LIGAND].embedding[ptr_l[i]:ptr_l[i + 1]], one[LIGAND].embedding)
                assert ep < tol and el < tol, (i, ep, el)
    finally:
        ops.set_precision("fp32")


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_gradients_are_linear_in_the_upstream_gradient(precision):
    import bench
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    try:
        model = build_model(DEV, 4, precision)
        batch, _ = bench.make_batch(0, 4, None)
        g = {k: v.to(DEV) for k, v in batch.items()}
        grads = []
        for scale in (1.0, 2.0):
            for p in model.parameters():
                p.grad = None
            torch.manual_seed(5)
            out = model(syn.GraphView(g))
            loss = scale * (out[PROTEIN].embedding.sum() + out[LIGAND].embedding.sum())
            loss.backward()
            grads.append({n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None})
    finally:
        ops.set_precision("fp32")
    assert grads[0].keys() == grads[1].keys() and len(grads[0]) > 50
    # scaling by 2 commutes with every rounding, so the gradients are bit-identical except where flush-to-zero of
    # subnormal intermediates (fast-math kernels) differs between the two scales: allow 1e-6 of the tensor's max
    worst = {n: rel_err(grads[1][n], 2.0 * grads[0][n]) for n in grads[0]}
    bad = {n: e for n, e in worst.items() if e > 1e-6}
    assert not bad, bad
    exact = sum(torch.equal(grads[1][n], 2.0 * grads[0][n]) for n in grads[0])
    assert exact >= len(grads[0]) // 2, (exact, len(grads[0]))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_training_step_is_bitwise_reproducible(precision):
    """No float atomics anywhere on the step (segment kernels, split-K planes, per-CTA partials and the embedding
    gradient all reduce in a fixed order): two identical forward+backward passes give identical bits."""
    import bench
    from singa_b200 import ops
    from singa_b200 import synthetic as syn
    try:
        model = build_model(DEV, 4, precision)
        batch, _ = bench.make_batch(0, 4, None)
        g = {k: v.to(DEV) for k, v in batch.items()}
        runs = []
        for _ in range(2):
            for p in model.parameters():
                p.grad = None
            torch.manual_seed(5)
            out = model(syn.GraphView(g))
            (out[PROTEIN].embedding.sum() + out[LIGAND].embedding.sum()).backward()
            runs.append(({k: out[k].embedding.detach().clone() for k in (PROTEIN, LIGAND)},
                         {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}))
    finally:
        ops.set_precision("fp32")
    for k in runs[0][0]:
        assert torch.equal(runs[0][0][k], runs[1][0][k]), k
    for n in runs[0][1]:
        assert torch.equal(runs[0][1][n], runs[1][1][n]), n
