"""GPU parity tests proper: the CUDA path (through the C ABI) against the reference-generated goldens and the
CPU oracle.  fp32 mode tolerances: outputs <= 1e-5 relative (||d||_inf/||ref||_inf), gradients <= 1e-4
(SURVEY.md 8c); bf16 tensor-core mode: stated looser bounds, 2e-2 / 5e-2."""
import numpy as np
import pytest
import torch

from tests.parity_util import (LIGAND, PROTEIN, build_model, digest_err, load_golden, oracle_tables, rel_err,
                               run_attention_case)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_OUT_FP32, TOL_GRAD_FP32 = 1e-5, 1e-4


def test_native_library_is_loaded():
    from singa_b200 import _lib
    assert _lib.LIB.efl_version() >= 100
    assert torch.cuda.get_device_capability(0)[0] == 10, "built for sm_100a"


@pytest.mark.parametrize("lmax,key", [(6, "rows_l6_m2"), (8, "rows_l8_m2")])
def test_wigner_kernel_matches_reference_rows(lmax, key):
    from singa_b200 import ops
    fx = load_golden("wigner_small.pt")
    wig = ops.wigner_from_rotmat(fx["rot"].to(DEV), lmax, 2).cpu()
    rows = fx[key]  # [E, M, L] rows of the dense block-diagonal matrix
    off = 0
    r = 0
    for l in range(lmax + 1):
        nr = min(2 * l + 1, 5)
        blk = rows[:, r:r + nr, l * l:(l + 1) ** 2].reshape(rows.shape[0], -1)
        rl = (2 * l + 1 + 3) // 4 * 4   # rows are padded to 16 bytes in the Wigner store
        got = wig[:, off:off + nr * rl].reshape(-1, nr, rl)
        assert float(got[:, :, 2 * l + 1:].abs().max() if rl > 2 * l + 1 else 0.0) == 0.0
        got = got[:, :, :2 * l + 1].reshape(rows.shape[0], -1)
        assert float((got - blk).abs().max()) < 5e-6, (l, float((got - blk).abs().max()))
        off += nr * rl
        r += nr
    assert off == wig.shape[1]


@pytest.mark.parametrize("lmax", [6, 8])
def test_wigner_kernel_ragged_edge_counts(lmax):
    """Edge counts that are not a multiple of the warp / CTA size (the staged stores are warp-cooperative): every
    prefix of the golden edge set gives bit-identical rows, and nothing is written past the last edge."""
    from singa_b200 import ops
    rot = load_golden("wigner_small.pt")["rot"].to(DEV)
    rot = torch.cat([rot, rot.flip(0), rot], dim=0)  # 192 edges: several warps and CTAs
    full = ops.wigner_from_rotmat(rot, lmax, 2)
    for n in (1, 31, 32, 33, 63, 65, 127, 191):
        part = ops.wigner_from_rotmat(rot[:n].contiguous(), lmax, 2)
        assert part.shape == (n, full.shape[1])
        assert torch.equal(part, full[:n]), n
    from singa_b200._lib import call, ptr
    guard = torch.full((35, full.shape[1]), 7.0, device=DEV)
    call("efl_wigner_from_rotmat", lmax, 2, ptr(rot), ptr(ops.jd_device(lmax, rot.device)), ptr(guard), 33)
    assert torch.equal(guard[:33], full[:33]) and bool((guard[33:] == 7.0).all())


@pytest.mark.parametrize("name", ["homo", "hetero"])
def test_transblock_fwd_bwd_matches_reference_golden(name):
    errs = run_attention_case(name, DEV, "fp32", return_all=True)
    bad = {k: v for k, v in errs.items()
           if v > (TOL_OUT_FP32 if k in ("block_out", "edge_degree_out") else TOL_GRAD_FP32)}
    assert not bad, bad


def _run_example(model, fx, dev):
    from singa_b200 import synthetic as syn
    g = {k: v.to(dev) for k, v in fx["graph"].items()}
    frames = {k: v.to(dev) for k, v in fx["frames"].items()}
    out = model(syn.GraphView(g), frames=frames)
    return out


def test_example_complex_forward_backward_matches_reference():
    """BASELINE.json config 1: example/3wi2_4tpp.pt through all four passes, outputs and parameter gradients."""
    fx = load_golden("example_3wi2.pt")
    model = build_model(DEV, fx["param_seed"], "fp32")
    out = _run_example(model, fx, DEV)
    for k, ref in fx["outputs"].items():
        assert rel_err(out[k].embedding, ref) < 2e-5, (k, rel_err(out[k].embedding, ref))

    def ptensor(shape, s):
        return torch.from_numpy(np.random.RandomState(s).normal(size=tuple(shape)).astype(np.float32)).to(DEV)

    loss = (out[PROTEIN].embedding * ptensor(out[PROTEIN].embedding.shape, 101)).sum() + \
           (out[LIGAND].embedding * ptensor(out[LIGAND].embedding.shape, 102)).sum()
    assert abs(float(loss) - fx["loss"]) < 2e-4 * max(1.0, abs(fx["loss"]))
    loss.backward()
    params = dict(model.named_parameters())
    for n in fx["params_without_grad"]:
        assert params[n].grad is None, n
    bad = {}
    for n, dig in fx["grads"].items():
        e = digest_err(params[n].grad, dig)
        if e > 2e-4:
            bad[n] = e
    assert not bad, bad


def test_dead_block_skip_changes_nothing():
    fx = load_golden("example_3wi2.pt")
    a = _run_example(build_model(DEV, 0, "fp32", compute_dead_blocks=True), fx, DEV)
    b = _run_example(build_model(DEV, 0, "fp32", compute_dead_blocks=False), fx, DEV)
    for k in a:
        assert torch.equal(a[k].embedding, b[k].embedding), k


def test_bitwise_determinism_and_chunk_invariance():
    """Deterministic segment kernels: two runs are bit-identical, and so is a run with tiny edge chunks
    (forward; weight gradients differ only by the chunked summation order)."""
    from singa_b200 import ops
    from singa_b200 import EF_layers as ef
    fx = load_golden("example_3wi2.pt")
    model = build_model(DEV, 0, "fp32")
    with torch.no_grad():
        a = _run_example(model, fx, DEV)
        b = _run_example(model, fx, DEV)
        old = ops.MAX_CHUNK_EDGES
        try:
            ops.MAX_CHUNK_EDGES = 37
            ef._CSR_CACHE.clear()
            c = _run_example(model, fx, DEV)
        finally:
            ops.MAX_CHUNK_EDGES = old
            ef._CSR_CACHE.clear()
    for k in a:
        assert torch.equal(a[k].embedding, b[k].embedding), k
        assert torch.equal(a[k].embedding, c[k].embedding), k


def test_edge_order_permutation_invariance():
    """Shuffling the edge list (with its frames) leaves the result bit-identical: the CSR sort is stable per
    destination only up to the given order, so sources are summed in a different order -> tolerance, not bits."""
    fx = load_golden("attention_small.pt")
    case = fx["cases"]["homo"]
    from singa_b200 import EF_layers as ef
    model = build_model(DEV, fx["param_seed"], "fp32")
    blk = model.blocks[0]
    E = case["edge_index"].shape[1]
    perm = torch.from_numpy(np.random.RandomState(0).permutation(E))
    outs = []
    for p in (torch.arange(E), perm):
        model.SO3_rotation[0].set_wigner(case["rot"][p].to(DEV))
        x = ef.SO3_Embedding(0, [6], 16, dtype=torch.float32, device=DEV)
        x.set_embedding(case["x_src"].to(DEV).clone())
        z = torch.full((x.embedding.shape[0],), 6, device=DEV)
        with torch.no_grad():
            outs.append(blk(x=x, atomic_numbers=z, edge_distance=case["x_edge"][p].to(DEV),
                            edge_index=case["edge_index"][:, p].to(DEV).contiguous(), batch=1, hetero=False).embedding)
    assert rel_err(outs[1], outs[0]) < 2e-6


def test_oracle_and_cuda_agree_on_synthetic_radius_graph():
    """Degree ~ 60 radius-graph complex (config-2 shaped, one small complex) against the CPU oracle."""
    from oracle import ef_oracle as eo
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import DEFAULT_CONFIG
    g = syn.make_complex(11, n_protein=90, n_ligand=14, cutoff=8.0)
    rnd = syn.frame_randoms(g, 4)
    frames = {}
    for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                lp=("pos_l", "pos_p", "ei_lp")).items():
        frames[key] = eo.edge_rot_mat(g[s][g[ei][0]] - g[d][g[ei][1]], rnd[key])
    model = build_model(DEV, 2, "fp32")
    out = model(syn.GraphView({k: v.to(DEV) for k, v in g.items()}), frames={k: v.to(DEV) for k, v in frames.items()})
    p = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    with torch.no_grad():
        ref = eo.embedding_forward(p, g, frames, oracle_tables(), dict(DEFAULT_CONFIG))
    for k in ref:
        assert rel_err(out[k].embedding, ref[k]) < 2e-5, (k, rel_err(out[k].embedding, ref[k]))


def test_second_model_in_one_process_uses_its_own_weights():
    """Round-2 review, weak #1: model(seed A) runs and is freed, model(seed B) is built in the blocks the allocator
    hands back and must run with ITS weights: both compared with the CPU oracle on their own state_dict, and the two
    outputs must differ.  The reference reads its weights on every call (EF_layers.py:807-875)."""
    import gc
    from oracle import ef_oracle as eo
    from singa_b200 import synthetic as syn
    from singa_b200.embedding import DEFAULT_CONFIG
    g = syn.make_complex(5, n_protein=40, n_ligand=9, cutoff=8.0)
    rnd = syn.frame_randoms(g, 7)
    frames = {}
    for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                lp=("pos_l", "pos_p", "ei_lp")).items():
        frames[key] = eo.edge_rot_mat(g[s][g[ei][0]] - g[d][g[ei][1]], rnd[key])
    gd = {k: v.to(DEV) for k, v in g.items()}
    fd = {k: v.to(DEV) for k, v in frames.items()}
    outs = []
    for seed in (3, 3, 8):
        model = build_model(DEV, seed, "fp32")
        out = model(syn.GraphView(gd), frames=fd)
        p = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        with torch.no_grad():
            ref = eo.embedding_forward(p, g, frames, oracle_tables(), dict(DEFAULT_CONFIG))
        for k in ref:
            assert rel_err(out[k].embedding, ref[k]) < 2e-5, (seed, k, rel_err(out[k].embedding, ref[k]))
        outs.append({k: v.embedding.detach().clone() for k, v in out.items()})
        del model, out
        gc.collect()
    k0 = sorted(outs[0])[0]
    assert rel_err(outs[0][k0], outs[1][k0]) < 1e-6 and rel_err(outs[2][k0], outs[1][k0]) > 1e-2


TOL_OUT_BF16, TOL_GRAD_BF16 = 2e-2, 6e-2


@pytest.mark.parametrize("name", ["homo", "hetero"])
def test_transblock_bf16_tensor_core_mode_within_stated_bound(name):
    """tcgen05 path: bf16 GEMM operands / edge activations, fp32 accumulation and fp32 everywhere else.
    Stated looser bound (north star): 2e-2 on outputs, 6e-2 on gradients, relative to ||ref||_inf."""
    errs = run_attention_case(name, DEV, "bf16", return_all=True)
    bad = {k: v for k, v in errs.items()
           if v > (TOL_OUT_BF16 if k in ("block_out", "edge_degree_out") else TOL_GRAD_BF16)}
    assert not bad, bad


TOL_OUT_X3, TOL_GRAD_X3 = 1e-3, 5e-3


@pytest.mark.parametrize("name", ["homo", "hetero"])
def test_transblock_bf16x3_accurate_tensor_core_mode(name):
    """The accurate tensor-core mode: tcgen05 GEMMs on error-compensated bf16 (hi, lo) operand
    splits, fp32 activations.  Stated bound 1e-3 on outputs, 5e-3 on gradients; measured values are far inside it
    (profiles/r2_parity_config2.json)."""
    from singa_b200 import ops
    try:
        errs = run_attention_case(name, DEV, "bf16x3", return_all=True)
    finally:
        ops.set_precision("fp32")
    bad = {k: v for k, v in errs.items()
           if v > (TOL_OUT_X3 if k in ("block_out", "edge_degree_out") else TOL_GRAD_X3)}
    assert not bad, bad
    assert max(errs.values()) < 1e-3, errs     # in practice ~1e-5: a regression to plain bf16 would show here


def test_example_complex_bf16_mode():
    fx = load_golden("example_3wi2.pt")
    model = build_model(DEV, fx["param_seed"], "bf16")
    with torch.no_grad():
        out = _run_example(model, fx, DEV)
    for k, ref in fx["outputs"].items():
        assert rel_err(out[k].embedding, ref) < 3e-2, (k, rel_err(out[k].embedding, ref))


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_saved_activations_and_recompute_give_identical_gradients(precision):
    """The saved-for-backward policy (ops.set_save_policy) only chooses between keeping a chunk's edge activations
    and re-running the same forward kernels in backward: gradients must be bit-identical."""
    from singa_b200 import ops
    fx = load_golden("example_3wi2.pt")
    grads = {}
    old = ops._SAVE_POLICY
    try:
        for policy in ("save", "recompute"):
            ops.set_save_policy(policy)
            model = build_model(DEV, 0, precision)
            out = _run_example(model, fx, DEV)
            (out[PROTEIN].embedding.sum() + 2.0 * out[LIGAND].embedding.sum()).backward()
            grads[policy] = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
    finally:
        ops.set_save_policy(old)
        ops.set_precision("fp32")
    assert grads["save"].keys() == grads["recompute"].keys()
    for n in grads["save"]:
        assert torch.equal(grads["save"][n], grads["recompute"][n]), n


@pytest.mark.parametrize("name", ["homo", "hetero"])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_transblock_lmax8_matches_reference_golden(name, precision):
    """BASELINE.json config 5 coefficient layout (lmax=8: L=81, M=39) against a golden written by the unmodified
    reference (oracle/make_golden.py: attention_l8.pt)."""
    from singa_b200 import ops
    try:
        errs = run_attention_case(name, DEV, precision, return_all=True, fixture="attention_l8.pt")
    finally:
        ops.set_precision("fp32")
    to, tg = (TOL_OUT_FP32, TOL_GRAD_FP32) if precision == "fp32" else (TOL_OUT_BF16, TOL_GRAD_BF16)
    bad = {k: v for k, v in errs.items() if v > (to if k in ("block_out", "edge_degree_out") else tg)}
    assert not bad, bad
