"""The CUDA path against the LIVE, unmodified reference on fresh inputs (not a frozen golden).

The reference's three files of this path are copied at build time into the git-ignored oracle/_ref/ and travel to the
GPU box with the snapshot (``__graft_entry__.build_oracle_ref``); oracle/ref_loader.py imports them there.  The
reference runs on the host in fp32, the B200 modules on the device, with the same state_dict (strict load, the
surface is identical), the same edge-frame random draws and the same loss.  Reference span: model/Embedding.py:205-480
over model/EF_layers.py."""
import pytest
import torch

from oracle import ref_loader as rl
from singa_b200 import synthetic as syn
from tests.parity_util import build_model, rel_err

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rl.reference_available(), reason="reference files neither in /root/reference nor oracle/_ref")]
DEV = "cuda:0"


def _reference_run(g, state, frame_seed, Rp, Rl):
    ef, emb = rl.load_reference()
    torch.manual_seed(0)
    ref = emb.EquivariantEmbedding(rl.EasyDict(rl.DEFAULT_EMBEDDING_CONFIG), device="cpu")
    ref.load_state_dict({k: v.detach().cpu() for k, v in state.items()}, strict=True)
    rnd = syn.frame_randoms(g, frame_seed)
    draws = iter([rnd["pp"], rnd["ll"], rnd["lp"]])
    orig = torch.rand_like
    torch.rand_like = lambda x, **k: next(draws)
    frames = {}
    try:
        out = ref(syn.GraphView(g))
        for key, (s, d, ei) in dict(pp=("pos_p", "pos_p", "ei_pp"), ll=("pos_l", "pos_l", "ei_ll"),
                                    lp=("pos_l", "pos_p", "ei_lp")).items():
            torch.rand_like = lambda x, _r=rnd[key], **k: _r
            frames[key] = ef.init_edge_rot_mat(g[s][g[ei][0]] - g[d][g[ei][1]])
    finally:
        torch.rand_like = orig
    loss = (out[syn.PROTEIN].embedding * Rp).sum() + (out[syn.LIGAND].embedding * Rl).sum()
    loss.backward()
    return ({k: v.embedding.detach() for k, v in out.items()},
            {n: p.grad for n, p in ref.named_parameters() if p.grad is not None}, frames)


@pytest.mark.parametrize("seed,n_p,n_l,cutoff", [(21, 70, 11, 6.0), (22, 45, 16, 8.0)])
def test_fresh_complex_matches_the_live_reference_forward_and_backward(seed, n_p, n_l, cutoff):
    """fp32 mode; stated tolerance 2e-5 on outputs and 2e-4 on parameter gradients (relative to the tensor's max), the
    bounds of the golden tests."""
    g = syn.make_complex(seed, n_protein=n_p, n_ligand=n_l, cutoff=cutoff)
    model = build_model(DEV, seed, "fp32")
    gen = torch.Generator().manual_seed(seed)
    Rp, Rl = torch.randn(n_p, 49, 16, generator=gen), torch.randn(n_l, 49, 16, generator=gen)
    outs, grads, frames = _reference_run(g, model.state_dict(), seed + 100, Rp, Rl)
    out = model(syn.GraphView({k: v.to(DEV) for k, v in g.items()}), frames={k: v.to(DEV) for k, v in frames.items()})
    loss = (out[syn.PROTEIN].embedding * Rp.to(DEV)).sum() + (out[syn.LIGAND].embedding * Rl.to(DEV)).sum()
    loss.backward()
    for k in outs:
        assert rel_err(out[k].embedding, outs[k]) < 2e-5, (k, rel_err(out[k].embedding, outs[k]))
    checked = 0
    for n, p in model.named_parameters():
        if p.grad is not None and n in grads:
            assert rel_err(p.grad, grads[n]) < 2e-4, (n, rel_err(p.grad, grads[n]))
            checked += 1
    assert checked >= 60
