"""Host-side mirror of the hot path's caller, ``EquivariantEmbedding`` (reference model/Embedding.py:52-480).

The reference's own ``model/Embedding.py`` runs unmodified on top of ``singa_b200.EF_layers`` (that is the drop-in
boundary).  This class exists because the reference file cannot travel to the GPU box and because
its per-atom Python loops (Embedding.py:250-253, 328-331: a ``.tolist()`` device sync per atom) would dominate a
step at GPU speed.  Same constructor, same parameter/buffer names (a reference ``state_dict``
loads with ``strict=True``), same four passes and the same quirks:

* l=0 seed features are written into an int64 tensor (truncation toward zero; ``sphere_embedding*`` get no
  gradient) -- Embedding.py:241-247, 262;
* hetero blocks mutate both node sets with ``norm_1`` and only the last block's output is kept
  (Embedding.py:415-428); the discarded outputs of blocks 0..n-2 are dead values, so by default they are not
  computed (``compute_dead_blocks=False``; results are identical either way);
* pass 4 reuses the Wigner matrices of pass 3 (Embedding.py:437-471).

Differences: the bit-string feature code is a vectorised bit-pack, and ``forward`` accepts pre-computed edge
frames so that parity runs can feed both implementations the same random gauge.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn as nn

from .EF_layers import (CoefficientMappingModule, EdgeDegreeEmbedding, GaussianSmearing, ModuleListInfo,
                        SO3_Embedding, SO3_Grid, SO3_Rotation, TransBlockV2, get_normalization_layer,
                        init_edge_rot_mat)
from . import EF_layers as _ef
from .synthetic import LIGAND, LL, LP, PL, PP, PROTEIN

_AVG_NUM_MODES = 77.81317
_AVG_DEGREE = 23.395238876342773


class _Cfg(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


DEFAULT_CONFIG = _Cfg(  # reference config/train.yml:27-49
    edge_channels=16, sphere_channels=16, attn_hidden_channels=128, attn_alpha_channels=32, attn_value_channels=16,
    ffn_hidden_channels=512, lmax_list=[6], mmax_list=[2], cutoff=10.0, max_num_elements=43, num_heads=7,
    num_layers=3, norm_type='rms_norm_sh', activation_type='scaled_silu', use_atom_edge_embedding=True,
    share_atom_edge_embedding=True, grid_resolution=None, alpha_drop=0.0, proj_drop=0.0, drop_path_rate=0.0)


def pack_feature_bits(x: torch.Tensor) -> torch.Tensor:
    """Last 15 feature columns cast to long and read as a big-endian bit string (Embedding.py:250-253)."""
    weights = 2 ** torch.arange(14, -1, -1, dtype=torch.long, device=x.device)
    return (x[:, -15:].to(torch.long) * weights).sum(dim=1)


def batch_vector_from_ptr(ptr: torch.Tensor, num_nodes: Optional[int] = None, device=None) -> torch.Tensor:
    """Node -> complex id vector from the pyg ``ptr`` offsets: the vectorised form of the two Python loops of
    ``SINGA.forward`` (GAN.py:48-55: ``batch.extend([i] * (ptr[i+1] - ptr[i]))``), which cost one device read per
    complex.  ``num_nodes`` (= ``ptr[-1]``, known to the caller as ``len(x)``) avoids the one remaining read."""
    counts = ptr[1:] - ptr[:-1]
    ids = torch.arange(counts.numel(), dtype=torch.long, device=ptr.device)
    out = torch.repeat_interleave(ids, counts, output_size=num_nodes)
    return out if device is None else out.to(device)


def stack_props(dic: Dict[str, torch.Tensor], props) -> torch.Tensor:
    """``torch.tensor(list(zip(*[dic[p] for p in props])))`` (GAN.py:42) without the per-element host round trip:
    [num_complexes, len(props)] float32."""
    return torch.stack([dic[p].reshape(-1).to(torch.float32) for p in props], dim=1)


class EquivariantEmbedding(nn.Module):
    def __init__(self, config, device: str = 'cuda', compute_dead_blocks: bool = False) -> None:
        super().__init__()
        cfg = _Cfg(config)
        self.device = device
        self.compute_dead_blocks = compute_dead_blocks
        self.edge_channels = cfg.edge_channels
        self.sphere_channels = cfg.sphere_channels
        self.attn_hidden_channels = cfg.attn_hidden_channels
        self.attn_alpha_channels = cfg.attn_alpha_channels
        self.attn_value_channels = cfg.attn_value_channels
        self.ffn_hidden_channels = cfg.ffn_hidden_channels
        self.lmax_list = [int(i) for i in cfg.lmax_list]
        self.mmax_list = [int(i) for i in cfg.mmax_list]
        self.num_resolutions = len(self.lmax_list)
        self.sphere_channels_all = self.num_resolutions * self.sphere_channels
        self.max_num_elements = cfg.max_num_elements
        self.num_heads = cfg.num_heads
        self.num_layers = cfg.num_layers
        self.offset = 0
        self.offset_res = 0
        self.cutoff = cfg.cutoff
        self.alpha_drop, self.proj_drop, self.drop_path_rate = cfg.alpha_drop, cfg.proj_drop, cfg.drop_path_rate
        self.norm_type = cfg.norm_type
        self.activation_type = cfg.activation_type
        self.grid_resolution = None  # the reference ignores config.grid_resolution (Embedding.py:90)
        self.share_atom_edge_embedding = cfg.share_atom_edge_embedding
        self.use_atom_edge_embedding = cfg.use_atom_edge_embedding
        if self.share_atom_edge_embedding:
            assert self.use_atom_edge_embedding
            self.block_use_atom_edge_embedding = False
        else:
            self.block_use_atom_edge_embedding = self.use_atom_edge_embedding

        self.SO3_rotation = nn.ModuleList(SO3_Rotation(l, device=device) for l in self.lmax_list)
        self.sphere_embedding = nn.Embedding(self.max_num_elements, self.sphere_channels_all, device=device)
        self.sphere_embedding_2 = nn.Embedding(32767, self.sphere_channels_all, device=device)
        self.distance_expansion = GaussianSmearing(start=0.0, stop=self.cutoff, num_gaussians=self.edge_channels,
                                                   basis_width_scalar=20, device=device)
        self.edge_channels_list = [int(self.distance_expansion.num_output)] + [self.edge_channels] * 2
        if self.share_atom_edge_embedding and self.use_atom_edge_embedding:
            self.source_embedding = nn.Embedding(self.max_num_elements, self.edge_channels_list[-1], device=device)
            self.target_embedding = nn.Embedding(self.max_num_elements, self.edge_channels_list[-1], device=device)
            self.edge_channels_list[0] = self.edge_channels_list[0] + 2 * self.edge_channels_list[-1]
        else:
            self.source_embedding, self.target_embedding = None, None
        self.mappingReduced = CoefficientMappingModule(self.lmax_list, self.mmax_list, device=device)
        top = max(self.lmax_list)
        self.SO3_grid = ModuleListInfo('({}, {})'.format(top, top))
        for l in range(top + 1):
            self.SO3_grid.append(nn.ModuleList(
                SO3_Grid(l, m, resolution=self.grid_resolution, normalization='component', device=device)
                for m in range(top + 1)))
        self.edge_degree_embedding = EdgeDegreeEmbedding(
            sphere_channels=self.sphere_channels, lmax_list=self.lmax_list, mmax_list=self.mmax_list,
            SO3_rotation=self.SO3_rotation, mappingReduced=self.mappingReduced,
            max_num_elements=self.max_num_elements, edge_channels_list=self.edge_channels_list,
            use_atom_edge_embedding=False, rescale_factor=_AVG_DEGREE, device=device)
        self.blocks = nn.ModuleList()
        for _ in range(self.num_layers):
            self.blocks.append(TransBlockV2(
                SO3_rotation=self.SO3_rotation, SO3_grid=self.SO3_grid, mappingReduced=self.mappingReduced,
                sphere_channels=self.sphere_channels, attn_hidden_channels=self.attn_hidden_channels,
                attn_alpha_channels=self.attn_alpha_channels, attn_value_channels=self.attn_value_channels,
                ffn_hidden_channels=self.ffn_hidden_channels, output_channels=self.sphere_channels,
                lmax_list=self.lmax_list, mmax_list=self.mmax_list, num_heads=self.num_heads,
                max_num_elements=self.max_num_elements, edge_channels_list=self.edge_channels_list,
                use_atom_edge_embedding=self.block_use_atom_edge_embedding, norm_type=self.norm_type,
                attn_activation=self.activation_type, ffn_activation=self.activation_type, use_m_share_rad=False,
                use_s2_act_attn=False, use_attn_renorm=True, use_gate_act=False, use_grid_mlp=False,
                use_sep_s2_act=True, alpha_drop=self.alpha_drop, proj_drop=self.proj_drop,
                drop_path_rate=self.drop_path_rate, device=device))
        self.norm = get_normalization_layer(self.norm_type, lmax=top, num_channels=self.sphere_channels, device=device)

    # -- helpers ----------------------------------------------------------------------------
    def _set_frames(self, edge_vec, frames, key):
        rot = frames[key] if frames is not None and key in frames else init_edge_rot_mat(edge_vec, device=edge_vec.device,
                                                                                     defer_checks=True)
        for r in self.SO3_rotation:
            r.set_wigner(rot)

    def _edge_scalars(self, dist, z_src, z_dst, ei):
        return torch.cat((self.distance_expansion(dist), _ef.embedding_lookup(self.source_embedding, z_src[ei[0]]),
                          _ef.embedding_lookup(self.target_embedding, z_dst[ei[1]])), dim=1)

    def _seed_nodes(self, z, feats):
        x = SO3_Embedding(len(z), self.lmax_list, self.sphere_channels, device=z.device, dtype=torch.long)
        # feats: the [N, 59] feature rows, or the already packed integer codes of the dataset format (dataset.py)
        bits = feats if feats.dim() == 1 else pack_feature_bits(feats)
        seed = self.sphere_embedding(z) + self.sphere_embedding_2(bits)
        x.embedding[:, self.offset_res, :] = seed  # float -> int64: truncation, no gradient (Embedding.py:262)
        return x

    def _homo_pass(self, z, feats, pos, ei, frames, key):
        vec = pos[ei[0]] - pos[ei[1]]
        dist = torch.norm(vec, dim=-1, p=2)
        self._set_frames(vec, frames, key)
        x = self._seed_nodes(z, feats)
        x_edge = self._edge_scalars(dist, z, z, ei)
        deg = self.edge_degree_embedding(z, x_edge, ei, hetero=False)
        x.embedding = x.embedding + deg.embedding
        for blk in self.blocks:
            x = blk(x=x, atomic_numbers=z, edge_distance=x_edge, edge_index=ei, batch=len(z), hetero=False)
        x.embedding = self.norm(x.embedding)
        return x

    def _hetero_pass(self, x_dict, zs, pos, ei, src, dst, batch):
        vec = pos[src][ei[0]] - pos[dst][ei[1]]
        dist = torch.norm(vec, dim=-1, p=2)
        x_edge = self._edge_scalars(dist, zs[src], zs[dst], ei)
        deg = self.edge_degree_embedding(zs, x_edge, ei, hetero=True, source_target=(src, dst))
        x_dict[dst].embedding = x_dict[dst].embedding + deg.embedding
        out = None
        for i, blk in enumerate(self.blocks):
            if self.compute_dead_blocks or i == len(self.blocks) - 1:
                out = blk(x=x_dict, atomic_numbers=zs, edge_distance=x_edge, edge_index=ei, batch=batch, hetero=True,
                          source_target=(src, dst))
            else:  # only the in-place norm_1 side effects of blocks 0..n-2 are observable (Embedding.py:415-424)
                x_dict[src].embedding = blk.norm_1(x_dict[src].embedding)
                x_dict[dst].embedding = blk.norm_1(x_dict[dst].embedding)
        out.embedding = self.norm(out.embedding)
        return out

    # -- forward ----------------------------------------------------------------------------
    def forward(self, g, batch: Optional[int] = None, gen_mode: bool = False, frames: Optional[Dict] = None) -> Dict:
        """g: anything indexable like the reference's HeteroData (g['atomicnum'][nt], g[nt]['x'|'pos'],
        g[(s, rel, d)]['edge_index']).  frames: optional dict(pp=, ll=, lp=) of [E,3,3] edge frames."""
        x_dict = {}
        _ef.FRAME_CHECKS.flush(wait=True)  # frame checks of the previous call (long complete: no stall)
        zs = {PROTEIN: g['atomicnum'][PROTEIN], LIGAND: g['atomicnum'][LIGAND]}
        pos = {PROTEIN: g[PROTEIN]['pos'], LIGAND: g[LIGAND]['pos']}
        if not gen_mode:  # all four edge lists are known here: one device->host read for their CSRs, none mid-step
            n_p, n_l = len(zs[PROTEIN]), len(zs[LIGAND])
            _ef.prefetch_edge_csr([(g[PP]['edge_index'], n_p, n_p), (g[LL]['edge_index'], n_l, n_l),
                                   (g[LP]['edge_index'], n_l, n_p), (g[PL]['edge_index'], n_p, n_l)])
        x_dict[PROTEIN] = self._homo_pass(zs[PROTEIN], g[PROTEIN]['x'], pos[PROTEIN], g[PP]['edge_index'], frames, 'pp')
        if gen_mode:
            return x_dict
        x_dict[LIGAND] = self._homo_pass(zs[LIGAND], g[LIGAND]['x'], pos[LIGAND], g[LL]['edge_index'], frames, 'll')
        if batch is None:
            batch = len(zs[PROTEIN])
        lp_ei = g[LP]['edge_index']
        self._set_frames(pos[LIGAND][lp_ei[0]] - pos[PROTEIN][lp_ei[1]], frames, 'lp')
        x_dict['lp_edge'] = self._hetero_pass(x_dict, zs, pos, lp_ei, LIGAND, PROTEIN, batch)
        # pass 4 keeps the Wigner matrices of pass 3 (no set_wigner in Embedding.py:437-471)
        x_dict['pl_edge'] = self._hetero_pass(x_dict, zs, pos, g[PL]['edge_index'], PROTEIN, LIGAND, batch)
        x_dict[PROTEIN].embedding = x_dict[PROTEIN].embedding + x_dict['lp_edge'].embedding
        x_dict[LIGAND].embedding = x_dict[LIGAND].embedding + x_dict['pl_edge'].embedding
        return x_dict
