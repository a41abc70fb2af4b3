"""TEST INFRASTRUCTURE ONLY -- loads the UNMODIFIED reference model/EF_layers.py
and model/Embedding.py from /root/reference on CPU.

Usable where /root/reference exists (the build container) or where oracle/_ref/ holds the build-time copy of the
path's three files (the GPU box; see REFERENCE_ROOT below).  Used by bench.py's reference arm / cpu_baseline, by
oracle/make_golden.py to generate tests/golden/*.pt and by the not-gpu tests to
pin oracle/ef_oracle.py (the travelling restatement) against the real code.
Nothing in singa_b200/ may import this file.

The reference imports three packages that are absent and un-vendored
(environment.yml:90,171-175).  They are replaced by shims that restate the
published behaviour of exactly the symbols used on the path:

  e3nn==0.5.1        o3.xyz_to_angles, o3.angles_to_matrix      EF_layers.py:510-515
                     o3.ToS2Grid(.shb,.sha), o3.FromS2Grid      EF_layers.py:562-587
  torch_geometric==2.3.1  utils.softmax                         EF_layers.py:1180
                     data.HeteroData / Batch, loader.DataLoader Embedding.py:7-10 (types only)
  easydict           EasyDict                                   Embedding.py:1

Parity status of the shims: the S2-grid per-degree normalisation constants are
"parity unpinned" (no e3nn here, no reference checkpoint); self-checks in
tests/test_constants.py: from_grid o to_grid = I, Y(R r) = D(R) Y(r) against
the shipped Jd.pt.  The Legendre values here come from scipy.special.lpmv, an
independent route from the recurrence used by singa_b200/so3_constants.py.
"""
from __future__ import annotations

import importlib
import importlib.util
import math
import os
import pickle
import sys
import types

import numpy as np
import torch

# Where the unmodified reference files are read from: /root/reference in the build container; on the GPU box, which
# has no /root/reference, the three files of the path that __graft_entry__.build_oracle_ref() copied (unchanged, at
# build time, never committed) into the git-ignored oracle/_ref/ -- they travel with the snapshot like libefl.so does.
_SHIPPED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REFERENCE_FILES = ("model/EF_layers.py", "model/Embedding.py", "model/Jd.pt")


def _pick_root() -> str:
    env = os.environ.get("SINGA_REFERENCE_ROOT")
    if env:
        return env
    for root in ("/root/reference", _SHIPPED_ROOT):
        if os.path.isfile(os.path.join(root, "model", "EF_layers.py")):
            return root
    return "/root/reference"


REFERENCE_ROOT = _pick_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "model", "EF_layers.py"))


# ---------------------------------------------------------------------------
# e3nn shim
# ---------------------------------------------------------------------------
def _xyz_to_angles(xyz):
    xyz = torch.nn.functional.normalize(xyz, p=2, dim=-1)
    xyz = xyz.clamp(-1, 1)
    beta = torch.acos(xyz[..., 1])
    alpha = torch.atan2(xyz[..., 0], xyz[..., 2])
    return alpha, beta


def _matrix_x(angle):
    c, s = angle.cos(), angle.sin()
    o, z = torch.ones_like(angle), torch.zeros_like(angle)
    return torch.stack([torch.stack([o, z, z], -1), torch.stack([z, c, -s], -1), torch.stack([z, s, c], -1)], -2)


def _matrix_y(angle):
    c, s = angle.cos(), angle.sin()
    o, z = torch.ones_like(angle), torch.zeros_like(angle)
    return torch.stack([torch.stack([c, z, s], -1), torch.stack([z, o, z], -1), torch.stack([-s, z, c], -1)], -2)


def _angles_to_matrix(alpha, beta, gamma):
    alpha, beta, gamma = torch.broadcast_tensors(alpha, beta, gamma)
    return _matrix_y(alpha) @ _matrix_x(beta) @ _matrix_y(gamma)


def _legendre_scipy(lmax, betas):
    from scipy.special import lpmv
    z = np.cos(betas)
    out = np.zeros((len(betas), (lmax + 1) ** 2))
    for l in range(lmax + 1):
        for m in range(0, l + 1):
            # scipy includes the Condon-Shortley phase (-1)^m; e3nn does not
            p = lpmv(m, l, z) * (-1.0) ** m
            norm = math.sqrt((2 * l + 1) / (4 * math.pi) * math.factorial(l - m) / math.factorial(l + m))
            out[:, l * l + l + m] = norm * p
            out[:, l * l + l - m] = norm * p
    return out


def _sha(lmax, alphas):
    cols = [math.sqrt(2) * np.sin(m * alphas) for m in range(lmax, 0, -1)]
    cols.append(np.ones_like(alphas))
    cols += [math.sqrt(2) * np.cos(m * alphas) for m in range(1, lmax + 1)]
    return np.stack(cols, axis=-1)


def _quad_w(b):
    k = np.arange(b)
    w = np.array([(2.0 / b) * math.sin(math.pi * (2.0 * j + 1.0) / (4.0 * b))
                  * ((1.0 / (2 * k + 1)) * np.sin((2 * j + 1) * (2 * k + 1) * math.pi / (4.0 * b))).sum()
                  for j in range(2 * b)])
    return w / (2.0 * (2 * b) ** 2)


def _shb(lmax, res_beta, n, qw=None):
    betas = (np.arange(res_beta) + 0.5) / res_beta * math.pi
    leg = _legendre_scipy(lmax, betas)
    shb = np.zeros((2 * lmax + 1, res_beta, (lmax + 1) ** 2))
    for l in range(lmax + 1):
        for m in range(-l, l + 1):
            i = l * l + l + m
            v = leg[:, i] * n[l]
            if qw is not None:
                v = v * qw
            shb[lmax + m, :, i] = v
    return shb


class _ToS2Grid(torch.nn.Module):
    def __init__(self, lmax=None, res=None, normalization="component", dtype=None, device=None):
        super().__init__()
        res_beta, res_alpha = res
        if normalization == "component":
            n = [math.sqrt(4 * math.pi) / math.sqrt(2 * l + 1) / math.sqrt(lmax + 1) for l in range(lmax + 1)]
        elif normalization == "norm":
            n = [math.sqrt(4 * math.pi) / math.sqrt(lmax + 1)] * (lmax + 1)
        else:
            n = [1.0] * (lmax + 1)
        alphas = np.arange(res_alpha) / res_alpha * 2 * math.pi
        self.register_buffer("sha", torch.tensor(_sha(lmax, alphas), dtype=torch.float32, device=device))
        self.register_buffer("shb", torch.tensor(_shb(lmax, res_beta, n), dtype=torch.float32, device=device))


class _FromS2Grid(torch.nn.Module):
    def __init__(self, res=None, lmax=None, normalization="component", lmax_in=None, dtype=None, device=None):
        super().__init__()
        res_beta, res_alpha = res
        if normalization == "component":
            n = [math.sqrt(4 * math.pi) * math.sqrt(2 * l + 1) * math.sqrt(lmax + 1) for l in range(lmax + 1)]
        elif normalization == "norm":
            n = [math.sqrt(4 * math.pi) * math.sqrt(lmax + 1)] * (lmax + 1)
        else:
            n = [4 * math.pi] * (lmax + 1)
        assert res_beta % 2 == 0
        qw = _quad_w(res_beta // 2) * res_beta ** 2 / res_alpha
        alphas = np.arange(res_alpha) / res_alpha * 2 * math.pi
        self.register_buffer("sha", torch.tensor(_sha(lmax, alphas), dtype=torch.float32, device=device))
        self.register_buffer("shb", torch.tensor(_shb(lmax, res_beta, n, qw), dtype=torch.float32, device=device))


def _install_e3nn():
    e3nn = types.ModuleType("e3nn")
    o3 = types.ModuleType("e3nn.o3")
    o3.xyz_to_angles = _xyz_to_angles
    o3.angles_to_matrix = _angles_to_matrix
    o3.ToS2Grid = _ToS2Grid
    o3.FromS2Grid = _FromS2Grid
    e3nn.o3 = o3
    sys.modules.setdefault("e3nn", e3nn)
    sys.modules.setdefault("e3nn.o3", o3)


# ---------------------------------------------------------------------------
# torch_geometric shim
# ---------------------------------------------------------------------------
def _pyg_softmax(src, index=None, ptr=None, num_nodes=None, dim=0):
    n = int(index.max()) + 1 if num_nodes is None else num_nodes
    shape = (n,) + tuple(src.shape[1:])
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    mx = torch.full(shape, float("-inf"), dtype=src.dtype).scatter_reduce(0, idx, src.detach(), "amax", include_self=True)
    out = (src - mx.gather(0, idx)).exp()
    den = torch.zeros(shape, dtype=src.dtype).scatter_add(0, idx, out) + 1e-16
    return out / den.gather(0, idx)


class _Store(dict):
    """dict-like stand-in for pyg storages (attribute and item access)."""
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class HeteroDataStub:
    """Minimal HeteroData: g['node_type']['x'], g[(s, r, d)]['edge_index'], g['atomicnum'][...]"""
    def __init__(self):
        self._stores = {}

    def __getitem__(self, key):
        return self._stores[key]

    def __setitem__(self, key, value):
        self._stores[key] = value

    def keys(self):
        return self._stores.keys()


class BatchStub(HeteroDataStub):
    pass


def _install_pyg():
    pyg = types.ModuleType("torch_geometric")
    utils = types.ModuleType("torch_geometric.utils")
    utils.softmax = _pyg_softmax
    data = types.ModuleType("torch_geometric.data")
    data.HeteroData = HeteroDataStub
    data.Batch = BatchStub
    loader = types.ModuleType("torch_geometric.loader")
    loader.DataLoader = object
    pyg.utils, pyg.data, pyg.loader = utils, data, loader
    for name, mod in (("torch_geometric", pyg), ("torch_geometric.utils", utils),
                      ("torch_geometric.data", data), ("torch_geometric.loader", loader)):
        sys.modules.setdefault(name, mod)


class EasyDict(dict):
    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = EasyDict(v) if isinstance(v, dict) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    __setattr__ = dict.__setitem__


def _install_easydict():
    m = types.ModuleType("easydict")
    m.EasyDict = EasyDict
    sys.modules.setdefault("easydict", m)


# ---------------------------------------------------------------------------
# loader
# ---------------------------------------------------------------------------
_CACHE = {}


def load_reference():
    """Returns (EF_layers module, Embedding module) of the unmodified reference, CPU."""
    if "mods" in _CACHE:
        return _CACHE["mods"]
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    _install_e3nn()
    _install_pyg()
    _install_easydict()
    orig_load = torch.load

    def _load(*a, **k):  # Jd.pt is a pickled list (EF_layers.py:2198); torch>=2.6 defaults weights_only=True
        k.setdefault("weights_only", False)
        return orig_load(*a, **k)

    torch.load = _load
    try:
        pkg = types.ModuleType("ref_model")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "model")]
        sys.modules["ref_model"] = pkg
        spec = importlib.util.spec_from_file_location(
            "ref_model.EF_layers", os.path.join(REFERENCE_ROOT, "model", "EF_layers.py"))
        ef = importlib.util.module_from_spec(spec)
        sys.modules["ref_model.EF_layers"] = ef
        spec.loader.exec_module(ef)
        # Embedding.py does `from EF_layers import ...` first (Embedding.py:12-22)
        sys.modules["EF_layers"] = ef
        spec2 = importlib.util.spec_from_file_location(
            "ref_model.Embedding", os.path.join(REFERENCE_ROOT, "model", "Embedding.py"))
        emb = importlib.util.module_from_spec(spec2)
        sys.modules["ref_model.Embedding"] = emb
        spec2.loader.exec_module(emb)
    finally:
        torch.load = orig_load
        sys.modules.pop("EF_layers", None)
    _CACHE["mods"] = (ef, emb)
    return ef, emb


DEFAULT_EMBEDDING_CONFIG = dict(  # config/train.yml:27-49
    edge_channels=16, sphere_channels=16, attn_hidden_channels=128, attn_alpha_channels=32,
    attn_value_channels=16, ffn_hidden_channels=512, lmax_list=[6], mmax_list=[2], cutoff=10.0,
    max_num_elements=43, num_heads=7, num_layers=3, norm_type="rms_norm_sh",
    activation_type="scaled_silu", use_atom_edge_embedding=True, share_atom_edge_embedding=True,
    grid_resolution=None, alpha_drop=0.0, proj_drop=0.0, drop_path_rate=0.0,
)


# ---------------------------------------------------------------------------
# example/*.pt reader (pickled pyg HeteroData; classes stubbed)
# ---------------------------------------------------------------------------
class _Anything:
    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        self._state = state


class _StubUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module.startswith("torch_geometric"):
            return type(name, (_Anything,), {})
        if module.startswith("numpy.core"):
            module = module.replace("numpy.core", "numpy._core")
        return super().find_class(module, name)


class _StubPickle:
    Unpickler = _StubUnpickler
    __name__ = "pickle"

    @staticmethod
    def load(f, **kw):
        return _StubUnpickler(f, **kw).load()


def _mapping(store):
    st = getattr(store, "_state", None) or store.__dict__
    return st["_mapping"]


def load_example(path):
    """Reads one example/*.pt into a HeteroDataStub with plain tensors."""
    obj = torch.load(path, map_location="cpu", weights_only=False, pickle_module=_StubPickle)
    st = getattr(obj, "_state", None) or obj.__dict__
    g = HeteroDataStub()
    for nt, store in st["_node_store_dict"].items():
        g[nt] = _Store({k: v for k, v in _mapping(store).items()})
    for et, store in st["_edge_store_dict"].items():
        g[tuple(et)] = _Store({k: v for k, v in _mapping(store).items()})
    for k, v in _mapping(st["_global_store"]).items():
        g[k] = v
    return g
