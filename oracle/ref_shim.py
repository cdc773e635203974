"""TEST INFRASTRUCTURE ONLY - import the UNMODIFIED reference from /root/reference.

This shim lets the reference's hot-path modules (rl4co/envs/**, rl4co/utils/{ops,decoding}.py,
rl4co/models/zoo/am/**, rl4co/models/nn/**) import and run in THIS container, where
tensordict / torchrl / lightning / hydra / matplotlib are absent.  It is used by
``oracle/gen_golden.py`` to produce the committed fixtures under ``tests/golden/`` and by
``oracle/check_port.py`` to validate ``oracle/port.py`` against the real reference code.
It cannot travel to the GPU box (/root/reference does not exist there); nothing in
``-m gpu`` tests, smoke() or bench.py imports it.  No reference source is copied: the
modules are executed from where they lie.

Third-party behaviour restated here (SURVEY.md 8c "parity unpinned" for item 1):
  1. torchrl.envs.EnvBase.reset (torchrl>=0.4, pyproject.toml:77; call site
     rl4co/envs/common/base.py:143): calls ``_reset``, adds done/terminated [B,1] False,
     merges the result into the input td.
  2. tensordict.TensorDict batch-dim view/expand/permute (rl4co/utils/ops.py:16,42).
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types

import torch
import torch.nn as nn

REF_ROOT = os.environ.get("RL4CO_REF_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))

from rl4co_slap_b200.tensordict_compat import CompatTensorDict  # noqa: E402

_STUB_TOPLEVEL = (
    "tensordict", "torchrl", "lightning", "matplotlib", "robust_downloader", "pyrootutils",
    "hydra", "omegaconf", "vrplib", "pyvrp", "torch_geometric", "wandb", "rich",
)


class _Anything:
    """Permissive dummy: any attribute / call / subscript yields another dummy."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return _Anything()

    def __getitem__(self, k):
        return _Anything()

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (object,)


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        cls = type(name, (_Anything,), {})
        setattr(self, name, cls)
        return cls


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in _STUB_TOPLEVEL and fullname not in sys.modules:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


class ShimEnvBase(nn.Module):
    """Stand-in for torchrl.envs.EnvBase (behaviour documented in the module header)."""

    batch_locked = False

    def __init__(self, device="cpu", batch_size=None, run_type_checks=False,
                 allow_done_after_reset=False, **_):
        super().__init__()
        self.__dict__["_device"] = torch.device(device)
        self.__dict__["_bs"] = torch.Size([]) if batch_size is None else torch.Size(batch_size)

    @property
    def device(self):
        return self.__dict__["_device"]

    @device.setter
    def device(self, v):
        self.__dict__["_device"] = torch.device(v)

    @property
    def batch_size(self):
        return self.__dict__["_bs"]

    @batch_size.setter
    def batch_size(self, v):
        self.__dict__["_bs"] = torch.Size(v)

    def __setattr__(self, name, value):
        # spec objects are inert stubs; nn.Module.__setattr__ is fine with plain objects
        if isinstance(value, _Anything):
            self.__dict__[name] = value
            return
        super().__setattr__(name, value)

    def set_seed(self, seed=None, static_seed=False):
        self._set_seed(seed)
        return seed

    def to(self, device):
        self.__dict__["_device"] = torch.device(device)
        return self

    def reset(self, tensordict=None, batch_size=None, **kwargs):
        td = tensordict
        if batch_size is None:
            batch_size = td.batch_size if td is not None else self.batch_size
        batch_size = [batch_size] if isinstance(batch_size, int) else list(batch_size)
        out = self._reset(td, batch_size=batch_size)
        dev = out.device
        for key in ("done", "terminated"):
            if key not in out.keys():
                out.set(key, torch.zeros(*batch_size, 1, dtype=torch.bool, device=dev))
        if td is not None:
            merged = td.copy() if hasattr(td, "copy") else td
            merged.update(out)
            merged.batch_size = batch_size
            return merged
        return out


def install():
    """Idempotently install stubs and register the reference as package ``rl4co``."""
    if "rl4co" in sys.modules and getattr(sys.modules["rl4co"], "_shimmed", False):
        return
    if not os.path.isdir(os.path.join(REF_ROOT, "rl4co")):
        raise RuntimeError(f"reference not present at {REF_ROOT}; the shim only works in the build container")
    sys.meta_path.insert(0, _StubFinder())

    td_mod = _StubModule("tensordict")
    td_mod.__path__ = []
    td_mod.TensorDict = CompatTensorDict
    td_mod.TensorDictBase = CompatTensorDict
    sys.modules["tensordict"] = td_mod
    tdt = _StubModule("tensordict.tensordict")
    tdt.__path__ = []
    tdt.TensorDict = CompatTensorDict
    tdt.TensorDictBase = CompatTensorDict
    sys.modules["tensordict.tensordict"] = tdt

    trl = _StubModule("torchrl")
    trl.__path__ = []
    sys.modules["torchrl"] = trl
    trl_envs = _StubModule("torchrl.envs")
    trl_envs.__path__ = []
    trl_envs.EnvBase = ShimEnvBase
    sys.modules["torchrl.envs"] = trl_envs
    trl.envs = trl_envs

    lit = _StubModule("lightning")
    lit.__path__ = []

    class LightningModule(nn.Module):
        def save_hyperparameters(self, *a, **k):
            pass

    lit.LightningModule = LightningModule
    sys.modules["lightning"] = lit
    rz = _StubModule("lightning.pytorch.utilities.rank_zero")
    rz.__path__ = []
    rz.rank_zero_only = lambda f: f
    for name in ("lightning.pytorch", "lightning.pytorch.utilities"):
        m = _StubModule(name)
        m.__path__ = []
        sys.modules[name] = m
    sys.modules["lightning.pytorch.utilities.rank_zero"] = rz
    sys.modules["lightning.pytorch.utilities"].rank_zero = rz
    sys.modules["lightning.pytorch.utilities"].rank_zero_only = rz.rank_zero_only

    pkg = types.ModuleType("rl4co")
    pkg.__path__ = [os.path.join(REF_ROOT, "rl4co")]
    pkg.__version__ = "0.5.0.dev1"
    pkg._shimmed = True
    sys.modules["rl4co"] = pkg


def upstream_tsp_context(embed_dim: int):
    """Build the upstream TSP context the fork left only as a comment
    (rl4co/models/nn/env_embeddings/context.py:102-137) on top of the reference's own
    ``EnvContext`` base (context.py:47-71).  The live ``TSPContext`` (context.py:176-248)
    needs EVRP keys TSPEnv never produces."""
    install()
    from rl4co.models.nn.env_embeddings.context import EnvContext
    from rl4co.utils.ops import gather_by_index

    class UpstreamTSPContext(EnvContext):
        def __init__(self, embed_dim):
            super().__init__(embed_dim, 2 * embed_dim)
            self.W_placeholder = nn.Parameter(torch.Tensor(2 * self.embed_dim).uniform_(-1, 1))

        def forward(self, embeddings, td):
            bsz = embeddings.size(0)
            node_dim = (-1,) if td["first_node"].dim() == 1 else (td["first_node"].size(-1), -1)
            if td["i"][(0,) * td["i"].dim()].item() < 1:
                if len(td.batch_size) < 2:
                    ctx = self.W_placeholder[None, :].expand(bsz, self.W_placeholder.size(-1))
                else:
                    ctx = self.W_placeholder[None, None, :].expand(
                        bsz, td.batch_size[1], self.W_placeholder.size(-1))
            else:
                ctx = gather_by_index(
                    embeddings,
                    torch.stack([td["first_node"], td["current_node"]], -1).view(bsz, -1),
                ).view(bsz, *node_dim)
            return self.project_context(ctx)

    return UpstreamTSPContext(embed_dim)


def slap_modules(embed_dim: int):
    """The three SLAP policy modules of examples/slap.py:11-71, loaded from that file
    (its __main__ block is not executed; its debug print at :29 is silenced)."""
    install()
    import contextlib
    import io
    import runpy

    # examples/slap.py imports rl4co.models (whole zoo) - heavy but works under the stubs.
    with contextlib.redirect_stdout(io.StringIO()):
        ns = runpy.run_path(os.path.join(REF_ROOT, "examples", "slap.py"), run_name="slap_example")
    init_cls, ctx_cls, dyn_cls = ns["SLAPInitEmbedding"], ns["SLAPContext"], ns["StaticEmbedding"]

    class QuietInit(init_cls):
        def forward(self, td):
            with contextlib.redirect_stdout(io.StringIO()):
                return super().forward(td)

    return QuietInit(embed_dim), ctx_cls(embed_dim), dyn_cls(embed_dim)
