"""Graph-attention encoder (stays in PyTorch per the north star).

Runs once per batch before the rollout kernels.  Module / parameter names are laid out so
that ``state_dict()`` keys equal the reference's (``encoder.init_embedding.init_embed.weight``,
``encoder.net.layers.<l>.<0..3>...``), i.e. reference checkpoints load unchanged.

Follows: rl4co/models/zoo/am/encoder.py:14-91, rl4co/models/nn/graph/attnnet.py:16-103,
rl4co/models/nn/attention.py:65-135, rl4co/models/nn/ops.py:12-56, rl4co/models/nn/mlp.py,
rl4co/models/nn/env_embeddings/init.py:51-64,111-134, examples/slap.py:11-47.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


# ---- init embeddings ------------------------------------------------------------------
def narrow_linear(x: torch.Tensor, lin: nn.Linear) -> torch.Tensor:
    """``lin(x)`` for the init embeddings (2..8 input features, 128 outputs).  On a CUDA device, outside autograd, the
    library's write-stream kernel (rl4co_linear_narrow); everywhere else the PyTorch op."""
    w = lin.weight
    if not (x.is_cuda and not torch.is_grad_enabled() and w.dtype == torch.float32 and x.dtype == torch.float32
            and w.shape[0] == 128 and 1 <= w.shape[1] <= 8 and getattr(lin, "tensor_core", True)):
        return lin(x)
    from . import _lib
    x2 = x.reshape(-1, w.shape[1]).contiguous()
    y = torch.empty(x2.shape[0], 128, dtype=torch.float32, device=x.device)
    _lib.call("rl4co_linear_narrow", x2.shape[0], w.shape[1], _lib.ptr(x2), _lib.ptr(w.detach().contiguous()),
              _lib.ptr(lin.bias.detach().contiguous()) if lin.bias is not None else None, _lib.ptr(y),
              _lib.stream_ptr(x.device))
    return y.view(*x.shape[:-1], 128)


class TSPInitEmbedding(nn.Module):
    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        return narrow_linear(td["locs"], self.init_embed)


class VRPInitEmbedding(nn.Module):
    def __init__(self, embed_dim: int, linear_bias: bool = True, node_dim: int = 3):
        super().__init__()
        self.init_embed = nn.Linear(node_dim, embed_dim, linear_bias)
        self.init_embed_depot = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"]
        depot, cities = locs[:, :1, :], locs[:, 1:, :]
        feats = torch.cat((cities, td["demand"][..., None]), -1)
        return torch.cat((narrow_linear(depot, self.init_embed_depot), narrow_linear(feats, self.init_embed)), -2)


class OPInitEmbedding(nn.Module):
    """x, y, prize for the customers, x, y for the depot (nn/env_embeddings/init.py:207-233)."""

    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(3, embed_dim, linear_bias)
        self.init_embed_depot = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"]
        depot, cities = locs[:, :1, :], locs[:, 1:, :]
        feats = torch.cat((cities, td["prize"][..., 1:, None]), -1)
        return torch.cat((narrow_linear(depot, self.init_embed_depot), narrow_linear(feats, self.init_embed)), -2)


class PCTSPInitEmbedding(nn.Module):
    """x, y, expected prize, penalty for the customers, x, y for the depot (nn/env_embeddings/init.py:174-204)."""

    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(4, embed_dim, linear_bias)
        self.init_embed_depot = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"]
        depot, cities = locs[:, :1, :], locs[:, 1:, :]
        feats = torch.cat((cities, td["expected_prize"][..., None], td["penalty"][..., 1:, None]), -1)
        return torch.cat((narrow_linear(depot, self.init_embed_depot), narrow_linear(feats, self.init_embed)), -2)


class SLAPInitEmbedding(nn.Module):
    """x, y and Manhattan distance to node 0 (examples/slap.py:11-47)."""

    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(3, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"].to(torch.float32)
        d0 = (locs - locs[:, :1, :]).abs().sum(-1)
        return narrow_linear(torch.cat((locs, d0[..., None]), -1), self.init_embed)


INIT_EMBEDDINGS = {"tsp": TSPInitEmbedding, "cvrp": VRPInitEmbedding, "sdvrp": VRPInitEmbedding, "slap": SLAPInitEmbedding,
                   "op": OPInitEmbedding, "pctsp": PCTSPInitEmbedding, "spctsp": PCTSPInitEmbedding}


# ---- linears (widths multiples of 128) on the tensor cores ----------------------------------------------
def _tc_ok(x: torch.Tensor, lin: nn.Linear) -> bool:
    w = lin.weight
    return (x.is_cuda and not torch.is_grad_enabled() and w.dtype == torch.float32 and x.dtype == torch.float32
            and w.shape[1] % 128 == 0 and w.shape[0] % 128 == 0 and max(w.shape) <= 2048
            and getattr(lin, "tensor_core", True))      # wider inputs: K chunks of 128 accumulate in TMEM


def _bn_affine(bn: nn.BatchNorm1d):
    """Eval-mode batch norm as a per-channel (scale, shift), cached until its buffers change."""
    tensors = (bn.weight, bn.bias, bn.running_mean, bn.running_var)
    vers = []
    for p in tensors:
        try:
            vers.append(p._version)
        except RuntimeError:        # inference tensor
            vers.append(-1)
    key = tuple((p.data_ptr(), v) for p, v in zip(tensors, vers))
    if getattr(bn, "_affine_key", None) != key:
        scale = bn.weight.detach() * torch.rsqrt(bn.running_var + bn.eps)
        shift = bn.bias.detach() - bn.running_mean * scale
        bn._affine, bn._affine_key = (scale.contiguous(), shift.contiguous()), key
    return bn._affine


def linear128(x: torch.Tensor, lin: nn.Linear, relu: bool = False, residual: torch.Tensor = None,
              norm: nn.BatchNorm1d = None) -> torch.Tensor:
    """``norm(act(lin(x)) + residual)`` (residual / norm optional).  On a CUDA device, outside autograd, a
    Linear whose input and output widths are multiples of 128 runs on the library's fp32-accurate tensor-core
    kernel (rl4co_linear_tc[_ex]: tcgen05, 3-term hi / lo split - FP16 halves for K = 128, TF32 for K > 128 -, fp32 accumulation - the K3 projection kernel
    with a bias / ReLU / skip / eval-batch-norm epilogue); everything else is the plain PyTorch ops.  The
    packed weight image is cached on the module.  ``norm`` must be an eval-mode BatchNorm1d."""
    w = lin.weight
    if not _tc_ok(x, lin):
        y = F.linear(x, w, lin.bias)
        y = F.relu(y) if relu else y
        if residual is not None:
            y = y + residual
        if norm is not None:
            y = norm(y.reshape(-1, y.size(-1))).view(*y.size())
        return y
    from . import _lib
    try:
        ver = w._version
    except RuntimeError:        # inference tensor
        ver = -1
    n_out, n_in = w.shape
    key = (w.data_ptr(), ver, str(w.device))
    if getattr(lin, "_tc_key", None) != key:
        n_img = (n_in // 128) * int(_lib.load().rl4co_am_weight_image_floats(n_out // 128))
        image = torch.empty(n_img, dtype=torch.float32, device=w.device)
        _lib.call("rl4co_linear_pack", n_out, n_in, _lib.ptr(w.detach().contiguous()), _lib.ptr(image),
                  _lib.stream_ptr(w.device))
        lin._tc_image, lin._tc_key = image, key
    x2 = x.reshape(-1, n_in).contiguous()
    y = torch.empty(x2.shape[0], n_out, dtype=torch.float32, device=x.device)
    bias = lin.bias.detach().contiguous() if lin.bias is not None else None
    if residual is None and norm is None:
        _lib.call("rl4co_linear_tc", x2.shape[0], n_out, n_in, _lib.ptr(x2), _lib.ptr(lin._tc_image),
                  _lib.ptr(bias), int(relu), _lib.ptr(y), _lib.stream_ptr(x.device))
    else:
        assert norm is None or not norm.training, "fused batch norm needs eval mode (running statistics)"
        scale, shift = _bn_affine(norm) if norm is not None else (None, None)
        res = residual.reshape(-1, n_out).contiguous() if residual is not None else None
        _lib.call("rl4co_linear_tc_ex", x2.shape[0], n_out, n_in, _lib.ptr(x2), _lib.ptr(lin._tc_image),
                  _lib.ptr(bias), int(relu), _lib.ptr(res), _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(y),
                  _lib.stream_ptr(x.device))
    return y.view(*x.shape[:-1], n_out)


# ---- encoder blocks -------------------------------------------------------------------
class _Skip(nn.Module):
    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, x):
        return x + self.module(x)


class _SelfAttention(nn.Module):
    def __init__(self, embed_dim: int, num_heads: int):
        super().__init__()
        self.num_heads = num_heads
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim)
        self.out_proj = nn.Linear(embed_dim, embed_dim)

    def attend(self, x):
        """Heads' outputs before the output projection, [B, N, E]."""
        B, N, E = x.shape
        qkv = linear128(x, self.Wqkv)
        if (qkv.is_cuda and not torch.is_grad_enabled() and qkv.dtype == torch.float32 and E == 128
                and self.num_heads == 8 and N <= 128 and getattr(self, "fused_attention", True)):
            # fp32 attention kernels of the library for the path's graph sizes: Q K^T on the tensor cores
            # (csrc/encoder_attention_tc.cu) or all-SIMT (csrc/encoder_attention.cu, the cross-check)
            from . import _lib
            qkv = qkv.contiguous()
            o = torch.empty(B, N, E, dtype=torch.float32, device=x.device)
            fn = "rl4co_encoder_attention_tc" if getattr(self, "tensor_core", True) else "rl4co_encoder_attention"
            _lib.call(fn, B, N, _lib.ptr(qkv), _lib.ptr(o), _lib.stream_ptr(x.device))
            return o
        qkv = qkv.view(B, N, 3, self.num_heads, E // self.num_heads)
        q, k, v = qkv.permute(2, 0, 3, 1, 4).unbind(0)
        o = F.scaled_dot_product_attention(q, k, v)
        return o.transpose(1, 2).reshape(B, N, E)

    def forward(self, x):
        return linear128(self.attend(x), self.out_proj)


class _FeedForward(nn.Module):
    def __init__(self, embed_dim: int, hidden: int):
        super().__init__()
        dims = [embed_dim] + ([hidden] if hidden > 0 else []) + [embed_dim]
        self.lins = nn.ModuleList(nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:]))

    def forward(self, x):
        for lin in self.lins[:-1]:
            x = linear128(x, lin, relu=True)
        return linear128(x, self.lins[-1])


class _Norm(nn.Module):
    def __init__(self, embed_dim: int, kind: str):
        super().__init__()
        self.kind = kind
        if kind == "batch":
            self.normalizer = nn.BatchNorm1d(embed_dim, affine=True)
        elif kind == "instance":
            self.normalizer = nn.InstanceNorm1d(embed_dim, affine=True)
        elif kind == "layer":
            self.normalizer = None
        else:
            raise ValueError(f"unknown normalization {kind!r}")

    def forward(self, x):
        if self.kind == "batch":
            return self.normalizer(x.reshape(-1, x.size(-1))).view(*x.size())
        if self.kind == "instance":
            return self.normalizer(x.permute(0, 2, 1)).permute(0, 2, 1)
        mu = x.mean((1, 2)).view(-1, 1, 1)
        return (x - mu) / torch.sqrt(x.var((1, 2)).view(-1, 1, 1) + 1e-05)


class _Layer(nn.Sequential):
    def __init__(self, embed_dim, num_heads, hidden, normalization):
        super().__init__(
            _Skip(_SelfAttention(embed_dim, num_heads)),
            _Norm(embed_dim, normalization),
            _Skip(_FeedForward(embed_dim, hidden)),
            _Norm(embed_dim, normalization),
        )

    def forward(self, x):
        mha, norm1, ff, norm2 = self[0].module, self[1], self[2].module, self[3]
        fused = (x.is_cuda and not torch.is_grad_enabled() and not self.training
                 and norm1.kind == "batch" and norm2.kind == "batch"
                 and norm1.normalizer.track_running_stats and norm2.normalizer.track_running_stats
                 and _tc_ok(x, mha.out_proj) and _tc_ok(x, ff.lins[-1]) and getattr(self, "fused_epilogue", True))
        if not fused:
            return super().forward(x)
        # inference: skip connection + eval-mode batch norm ride the epilogue of the linear before them
        h = linear128(mha.attend(x), mha.out_proj, residual=x, norm=norm1.normalizer)
        f = h
        for lin in ff.lins[:-1]:
            f = linear128(f, lin, relu=True)
        return linear128(f, ff.lins[-1], residual=h, norm=norm2.normalizer)


class GraphAttentionNetwork(nn.Module):
    def __init__(self, num_heads, embed_dim, num_layers, normalization="batch", feedforward_hidden=512):
        super().__init__()
        self.layers = nn.Sequential(
            *(_Layer(embed_dim, num_heads, feedforward_hidden, normalization) for _ in range(num_layers))
        )

    def forward(self, x, mask=None):
        assert mask is None, "Mask not yet supported!"
        return self.layers(x)


class AttentionModelEncoder(nn.Module):
    def __init__(self, embed_dim: int = 128, init_embedding: nn.Module = None, env_name: str = "tsp",
                 num_heads: int = 8, num_layers: int = 3, normalization: str = "batch",
                 feedforward_hidden: int = 512, net: nn.Module = None, **_):
        super().__init__()
        self.env_name = env_name
        if init_embedding is None:
            if env_name not in INIT_EMBEDDINGS:
                raise ValueError(
                    f"Unknown environment name '{env_name}'. Available init embeddings: {INIT_EMBEDDINGS.keys()}")
            init_embedding = INIT_EMBEDDINGS[env_name](embed_dim)
        self.init_embedding = init_embedding
        self.net = (GraphAttentionNetwork(num_heads, embed_dim, num_layers, normalization, feedforward_hidden)
                    if net is None else net)

    def forward(self, td, mask=None):
        init_h = self.init_embedding(td)
        return self.net(init_h, mask), init_h
