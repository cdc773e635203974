"""AttentionModelPolicy with the reference's forward() surface over the rollout kernels.

Mirrors rl4co/models/zoo/am/policy.py:10-122 (constructor kwargs, state_dict names) and
rl4co/models/common/constructive/base.py:158-276 (forward signature and output dict).  The
encoder runs in PyTorch (encoder.py); everything after it - cache projection (K3), the whole
decode loop with the environment inside (K2), reward and log-likelihood - is two launches
into the C-ABI library per chunk of instances.  There is no step-by-step PyTorch decode path
and no CPU path: unsupported decoding options raise instead of silently degrading.
"""
from __future__ import annotations

import math
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from . import _lib
from .encoder import AttentionModelEncoder
from .ops import calculate_entropy, unbatchify, unbatchify_and_gather

_DECODE_TYPES = ("greedy", "sampling", "multistart_greedy", "multistart_sampling", "evaluate")


class PrecomputedCache(NamedTuple):
    """am/decoder.py:22-41 analogue: what ``return_hidden`` hands back."""
    node_embeddings: torch.Tensor
    graph_context: Optional[torch.Tensor]
    tables: torch.Tensor


class _ContextEmbedding(nn.Module):
    """Parameter holder with the reference's names: project_context [E, step_ctx + E]
    (context.py:47-71) and, for tsp, W_placeholder [2E] ~ U(-1,1) (context.py:111-113)."""

    def __init__(self, env_name: str, embed_dim: int):
        super().__init__()
        in_dim = {"tsp": 2 * embed_dim, "slap": embed_dim}.get(env_name, embed_dim + 1)    # cvrp, op, (s)pctsp: + 1 scalar
        self.project_context = nn.Linear(in_dim, embed_dim, bias=False)
        if env_name == "tsp":
            self.W_placeholder = nn.Parameter(torch.Tensor(2 * embed_dim).uniform_(-1, 1))


class _DynamicEmbedding(nn.Module):
    """Parameter holder of SDVRPDynamicEmbedding (dynamic.py:61-79)."""

    def __init__(self, embed_dim: int):
        super().__init__()
        self.projection = nn.Linear(1, 3 * embed_dim, bias=False)


class _Pointer(nn.Module):
    def __init__(self, embed_dim: int):
        super().__init__()
        self.project_out = nn.Linear(embed_dim, embed_dim, bias=False)


class AttentionModelDecoder(nn.Module):
    """Parameters of am/decoder.py:44-131 (no biases).  The arithmetic lives in
    csrc/am_project.cu and csrc/am_rollout.cu; this module only owns the weights and folds
    them into the table projection matrix."""

    def __init__(self, embed_dim: int = 128, num_heads: int = 8, env_name: str = "tsp",
                 use_graph_context: bool = True, **unused):
        super().__init__()
        if embed_dim != 128 or num_heads != 8:
            raise NotImplementedError("the rollout kernel is built for embed_dim=128, num_heads=8 "
                                      "(the reference defaults, am/policy.py:54-58)")
        if env_name not in _lib.ENV_IDS:
            raise ValueError(f"Unknown environment name '{env_name}'. Available: {list(_lib.ENV_IDS)}")
        self.env_name, self.embed_dim, self.num_heads = env_name, embed_dim, num_heads
        self.context_embedding = _ContextEmbedding(env_name, embed_dim)
        self.pointer = _Pointer(embed_dim)
        self.project_node_embeddings = nn.Linear(embed_dim, 3 * embed_dim, bias=False)
        self.project_fixed_context = nn.Linear(embed_dim, embed_dim, bias=False)
        self.use_graph_context = use_graph_context
        if env_name == "sdvrp":     # SDVRPDynamicEmbedding (nn/env_embeddings/dynamic.py:61-79): Linear(1 -> 3E), no bias
            self.dynamic_embedding = _DynamicEmbedding(embed_dim)

    def folded_weights(self):
        """Cached on the parameters' version counters (an optimiser step invalidates it)."""
        ps = [self.project_node_embeddings.weight, self.pointer.project_out.weight,
              self.context_embedding.project_context.weight, self.project_fixed_context.weight]
        if hasattr(self.context_embedding, "W_placeholder"):
            ps.append(self.context_embedding.W_placeholder)
        if hasattr(self, "dynamic_embedding"):
            ps.append(self.dynamic_embedding.projection.weight)
        def version(p):     # inference tensors (module built under torch.inference_mode) carry no counter
            try:
                return p._version
            except RuntimeError:
                return -1
        key = tuple((p.data_ptr(), version(p), str(p.device)) for p in ps)
        if getattr(self, "_fold_key", None) != key:
            self._fold = self._fold_weights() + (self.project_fixed_context.weight.detach().float().contiguous(),)
            self._fold_key = key
            self._wdyn = None
            if hasattr(self, "dynamic_embedding"):
                # [3][E]: glimpse-key column, glimpse-value column, Wout^T x logit-key column (project_out folded as in
                # table 2); fp32 matmul without TF32
                E = self.embed_dim
                w = self.dynamic_embedding.projection.weight.detach().float()[:, 0]
                prev = torch.backends.cuda.matmul.allow_tf32
                torch.backends.cuda.matmul.allow_tf32 = False
                try:
                    wl = self.pointer.project_out.weight.detach().float().t() @ w[2 * E:]
                finally:
                    torch.backends.cuda.matmul.allow_tf32 = prev
                self._wdyn = torch.stack((w[:E], w[E:2 * E], wl)).contiguous()
            self._image = None
        return self._fold

    def weight_image(self):
        """The folded table weights as the pre-split (hi / lo halves), pre-swizzled shared-memory image
        the tensor-core projection reads (csrc/am_project_tc.cu); packed once per weight update."""
        W = self.folded_weights()[0]
        if getattr(self, "_image", None) is None:
            n_tables = W.shape[0] // 128
            image = torch.empty(int(_lib.load().rl4co_am_weight_image_floats(n_tables)), dtype=torch.float32,
                                device=W.device)
            _lib.call("rl4co_am_pack_weights", n_tables, _lib.ptr(W), _lib.ptr(image), _lib.stream_ptr(W.device))
            self._image = image
        return self._image

    @torch.no_grad()
    def _fold_weights(self):
        """-> (W [n_tables*E, E], q0 [E] or None, wcap [E] or None).
        tables: 0 glimpse K, 1 glimpse V, 2 logit K with project_out folded in
        (<Wout o, l> = <o, Wout^T l>, attention.py:276-282), 3 per-node context (current node
        half of project_context, context.py:67-71), 4 (tsp) first-node half."""
        E = self.embed_dim
        Wn = self.project_node_embeddings.weight.float()
        Wk, Wv, Wl = Wn[:E], Wn[E:2 * E], Wn[2 * E:]
        Wout = self.pointer.project_out.weight.float()
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            parts = [Wk, Wv, Wout.t() @ Wl]
            q0 = wcap = None
            Wc = self.context_embedding.project_context.weight.float()
            if self.env_name == "tsp":
                parts += [Wc[:, E:], Wc[:, :E]]
                q0 = (Wc @ self.context_embedding.W_placeholder.float()).contiguous()
            elif self.env_name in ("cvrp", "sdvrp", "op", "pctsp", "spctsp"):    # [h_cur ; remaining capacity / length / prize]
                parts += [Wc[:, :E]]
                wcap = Wc[:, E].contiguous()
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        return torch.cat(parts, 0).contiguous(), q0, wcap


class AttentionModelPolicy(nn.Module):
    """Kool et al. attention model; constructor follows am/policy.py:54-83."""

    def __init__(self, encoder: nn.Module = None, decoder: nn.Module = None, embed_dim: int = 128,
                 num_encoder_layers: int = 3, num_heads: int = 8, normalization: str = "batch",
                 feedforward_hidden: int = 512, env_name: str = "tsp", encoder_network: nn.Module = None,
                 init_embedding: nn.Module = None, context_embedding: nn.Module = None,
                 dynamic_embedding: nn.Module = None, use_graph_context: bool = True,
                 linear_bias_decoder: bool = False, mask_inner: bool = True,
                 out_bias_pointer_attn: bool = False, check_nan: bool = True, temperature: float = 1.0,
                 tanh_clipping: float = 10.0, mask_logits: bool = True,
                 train_decode_type: str = "sampling", val_decode_type: str = "greedy",
                 test_decode_type: str = "greedy", moe_kwargs: dict = None,
                 rollout_chunk: int = 16384, **unused_kwargs):
        super().__init__()
        for name, bad in (("context_embedding", context_embedding is not None),
                          ("dynamic_embedding", dynamic_embedding is not None),
                          ("linear_bias_decoder", linear_bias_decoder),
                          ("out_bias_pointer_attn", out_bias_pointer_attn),
                          ("mask_inner=False", not mask_inner),
                          ("moe_kwargs", bool(moe_kwargs) and any(v for v in moe_kwargs.values()))):
            if bad:
                raise NotImplementedError(
                    f"{name}: the fused rollout kernel implements the stock tsp / cvrp / slap context "
                    "and static dynamic-embedding only (SURVEY.md 8b)")
        self.env_name = env_name
        self.encoder = encoder if encoder is not None else AttentionModelEncoder(
            embed_dim=embed_dim, num_heads=num_heads, num_layers=num_encoder_layers, env_name=env_name,
            normalization=normalization, feedforward_hidden=feedforward_hidden, net=encoder_network,
            init_embedding=init_embedding)
        self.decoder = decoder if decoder is not None else AttentionModelDecoder(
            embed_dim=embed_dim, num_heads=num_heads, env_name=env_name,
            use_graph_context=use_graph_context)
        self.temperature, self.tanh_clipping, self.mask_logits = temperature, tanh_clipping, mask_logits
        self.train_decode_type = train_decode_type
        self.val_decode_type = val_decode_type
        self.test_decode_type = test_decode_type
        self.rollout_chunk = rollout_chunk
        self.projection = "tensor"      # "tensor" | "simt": which K3 kernel builds the rollout tables
        self.unused_kwargs = unused_kwargs
        self.last_launches = 0  # rl4co_* kernel launches of the most recent forward()
        # tsp single start: project 4 tables and let the rollout kernel form table 4's one row ("auto": from 48 nodes on -
        # the row costs the kernel ~1.4 k clk per rollout, the table costs the projection ~36 clk x SMs per node; "always",
        # "never")
        self.first_context_in_kernel = "auto"
        self.profile = None     # set to a list to collect (rollouts, [ev0, ev1, ev2]) per chunk

    # ------------------------------------------------------------------------------
    def forward(self, td, env=None, phase: str = "train", calc_reward: bool = True,
                return_actions: bool = False, return_entropy: bool = False, return_hidden: bool = False,
                return_init_embeds: bool = False, return_sum_log_likelihood: bool = True,
                actions=None, max_steps=1_000_000, hidden: torch.Tensor = None, **decoding_kwargs) -> dict:
        """constructive/base.py:158-276.  ``hidden`` (extension): precomputed encoder output
        [B,N,E]; skips the encoder (used by parity tests and decode-only timing)."""
        if isinstance(env, str) or env is None:
            from .envs import get_env
            env = get_env(self.env_name if env is None else env)
        decode_type = decoding_kwargs.pop("decode_type", None)
        if actions is not None:
            decode_type = "evaluate"
        elif decode_type is None:
            decode_type = getattr(self, f"{phase}_decode_type")
        if decode_type not in _DECODE_TYPES:
            raise NotImplementedError(f"decode_type {decode_type!r}: the fused kernel implements {_DECODE_TYPES}")
        temperature = float(decoding_kwargs.pop("temperature", self.temperature))
        tanh_clipping = float(decoding_kwargs.pop("tanh_clipping", self.tanh_clipping))
        mask_logits = bool(decoding_kwargs.pop("mask_logits", self.mask_logits))
        store_all_logp = bool(decoding_kwargs.pop("store_all_logp", return_entropy))
        num_starts = decoding_kwargs.pop("num_starts", None)
        select_best = bool(decoding_kwargs.pop("select_best", False))
        for k in ("top_k", "top_p"):
            if decoding_kwargs.pop(k, 0):
                raise NotImplementedError(f"{k} > 0 is not implemented by the fused rollout kernel")
        # multisample (utils/decoding.py:271-308): the batch is tiled num_starts times and every copy is an independent
        # rollout of the decode type - no forced first action.  select_start_nodes_fn is only consulted by MULTISTART
        # decoding in the reference (:286-291); with multisample it is accepted and unused, exactly as there.
        multisample = bool(decoding_kwargs.pop("multisample", False))
        start_fn = decoding_kwargs.pop("select_start_nodes_fn", None)
        if decoding_kwargs.pop("improvement_method_mode", None):
            raise NotImplementedError("improvement_method_mode is not implemented by the fused rollout kernel")
        decoding_kwargs.pop("softmax_temp", None)       # tasks/eval.py:191 passes it; the reference's strategy ignores it
        decoding_kwargs.pop("key", None)
        # extension (CUDA-graph capture, GraphedRollout below): no host read of the status word inside forward(); it is
        # returned as out["status"] and checked by the caller after the replay
        defer_status = bool(decoding_kwargs.pop("defer_status", False))
        if decoding_kwargs:
            raise TypeError(f"unknown decoding kwargs: {sorted(decoding_kwargs)}")

        locs = td["locs"]
        if not locs.is_cuda:
            raise RuntimeError("AttentionModelPolicy.forward needs CUDA tensors (no CPU fallback)")
        dev, (B, N, _) = locs.device, locs.shape
        name = env.name
        if name != self.decoder.env_name:
            raise ValueError(f"policy was built for {self.decoder.env_name!r}, env is {name!r}")
        multistart = "multistart" in decode_type
        if multisample and multistart:
            multisample = False                         # decoding.py:286: the multistart branch wins
        S = 0
        if multistart or multisample:  # utils/decoding.py:265-284
            S = int(num_starts) if num_starts is not None else env.get_num_starts(td)
            if S < 1:
                multistart, multisample, S = False, False, 0
        mode = {"greedy": _lib.MODE_GREEDY, "sampling": _lib.MODE_SAMPLING,
                "evaluate": _lib.MODE_EVALUATE}[decode_type.replace("multistart_", "")]
        Sx = max(S, 1)
        R = B * Sx
        P = O = K = 0
        if name == "tsp":
            T = N
        elif name == "slap":
            P, O, K = td["freq"].shape[-2], td["picklist"].shape[1], td["picklist"].shape[2]
            T = P
        elif name == "sdvrp":
            # a customer visit either completes the customer or fills the vehicle, every fill costs one depot visit
            T = 3 * N
            if store_all_logp:
                raise NotImplementedError("sdvrp: per-node log-probs (return_entropy / store_all_logp) are not "
                                          "produced by the register-file rollout kernel")
        elif name in ("op", "pctsp", "spctsp"):
            # op at worst: depot first (not done at step 0), ..., every customer, depot; pctsp: every customer, depot
            T = N + 1 if name == "op" else N
            if store_all_logp:
                raise NotImplementedError(f"{name}: per-node log-probs (return_entropy / store_all_logp) are not "
                                          "produced by the register-file rollout kernel")
        else:  # every customer may be followed by a depot return
            T = 2 * (N - 1) + 1
        # multistart start nodes (utils/decoding.py:286-291): a caller-supplied select_start_nodes_fn(td, env, num_starts),
        # or - op with fewer feasible nodes than starts in some instance - the reference's own resampling of the starts
        # from the feasible nodes (utils/ops.py:166-176, torch.multinomial with replacement).  Otherwise the kernel's
        # built-in rule s % num_loc (+1 past a depot), which is the rest of utils/ops.py:139-163.
        start_nodes = None
        if multistart and start_fn is not None:
            start_nodes = start_fn(td, env, S)
        elif multistart and name == "op" and bool((td["action_mask"][..., 1:].float().sum(-1) < S).any()):
            sel = torch.multinomial(td["action_mask"][..., 1:].float(), S, replacement=True) + 1     # [B,S], depot re-added
            start_nodes = sel.t().reshape(-1)                                                        # "b n -> (n b)"
        if start_nodes is not None:
            start_nodes = torch.as_tensor(start_nodes, device=dev).reshape(-1)
            if start_nodes.numel() != R:
                raise ValueError(f"select_start_nodes_fn returned {start_nodes.numel()} nodes for {R} rollouts")
            if bool(((start_nodes < 0) | (start_nodes >= N)).any()):
                raise ValueError("select_start_nodes_fn returned a node outside [0, N)")
            start_nodes = start_nodes.to(torch.int32).contiguous()
        if actions is not None:
            T = actions.shape[-1]
            actions = actions.contiguous()
        T = min(T, int(max_steps) + 1)

        out_actions = torch.empty(R, T, dtype=torch.int64, device=dev)
        out_logp = torch.empty(R, T, dtype=torch.float32, device=dev)
        out_ll = torch.empty(R, dtype=torch.float32, device=dev)
        out_reward = torch.empty(R, dtype=torch.float32, device=dev)
        out_steps = torch.empty(R, dtype=torch.int32, device=dev)
        logp_all = torch.empty(R, T, N, dtype=torch.float32, device=dev) if store_all_logp else None
        assign = torch.empty(R, P, dtype=torch.int32, device=dev) if name == "slap" else None
        status = _lib.new_status(dev)
        W, q0, wcap, Wfixed = self.decoder.folded_weights()
        W_image = self.decoder.weight_image() if self.projection == "tensor" else None
        n_tables = W.shape[0] // 128
        # tsp, single start: a trajectory reads one row of table 4 (its first node's), so the projection leaves that table
        # out and the rollout kernel forms the row from h and the table's weight block at step 1 (include/: args h / w_first)
        w_first = None
        fck = self.first_context_in_kernel
        # "auto": where the compaction kernel (am_rollout_tsp.cu) will run - the register-file kernel does not take h / w_first
        # and the shared-memory kernel is the slow fallback
        auto = (fck == "auto" and 48 <= N <= 112 and mode != _lib.MODE_EVALUATE and not store_all_logp
                and tanh_clipping > 0 and temperature == 1.0 and mask_logits)
        if name == "tsp" and S == 0 and not return_hidden and (fck == "always" or auto):
            w_first = W[4 * 128:5 * 128]
            n_tables = 4
        prof = self.profile  # optional CUDA-event timing of the K3 / K2 launches (bench.py)
        seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if mode == _lib.MODE_SAMPLING else 0
        stream = _lib.stream_ptr(dev)
        hiddens, inits, gctxs, tabs = [], [], [], []
        launches = 0

        chunk = B if S > 0 else min(B, max(1, int(self.rollout_chunk)))
        # train-mode batch norm takes its statistics over the batch it is given: a chunked encoder pass would normalise
        # every chunk differently (and update the running statistics once per chunk), so in that case the encoder sees
        # the whole batch once and only the projection / rollout kernels are chunked
        whole_init = None
        if hidden is None and chunk < B and self.encoder.training and any(
                isinstance(m, nn.modules.batchnorm._BatchNorm) for m in self.encoder.modules()):
            hidden, whole_init = self.encoder(td)
        for c0 in range(0, B, chunk):
            c1 = min(B, c0 + chunk)
            Bc = c1 - c0
            if hidden is not None:
                h, init_h = hidden[c0:c1], (whole_init[c0:c1] if whole_init is not None else None)
            else:
                h, init_h = self.encoder(td[c0:c1] if (c0, c1) != (0, B) else td)
            h = h.detach().float().contiguous()
            tables = torch.empty(n_tables, Bc * N, 128, dtype=torch.float32, device=dev)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if prof is not None else None
            if ev:
                ev[0].record()
            if self.projection == "tensor":     # tcgen05, 3-term FP16 split with range scales, fp32-accurate
                _lib.call("rl4co_am_project_tc", Bc * N, n_tables, _lib.ptr(h), _lib.ptr(W_image), _lib.ptr(tables), stream)
            else:                               # fp32 SIMT kernel (kept as the in-library cross-check)
                _lib.call("rl4co_am_project", Bc * N, n_tables, _lib.ptr(h), _lib.ptr(W), _lib.ptr(tables), stream)
            launches += 1
            gctx = None
            if self.decoder.use_graph_context:
                gctx = torch.empty(Bc, 128, dtype=torch.float32, device=dev)
                _lib.call("rl4co_am_graph_context", Bc, N, _lib.ptr(h), _lib.ptr(Wfixed), _lib.ptr(gctx), stream)
                launches += 1
            if ev:
                ev[1].record()
            a = _lib.AmRolloutArgs()
            a.env, a.mode, a.B, a.S, a.multistart = _lib.ENV_IDS[name], mode, Bc, S, int(multistart)
            a.num_loc = int(getattr(env.generator, "num_loc", N - (0 if name == "tsp" else 1)))
            a.N, a.T, a.P, a.O, a.K, a.n_tables = N, T, P, O, K, n_tables
            a.temperature, a.tanh_clipping, a.mask_logits = temperature, tanh_clipping, int(mask_logits)
            a.tables, a.gctx = _lib.ptr(tables), _lib.ptr(gctx)
            a.q0, a.wcap = _lib.ptr(q0), _lib.ptr(wcap)
            a.locs = _lib.ptr(locs[c0:c1].contiguous(), torch.float32)
            if name in ("cvrp", "sdvrp"):
                a.demand = _lib.ptr(td["demand"][c0:c1].contiguous(), torch.float32)
                a.vehicle_capacity = _lib.ptr(td["vehicle_capacity"][c0:c1].contiguous(), torch.float32)
            if name == "sdvrp":
                a.wdyn = _lib.ptr(self.decoder._wdyn)
            if name == "op":
                a.prize = _lib.ptr(td["prize"][c0:c1].contiguous(), torch.float32)
                a.max_length = _lib.ptr(td["max_length"][c0:c1].contiguous(), torch.float32)
            if name in ("pctsp", "spctsp"):
                a.prize = _lib.ptr(td["real_prize"][c0:c1].contiguous(), torch.float32)
                a.penalty = _lib.ptr(td["penalty"][c0:c1].contiguous(), torch.float32)
                a.prize_required = _lib.ptr(td["prize_required"][c0:c1].contiguous(), torch.float32)
            if name == "slap":
                a.picklist = _lib.ptr(td["picklist"][c0:c1].contiguous(), torch.int64)
            # single start: rollouts of the chunk are rows c0..c1; multistart / multisample run unchunked
            sl = slice(c0, c1) if S == 0 else slice(0, R)
            a.actions_in = _lib.ptr(actions[sl]) if actions is not None else None
            a.seed, a.offset = seed, c0 * T
            a.actions, a.logp = _lib.ptr(out_actions[sl]), _lib.ptr(out_logp[sl])
            a.logp_all = _lib.ptr(logp_all[sl]) if logp_all is not None else None
            a.ll, a.reward, a.steps = _lib.ptr(out_ll[sl]), _lib.ptr(out_reward[sl]), _lib.ptr(out_steps[sl])
            a.assignment = _lib.ptr(assign[sl]) if assign is not None else None
            a.start_nodes = _lib.ptr(start_nodes, torch.int32) if start_nodes is not None else None
            if w_first is not None:
                a.h, a.w_first = _lib.ptr(h), _lib.ptr(w_first)
            a.status = _lib.ptr(status)
            _lib.check(_lib.load().rl4co_am_rollout(a, stream), "rl4co_am_rollout")
            launches += 1
            if ev:
                ev[2].record()
                prof.append((Bc * Sx, ev))
            if return_hidden:
                hiddens.append(h); gctxs.append(gctx); tabs.append(tables)
            if return_init_embeds and init_h is not None:
                inits.append(init_h)
        self.last_launches = launches

        # ---- the reference's output construction (base.py:253-276) --------------------
        if name in ("cvrp", "sdvrp", "op", "pctsp", "spctsp") and actions is None:
            # the reference loops until the slowest episode is done (base.py:230): T = max steps
            Tmax = int(out_steps.max().item())
            out_actions, out_logp = out_actions[:, :Tmax], out_logp[:, :Tmax]
            if logp_all is not None:
                logp_all = logp_all[:, :Tmax]
        if calc_reward and env.check_solution and mode == _lib.MODE_EVALUATE and name != "slap":
            env.check_solution_validity(_batchified_view(td, Sx), out_actions)
        if not defer_status:
            _lib.raise_for_status(status)
        logprobs = logp_all if store_all_logp else out_logp
        reward = out_reward
        if S > 0 and select_best:  # utils/decoding.py:399-407
            _, max_idxs = unbatchify(reward, S).max(dim=-1)
            out_actions = unbatchify_and_gather(out_actions, max_idxs, S)
            logprobs = unbatchify_and_gather(logprobs, max_idxs, S)
            out_logp = unbatchify_and_gather(out_logp, max_idxs, S)
            reward = unbatchify_and_gather(reward, max_idxs, S)
            out_ll = out_logp.sum(1)
        if not calc_reward:
            reward = torch.zeros(reward.shape, dtype=torch.bool, device=dev)
        outdict = {"reward": reward,
                   "log_likelihood": out_ll if return_sum_log_likelihood else out_logp}
        if return_actions:
            outdict["actions"] = out_actions
        if return_entropy:
            outdict["entropy"] = calculate_entropy(logprobs)
        if return_hidden:
            outdict["hidden"] = PrecomputedCache(torch.cat(hiddens), None if gctxs[0] is None else torch.cat(gctxs),
                                                 tabs[0] if len(tabs) == 1 else torch.cat(tabs, 1))
        if return_init_embeds:
            outdict["init_embeds"] = torch.cat(inits) if inits else None
        if name == "slap":
            outdict["assignment"] = assign
        if defer_status:
            outdict["status"] = status
        return outdict

    def graphed(self, td, env, **kwargs) -> "GraphedRollout":
        """Capture ``self(td, env, **kwargs)`` - encoder, projection, rollout - as ONE CUDA graph for inputs of this shape."""
        return GraphedRollout(self, td, env, **kwargs)


class GraphedRollout:
    """A rollout for one input shape, captured as a CUDA graph: small batches are launch bound (TSP50, B = 512:
    ~27 kernel launches and their Python dispatch cost 2 ms, the kernels ~0.3 ms), a replay costs one launch.

        fast = policy.graphed(td, env, decode_type="greedy", return_actions=True)
        out = fast(new_td)           # same keys as policy(new_td, env, ...); tensors are REUSED by the next call

    Works for what runs without a host read in ``forward``: tsp / slap, greedy or evaluate, eval-mode encoder.  Sampling
    draws its Philox seed on the host and the depot environments trim the tours to the longest one (a device -> host
    read), so those raise here.  The device status word (invalid action, NaN logits, ...) is checked after each replay and
    raises the reference's assertion texts like the eager path."""

    def __init__(self, policy: "AttentionModelPolicy", td, env, **kwargs):
        if isinstance(env, str) or env is None:
            from .envs import get_env
            env = get_env(policy.env_name if env is None else env)
        decode_type = kwargs.get("decode_type", policy.test_decode_type)
        if env.name not in ("tsp", "slap") or "sampling" in decode_type:
            raise NotImplementedError("GraphedRollout: tsp / slap with greedy or evaluate decoding only (sampling seeds and "
                                      "depot-environment tour trimming need a host read per call)")
        if policy.training:
            raise RuntimeError("GraphedRollout captures inference: call policy.eval() first")
        self.policy, self.env, self.kwargs = policy, env, dict(kwargs, defer_status=True)
        self.static_td = td.clone()
        with torch.inference_mode():
            side = torch.cuda.Stream(device=td["locs"].device)
            side.wait_stream(torch.cuda.current_stream(td["locs"].device))
            with torch.cuda.stream(side):               # warm-up off the capture: lazy initialisation, weight images
                for _ in range(2):
                    policy(self.static_td, env, **self.kwargs)
            torch.cuda.current_stream(td["locs"].device).wait_stream(side)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.out = policy(self.static_td, env, **self.kwargs)
        self.status = self.out.pop("status")

    def __call__(self, td, actions: Optional[torch.Tensor] = None) -> dict:
        if actions is not None:                         # evaluate mode: the tours to score
            self.kwargs["actions"].copy_(actions, non_blocking=True)
        for key, value in td.items():                   # new instance data into the captured buffers
            if key in self.static_td.keys() and isinstance(value, torch.Tensor):
                dst = self.static_td[key]
                if dst.shape != value.shape:
                    raise ValueError(f"GraphedRollout was captured for {key} {tuple(dst.shape)}, got {tuple(value.shape)}")
                if dst.data_ptr() != value.data_ptr():
                    dst.copy_(value, non_blocking=True)
        self.graph.replay()
        _lib.raise_for_status(self.status)
        self.status.zero_()
        return self.out


def _batchified_view(td, S: int):
    if S <= 1:
        return td
    from .ops import batchify
    return batchify(td, S)


class POMOPolicy(AttentionModelPolicy):
    """The policy POMO instantiates (zoo/pomo/model.py:58-68): 6 encoder layers, instance
    normalisation, no graph context, multistart decode types."""

    def __init__(self, env_name: str = "tsp", **kw):
        kw.setdefault("num_encoder_layers", 6)
        kw.setdefault("normalization", "instance")
        kw.setdefault("use_graph_context", False)
        kw.setdefault("train_decode_type", "multistart_sampling")
        kw.setdefault("val_decode_type", "multistart_greedy")
        kw.setdefault("test_decode_type", "multistart_greedy")
        super().__init__(env_name=env_name, **kw)
