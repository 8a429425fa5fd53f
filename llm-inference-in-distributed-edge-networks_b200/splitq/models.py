"""Layer-wise model wrappers with the reference's call surface, running the split-layer
boundary through the CUDA path.

Reference: `Pythia70Model` (Experiments/Pythia-70M/pythia_model.py:13-206) and
`QwenPointFiveBModel` (Experiments/Qwen2-0.5B/qwen_layer_wise.py:5-167).  Same constructor
arguments, attributes (`.model .tokenizer .device .num_layers .gpt_neox/.qwen .final_layer_norm
.rotary_emb .ratios`) and method signatures; `activation_quantization` returns the 0-dim NLL.

What differs, by design:
  * the in-place fake quant at `i == layer_of_interest` (pythia_model.py:167-188,
    qwen_layer_wise.py:54-70) becomes select -> quantise/pack -> transfer -> dequantise, the
    last step fused with the first norm of layer i+1 (or the final norm);
  * the two halves may live on two devices (`split_devices=(up, down)`); each device holds a full
    copy of the (small) model so that any `layer_of_interest` can be evaluated without moving
    modules, and the payload is what crosses NVLink;
  * `from_config(...)` builds the architecture with seeded random weights (no network here).
HF decoder layers, rotary embedding, LM head and the loss stay stock PyTorch/HF.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch

from . import _native as N
from . import boundary, compat


class _LayerWiseModel:
    family = ""
    hub_name = ""
    default_cfg: dict = {}

    # ---- construction ---------------------------------------------------------------------
    def _setup(self, device, model, tokenizer=None, split_devices=None, scale_mode="group", bits=4,
               fuse_norm=True, ladder_fracs=None):
        self.device = device
        self.tokenizer = tokenizer
        self.model = model.to(device).eval()
        self.scale_mode = N.SCALE_GROUP if scale_mode in ("group", N.SCALE_GROUP) else N.SCALE_TOKEN
        self.bits = bits
        self.ladder_fracs = ladder_fracs
        self.fuse_norm = fuse_norm
        self.num_layers = len(self.model.base_model.layers)
        self._bind(self.model)
        self.down_device = torch.device(device)
        self._down = self.model
        if split_devices is not None:
            up, down = (torch.device(d) for d in split_devices)
            if up != torch.device(device):
                raise ValueError("split_devices[0] must equal `device` (the upstream half)")
            if down != up:
                import copy
                self._down = copy.deepcopy(self.model).to(down).eval()
                self.down_device = down
        self.last_payload_bytes = 0

    @classmethod
    def from_config(cls, device, *args, cfg: Optional[dict] = None, seed: int = 0, **kw):
        """Offline constructor: same architecture, seeded random weights."""
        self = object.__new__(cls)
        model = compat.build_causal_lm(cls.family, cfg or cls.default_cfg, seed=seed, attn_implementation="sdpa")
        self._init_extra(*args)
        self._setup(device, model, **kw)
        return self

    @classmethod
    def from_model(cls, device, model, *args, **kw):
        self = object.__new__(cls)
        self._init_extra(*args)
        self._setup(device, model, **kw)
        return self

    def _from_pretrained(self, device, **kw):
        from transformers import AutoModelForCausalLM, AutoTokenizer
        try:
            model = AutoModelForCausalLM.from_pretrained(self.hub_name, attn_implementation="sdpa")
            tokenizer = AutoTokenizer.from_pretrained(self.hub_name)
        except Exception as e:  # no network / no cache
            raise RuntimeError(f"cannot load {self.hub_name} ({e}); offline use "
                               f"{type(self).__name__}.from_config(device, ...)") from e
        self._setup(device, model, tokenizer, **kw)

    def _init_extra(self, *args):
        pass

    # family hooks -------------------------------------------------------------------------
    def _bind(self, model):
        raise NotImplementedError

    def _embed(self, model, ids):
        raise NotImplementedError

    def _lm_head(self, model):
        raise NotImplementedError

    def _final_norm(self, model):
        raise NotImplementedError

    def _run_prenormed_layer(self, layer, resid, normed, normed2, pos):
        raise NotImplementedError

    def _first_norm_params(self, layer):
        raise NotImplementedError

    # ---- shared pieces ----------------------------------------------------------------------
    def move_to_device(self):  # pythia_model.py:147-148
        self.model.to(self.device)

    def remove_from_device(self):  # pythia_model.py:150-151
        self.model.to("cpu")

    def _positions(self, model, hidden, ids):
        position_ids = torch.arange(ids.size(1), dtype=torch.long, device=ids.device).unsqueeze(0).expand(
            ids.size(0), -1)
        return model.base_model.rotary_emb(hidden, position_ids)

    def last_portion(self, post_norm, target, model=None, per_sequence=False):
        """LM head + shifted cross entropy (qwen_layer_wise.py:28-40, pythia_model.py:193-203).
        per_sequence=True returns one mean NLL per batch row (used when several (method, ratio)
        variants of the same window are pushed through the downstream half as one batch)."""
        model = model or self._down
        logits = self._lm_head(model)(post_norm)
        target = target[:, 1:].to(logits.device)
        logits = logits[:, :-1, :]
        if not per_sequence:
            return torch.nn.functional.cross_entropy(logits.reshape(-1, logits.size(-1)), target.reshape(-1),
                                                     ignore_index=-100)
        tok = torch.nn.functional.cross_entropy(logits.reshape(-1, logits.size(-1)), target.reshape(-1),
                                                ignore_index=-100, reduction="none").view(target.shape)
        valid = (target != -100).sum(dim=1).clamp_min(1)
        return tok.sum(dim=1) / valid

    def ladder(self, ratio: float, S: int):
        if self.ladder_fracs is not None:  # extension: fixed multi-width ladder scaled by ratio
            return boundary.ladder_from_fractions([(f * ratio, b) for f, b in self.ladder_fracs], S)
        return boundary.ladder_for_ratio(ratio, S, self.bits)

    def upstream(self, batched_input_tokens, layer_of_interest):
        """Embedding + layers 0..layer_of_interest on the upstream device -> (B,S,D) fp32."""
        ids = batched_input_tokens.to(self.device)
        hidden = self._embed(self.model, ids)
        pos = self._positions(self.model, hidden, ids)
        layers = compat.decoder_layers(self.model.base_model)
        for i in range(layer_of_interest + 1):
            hidden = compat.layer_output(layers[i](hidden, position_embeddings=pos))
        return hidden

    def boundary_send(self, hidden, importance_values, ratio):
        """select -> quantise -> pack on the upstream device.  Returns (payload, S, D)."""
        B, S, D = hidden.shape
        imp = importance_values.to(hidden.device, torch.float32)
        if imp.dim() == 1:
            imp = imp.unsqueeze(0).expand(B, S)
        ladder = self.ladder(ratio, S)
        nbytes = boundary.payload_bytes(S, D, ladder, self.scale_mode)
        bits, row_off = boundary.select_bits(imp, ladder, D)
        payload = boundary.quant_pack(hidden, bits, row_off, self.scale_mode, nbytes)
        self.last_payload_bytes = B * nbytes
        return payload

    def payload_like(self, rows: int, S: int, ratio: float) -> torch.Tensor:
        """Receive buffer for what boundary_send(hidden (rows,S,D), ., ratio) produces on the other
        side of a GPU pair: the size depends only on (S, D, ladder, scale mode)."""
        D = self._down.config.hidden_size
        nbytes = boundary.payload_bytes(S, D, self.ladder(ratio, S), self.scale_mode)
        return torch.empty(rows, nbytes, dtype=torch.uint8, device=self.down_device)

    def hidden_like(self, rows: int, S: int) -> torch.Tensor:
        return torch.empty(rows, S, self._down.config.hidden_size, dtype=torch.float32, device=self.down_device)

    def boundary_hop(self, payload):
        if self.down_device == payload.device:
            return payload
        dst = torch.empty(payload.shape, dtype=torch.uint8, device=self.down_device)
        # `dst` comes from the downstream device's caching allocator and may be a block that an
        # earlier, still queued dequantise on that device reads; the copy is enqueued on the UPSTREAM
        # device's stream, which knows nothing of that work -- so order it behind it explicitly
        up, down = torch.cuda.current_stream(payload.device), torch.cuda.current_stream(self.down_device)
        free = torch.cuda.Event()
        free.record(down)
        up.wait_event(free)
        boundary.transfer(dst, payload)
        ev = torch.cuda.Event()
        ev.record(up)
        down.wait_event(ev)
        return dst

    def downstream(self, ids_shape, target, layer_of_interest, payload=None, hidden=None, per_sequence=False):
        """Layers layer_of_interest+1.. + final norm + loss on the downstream device, starting
        either from a payload (dequantise fused into the first norm) or from raw hidden states.
        `payload` may be a list of payloads of different widths (one per ratio); their rows are
        dequantised separately and run through the layers as ONE batch of ids_shape[0] rows."""
        model = self._down
        dev = self.down_device
        B, S = ids_shape
        D = model.config.hidden_size
        layers = compat.decoder_layers(model.base_model)
        ids_stub = torch.empty(B, S, dtype=torch.long, device=dev)
        nxt = layer_of_interest + 1
        payloads = None if payload is None else (list(payload) if isinstance(payload, (list, tuple)) else [payload])
        cat = lambda ts: ts[0] if len(ts) == 1 else torch.cat(ts, dim=0)  # noqa: E731
        if payloads is not None:
            assert sum(p.shape[0] for p in payloads) == B, "ids_shape[0] must count the rows of all payloads"
        if payloads is not None and self.fuse_norm:
            if nxt < self.num_layers:
                kind, g1, b1, eps, g2, b2 = self._first_norm_params(layers[nxt])
                outs = [boundary.dequant_norm(p, S, D, kind, g1, b1, eps, g2, b2) for p in payloads]
                resid, n1 = cat([o[0] for o in outs]), cat([o[1] for o in outs])
                n2 = cat([o[2] for o in outs]) if outs[0][2] is not None else None
                pos = self._positions(model, resid, ids_stub)
                hidden = self._run_prenormed_layer(layers[nxt], resid, n1, n2, pos)
                nxt += 1
            else:
                fn = self._final_norm(model)
                kind = N.NORM_LAYERNORM if isinstance(fn, torch.nn.LayerNorm) else N.NORM_RMSNORM
                eps = fn.eps if kind == N.NORM_LAYERNORM else fn.variance_epsilon
                post_norm = cat([boundary.dequant_norm(p, S, D, kind, fn.weight, getattr(fn, "bias", None), eps,
                                                       want_resid=False)[1] for p in payloads])
                return self.last_portion(post_norm, target, model, per_sequence)
        elif payloads is not None:
            hidden = cat([boundary.dequant(p, S, D) for p in payloads])
        else:
            hidden = hidden.to(dev)
        if nxt < self.num_layers:
            pos = self._positions(model, hidden, ids_stub)
            for i in range(nxt, self.num_layers):
                hidden = compat.layer_output(layers[i](hidden, position_embeddings=pos))
        post_norm = self._final_norm(model)(hidden)
        return self.last_portion(post_norm, target, model, per_sequence)

    # ---- reference API ------------------------------------------------------------------------
    def activation_quantization(self, batched_input_tokens, target, importance_values, ratio, layer_of_interest):
        """Same contract as pythia_model.py:153-206 / qwen_layer_wise.py:42-76."""
        with torch.no_grad():
            hidden = self.upstream(batched_input_tokens, layer_of_interest)
            shape = tuple(batched_input_tokens.shape)
            if ratio > 0:
                payload = self.boundary_hop(self.boundary_send(hidden, importance_values, ratio))
                return self.downstream(shape, target, layer_of_interest, payload=payload)
            return self.downstream(shape, target, layer_of_interest, hidden=hidden)

    def layer_by_layer_impl(self, batched_input_tokens, target):
        """No-quant baseline (qwen_layer_wise.py:78-104)."""
        with torch.no_grad():
            hidden = self.upstream(batched_input_tokens, self.num_layers - 1)
            return self.downstream(tuple(batched_input_tokens.shape), target, self.num_layers - 1, hidden=hidden)


class Pythia70Model(_LayerWiseModel):
    """Reference: Experiments/Pythia-70M/pythia_model.py:13 (`Pythia70Model(device, ratios)`)."""
    family, hub_name, default_cfg = "gpt_neox", "EleutherAI/pythia-70m", compat.PYTHIA_70M

    def __init__(self, device, ratios, **kw) -> None:
        self.ratios = ratios
        self._from_pretrained(device, **kw)

    def _init_extra(self, ratios=None):
        self.ratios = ratios

    def _bind(self, model):
        self.gpt_neox = model.base_model
        self.final_layer_norm = model.gpt_neox.final_layer_norm
        self.rotary_emb = self.gpt_neox.rotary_emb

    def _embed(self, model, ids):
        return model.base_model.embed_in(ids)

    def _lm_head(self, model):
        return model.embed_out

    def _final_norm(self, model):
        return model.gpt_neox.final_layer_norm

    def _first_norm_params(self, layer):
        ln1, ln2 = layer.input_layernorm, layer.post_attention_layernorm
        if layer.use_parallel_residual:
            return N.NORM_LAYERNORM, ln1.weight, ln1.bias, ln1.eps, ln2.weight, ln2.bias
        return N.NORM_LAYERNORM, ln1.weight, ln1.bias, ln1.eps, None, None

    def _run_prenormed_layer(self, layer, resid, normed, normed2, pos):
        # GPTNeoXLayer.forward with input_layernorm (and, for parallel residual, the second
        # LayerNorm) already applied by the fused dequant kernel
        attn_out, _ = layer.attention(normed, attention_mask=None, position_embeddings=pos)
        attn_out = layer.post_attention_dropout(attn_out)
        if layer.use_parallel_residual:
            mlp_out = layer.post_mlp_dropout(layer.mlp(normed2))
            return mlp_out + attn_out + resid
        attn_out = attn_out + resid
        mlp_out = layer.post_mlp_dropout(layer.mlp(layer.post_attention_layernorm(attn_out)))
        return mlp_out + attn_out

    # ---- the first experiment's batched entry points (initial_exp.py:112-131) --------------------
    # The batch holds one copy of the window per entry of self.ratios and the RATIO VALUE is the batch
    # row it acts on (pythia_model.py:59, `hidden_states[ratio, token, :]`), so the ratios are the
    # integers 0 .. B-1.  In the reference both methods die in `simulate_quantization` (a @staticmethod
    # declared with `self`, pythia_model.py:57 vs :112,:143), whose per-token affine qint8 would leave
    # two levels per token; the token-set rules are rebuilt here as intended, over this library's
    # quantiser (self.bits, self.scale_mode) and the real boundary (pack -> hop -> dequantise).
    def _batched_boundary(self, batched_input_tokens, target, quant_layer, select):
        with torch.no_grad():
            hidden = self.upstream(batched_input_tokens, quant_layer)
            B, S, D = hidden.shape
            target = target.to(hidden.device)
            if target.size(0) != B:
                target = target.expand(B, -1)
            shape = (B, S)
            if select is None:
                nll = self.downstream(shape, target, quant_layer, hidden=hidden, per_sequence=True)
            else:
                bits, row_off = select(B, S, D)
                nbytes = boundary.payload_bytes(S, D, [], self.scale_mode)  # rows differ per window: size for all-raw
                payload = boundary.quant_pack(hidden, bits, row_off, self.scale_mode, nbytes)
                self.last_payload_bytes = B * nbytes
                nll = self.downstream(shape, target, quant_layer, payload=self.boundary_hop(payload), per_sequence=True)
            return [nll[int(r)] for r in self.ratios]

    def batched_quantization_helper(self, batched_input_tokens, ordering, target, quant_layer=2, do_quant=True):
        """pythia_model.py:116-145: batch row `ratio` has the first int(0.1 * ratio * len(ordering)) tokens
        of `ordering` (ascending importance, from extract_attentions) quantised after layer
        `quant_layer`; returns one NLL per entry of self.ratios."""
        if not do_quant:
            return self._batched_boundary(batched_input_tokens, target, quant_layer, None)
        order = torch.as_tensor(ordering, dtype=torch.long)
        n = order.numel()

        def select(B, S, D):
            rank = torch.empty(S, dtype=torch.float32)
            rank[order] = torch.arange(n, dtype=torch.float32)  # rank as importance: the k lowest ranks = ordering[:k]
            counts = [0] * B
            for r in self.ratios:
                counts[int(r)] = int(0.1 * r * n)
            imp = rank.to(self.device).unsqueeze(0).expand(B, S).contiguous()
            return boundary.select_bits_per_window(imp, self.bits, D, counts=counts)
        return self._batched_boundary(batched_input_tokens, target, quant_layer, select)

    def batched_top_rho_quantization(self, batched_input_tokens, distribution, target, quant_layer=2, do_quant=True):
        """pythia_model.py:71-114 ("upto ratio"): batch row `ratio` keeps the fewest most-important tokens
        whose importance (`distribution`, the layer's column-mean scores) sums to >= 1 - 0.1 * ratio and
        has all the others quantised after layer `quant_layer`; returns one NLL per entry of self.ratios.
        The threshold search runs inside the selection kernel (no host sort, no sync)."""
        if not do_quant:
            return self._batched_boundary(batched_input_tokens, target, quant_layer, None)
        dist = torch.as_tensor(distribution, dtype=torch.float32).to(self.device).reshape(-1)

        def select(B, S, D):
            thresholds = [float(S + 1)] * B  # rows no ratio points at keep every token
            for r in self.ratios:
                thresholds[int(r)] = 1 - 0.1 * r
            return boundary.select_bits_per_window(dist.unsqueeze(0).expand(B, S).contiguous(), self.bits, D,
                                                   mass_thresholds=thresholds)
        return self._batched_boundary(batched_input_tokens, target, quant_layer, select)

    def calculate_batched_nll(self, hidden_states, target):
        """pythia_model.py:36-54: one NLL per entry of self.ratios, ratio used as batch index."""
        with torch.no_grad():
            logits = self.model.embed_out(self.final_layer_norm(hidden_states))
            target = target[:, 1:]
            logits = logits[:, :-1, :]
            return [torch.nn.functional.cross_entropy(logits[r, :, :].view(-1, logits.size(-1)), target.view(-1),
                                                      ignore_index=-100) for r in self.ratios]


class QwenPointFiveBModel(_LayerWiseModel):
    """Reference: Experiments/Qwen2-0.5B/qwen_layer_wise.py:5 (`QwenPointFiveBModel(device)`)."""
    family, hub_name, default_cfg = "qwen2", "Qwen/Qwen2-0.5B", compat.QWEN2_05B

    def __init__(self, device, **kw) -> None:
        self._from_pretrained(device, **kw)

    def _bind(self, model):
        self.qwen = model.base_model
        self.final_layer_norm = model.model.norm
        self.rotary_emb = self.qwen.rotary_emb

    def _embed(self, model, ids):
        return model.base_model.embed_tokens(ids)

    def _lm_head(self, model):
        return model.lm_head

    def _final_norm(self, model):
        return model.model.norm

    def _first_norm_params(self, layer):
        n = layer.input_layernorm
        return N.NORM_RMSNORM, n.weight, None, n.variance_epsilon, None, None

    def channel_wise(self, batched_input_tokens, target, layer_of_interest, method):
        """Channel-wise quantisation of the whole hidden state after `layer_of_interest`
        (qwen_layer_wise.py:106-167); method in {"channel_8","channel_4","channel_1_mean","channel_1_max"}."""
        with torch.no_grad():
            hidden = self.upstream(batched_input_tokens, layer_of_interest)
            hidden, _, _ = boundary.channel_fake_quant(hidden, method, out=hidden)
            return self.downstream(tuple(batched_input_tokens.shape), target, layer_of_interest, hidden=hidden)

    def _run_prenormed_layer(self, layer, resid, normed, normed2, pos):
        # Qwen2DecoderLayer.forward with input_layernorm already applied
        attn_out, _ = layer.self_attn(hidden_states=normed, attention_mask=None, position_embeddings=pos)
        hidden = resid + attn_out
        return hidden + layer.mlp(layer.post_attention_layernorm(hidden))
