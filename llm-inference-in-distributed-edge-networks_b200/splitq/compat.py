"""Offline model construction and transformers-version isolation.

The reference wrappers load `EleutherAI/pythia-70m` / `Qwen/Qwen2-0.5B` with
`from_pretrained` (Experiments/Pythia-70M/pythia_model.py:24-34,
Experiments/Qwen2-0.5B/qwen_layer_wise.py:16-25).  There is no network here, so the
same architectures are built from their configs with seeded random weights.  All
transformers API drift (5.x decoder layers return a bare tensor, attention-interface
registration) is confined to this module.
"""
from __future__ import annotations

import torch

PYTHIA_70M = dict(
    vocab_size=50304, hidden_size=512, num_hidden_layers=6, num_attention_heads=8,
    intermediate_size=2048, max_position_embeddings=2048, rotary_pct=0.25,
    use_parallel_residual=True, layer_norm_eps=1e-5, hidden_act="gelu",
    tie_word_embeddings=False,
)
QWEN2_05B = dict(
    vocab_size=151936, hidden_size=896, num_hidden_layers=24, num_attention_heads=14,
    num_key_value_heads=2, intermediate_size=4864, max_position_embeddings=131072,
    rms_norm_eps=1e-6, rope_theta=1e6, tie_word_embeddings=True,
)
# small stand-ins with the same structure, for CPU tests
PYTHIA_TINY = dict(PYTHIA_70M, vocab_size=512, hidden_size=64, num_hidden_layers=3,
                   num_attention_heads=4, intermediate_size=128, max_position_embeddings=256)
QWEN2_TINY = dict(QWEN2_05B, vocab_size=512, hidden_size=64, num_hidden_layers=3,
                  num_attention_heads=4, num_key_value_heads=2, intermediate_size=128,
                  max_position_embeddings=256)


def build_causal_lm(family: str, cfg: dict, seed: int = 0, attn_implementation: str = "sdpa"):
    """Random-init HF model of `family` in {"gpt_neox","qwen2"}; deterministic for a seed."""
    from transformers import GPTNeoXConfig, GPTNeoXForCausalLM, Qwen2Config, Qwen2ForCausalLM

    if family == "gpt_neox":
        config = GPTNeoXConfig(**cfg)
        cls = GPTNeoXForCausalLM
    elif family == "qwen2":
        config = Qwen2Config(**cfg)
        cls = Qwen2ForCausalLM
    else:
        raise ValueError(f"unknown model family: {family}")
    config._attn_implementation = attn_implementation
    rng_state = torch.get_rng_state()
    torch.manual_seed(seed)
    try:
        model = cls(config)
    finally:
        torch.set_rng_state(rng_state)
    return model.eval()


def clone_with_attention(model, attn_implementation: str):
    """Second copy sharing the same weights but another attention implementation (the
    reference keeps an "eager" copy for output_attentions, last_row_exp.py:68-70)."""
    import copy

    config = copy.deepcopy(model.config)
    config._attn_implementation = attn_implementation
    twin = type(model)(config)
    twin.load_state_dict(model.state_dict())
    return twin.eval()


def layer_output(out):
    """Decoder layers return a tuple on transformers 4.x and a tensor on 5.x."""
    return out[0] if isinstance(out, (tuple, list)) else out


def decoder_layers(base_model):
    return base_model.layers
