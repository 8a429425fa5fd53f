"""K2 through the HF attention hook vs the reference route (eager attentions + reductions)."""
import pytest
import torch

import splitq
from splitq import compat, hf_hook

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
QWEN_HD64 = dict(compat.QWEN2_TINY, hidden_size=256, num_attention_heads=4, num_key_value_heads=2,
                 intermediate_size=256, num_hidden_layers=3)


@pytest.mark.parametrize("family,cfg,S", [("gpt_neox", compat.PYTHIA_70M, 300), ("qwen2", QWEN_HD64, 200)])
def test_hook_importance_equals_eager_route(family, cfg, S):
    model = compat.build_causal_lm(family, cfg, seed=0).to(DEV)
    eager = compat.clone_with_attention(model, "eager").to(DEV)
    L = cfg["num_hidden_layers"]
    ids = torch.randint(0, cfg["vocab_size"], (1, S), generator=torch.Generator().manual_seed(1), device="cpu").to(DEV)
    methods = ["regular_importance", "last_row", "aggregate_till", "weighted_importance"]
    hw = (torch.randn(L, cfg["num_attention_heads"], generator=torch.Generator().manual_seed(2)) * 0.3 + 0.07).tolist()
    with torch.no_grad():
        ref_logits = model(input_ids=ids).logits
        att = eager(input_ids=ids, output_attentions=True).attentions
    got = hf_hook.fused_importance(model, ids, methods, L, hw)
    assert model.config._attn_implementation == "sdpa"  # restored
    with torch.no_grad():
        assert torch.equal(model(input_ids=ids).logits, ref_logits)  # the hook leaves the forward untouched
    for m in methods:
        want = splitq.get_importance_order(m, att, L, hw)
        for l in range(L):
            if m == "weighted_importance":  # signed head weights cancel: compare against the vector's scale
                err = ((got[m][l] - want[l]).abs().max() / want[l].abs().max()).item()
            else:
                err = ((got[m][l] - want[l]).abs() / want[l].abs().clamp_min(1e-30)).max().item()
            assert err <= 1e-5, (m, l, err)


def test_hook_layer_subset():
    model = compat.build_causal_lm("gpt_neox", compat.PYTHIA_70M, seed=0).to(DEV)
    ids = torch.randint(0, 1000, (1, 130), device=DEV)
    got = hf_hook.fused_importance(model, ids, ["regular_importance"], 6, layers=[2, 5])
    assert [g is not None for g in got["regular_importance"]] == [False, False, True, False, False, True]
