"""Host-side logic of the DP eval driver: window schedule vs the reference loop, world-size-2
gloo run vs single process (CPU, stub wrapper), and a GPU end-to-end check on Pythia-70M."""
import math
import os
import pickle

import pytest
import torch
import torch.multiprocessing as mp

import splitq_oracle as O
from splitq import evalloop


@pytest.mark.parametrize("n,max_len,stride", [(1000, 256, 32), (100, 256, 32), (257, 256, 32), (4096, 512, 32),
                                              (33, 32, 32), (64, 32, 7), (1, 8, 4)])
def test_closed_form_schedule_equals_reference_loop(n, max_len, stride):
    assert evalloop.window_schedule(n, max_len, stride) == O.window_schedule(n, max_len, stride)


class StubWrapper:
    """Deterministic stand-in: NLL depends on the window's tokens, the split layer and the ratio."""
    device = down_device = torch.device("cpu")
    num_layers = 4

    def upstream(self, ids, loi):
        return ids.float().mean() + loi

    def boundary_send(self, hidden, imp, ratio):
        return hidden + ratio * imp.sum()

    def boundary_hop(self, payload):
        return payload

    def downstream(self, shape, target, loi, payload=None, hidden=None):
        x = payload if payload is not None else hidden
        return (x * 1e-3 + (target != -100).sum() * 1e-4).float()


def stub_importance(ids):
    S = ids.shape[1]
    return {"regular_importance": [torch.full((S,), 1.0 / S)] * 4, "last_row": [torch.linspace(0, 1, S)] * 4}


PARAMS = {"ratios": [0, 0.5, 1], "layers_of_interest": [1, 3], "methods": ["regular_importance", "last_row"], "stride": 32}


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    ids = torch.randint(0, 1000, (1, 1500), generator=torch.Generator().manual_seed(3))
    res = evalloop.evaluate(PARAMS, StubWrapper(), ids, stub_importance, 256, out_prefix=out, log=lambda *a: None)
    if rank == 0:
        with open(out + "res.pkl", "wb") as f:
            pickle.dump((res.total_nll, res.n_tokens, res.windows), f)
    torch.distributed.destroy_process_group()


def test_two_rank_gloo_equals_single_process(tmp_path):
    ids = torch.randint(0, 1000, (1, 1500), generator=torch.Generator().manual_seed(3))
    single = evalloop.evaluate(PARAMS, StubWrapper(), ids, stub_importance, 256, log=lambda *a: None)
    out = str(tmp_path) + "/"
    mp.spawn(_worker, args=(2, 29611, out), nprocs=2, join=True)
    total, n_tokens, windows = pickle.load(open(out + "res.pkl", "rb"))
    assert n_tokens == single.n_tokens == 1500 - len(evalloop.window_schedule(1500, 256, 32))
    assert windows == single.windows == len(evalloop.window_schedule(1500, 256, 32))
    a, b = torch.tensor(total), torch.tensor(single.total_nll)
    assert torch.allclose(a, b, rtol=1e-12, atol=0)
    # the reference's output files, written by rank 0
    assert pickle.load(open(out + "n_tokens.pkl", "rb")) == n_tokens
    ppl = pickle.load(open(out + "avg_ppl_results.pkl", "rb"))
    assert ppl[0][0][0] == pytest.approx(math.exp(total[0][0][0] / n_tokens))
    # ratio 0 is method independent, as in the reference's own logs
    assert total[0][0][0] == total[1][0][0]


@pytest.mark.gpu
def test_eval_driver_matches_naive_reference_loop_on_pythia():
    """evaluate() (upstream once per layer, fused importance) == the reference's loop structure
    (full activation_quantization per (method, layer, ratio), materialised importance)."""
    from splitq import compat, get_importance_order
    from splitq.main import run
    from splitq.models import Pythia70Model

    params = {"ratios": [0, 0.5], "layers_of_interest": [2, 5], "methods": ["regular_importance", "last_row", "aggregate_till"],
              "max_length": 256}
    n = 256 + 3 * 32
    res = run("pythia", dict(params, importance_impl="fused"), n, "cuda:0")
    res_m = run("pythia", dict(params, importance_impl="materialised"), n, "cuda:0")
    # naive loop
    w = Pythia70Model.from_config("cuda:0", params["ratios"], cfg=compat.PYTHIA_70M, seed=0)
    eager = compat.clone_with_attention(w.model, "eager").to("cuda:0")
    ids_all = torch.randint(0, compat.PYTHIA_70M["vocab_size"], (1, n), generator=torch.Generator().manual_seed(1234))
    total = [[[0.0] * 2 for _ in range(2)] for _ in range(3)]
    ntok = 0
    for begin, end, trg in O.window_schedule(n, 256, 32):
        ids = ids_all[:, begin:end].to("cuda:0")
        tgt, nlt = O.window_targets(ids, trg)
        with torch.no_grad():
            att = eager(input_ids=ids, output_attentions=True).attentions
        for m, method in enumerate(params["methods"]):
            imp = get_importance_order(method, att, 6)
            for l, loi in enumerate(params["layers_of_interest"]):
                for r, ratio in enumerate(params["ratios"]):
                    total[m][l][r] += w.activation_quantization(ids, tgt, imp[loi], ratio, loi).item() * nlt
        ntok += nlt
    assert res.n_tokens == res_m.n_tokens == ntok and res.windows == 4
    for got in (res, res_m):
        for m in range(3):
            for l in range(2):
                for r in range(2):
                    want = math.exp(total[m][l][r] / ntok)
                    assert abs(got.ppl[m][l][r] - want) / want <= 1e-3


def _pair_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    from splitq.transport import PairExchange, pair_partner
    ex = PairExchange()
    src = torch.full((2, 64), rank, dtype=torch.uint8)
    dst = torch.zeros_like(src)
    ex.send(src, dst)
    assert ex.partner == pair_partner(rank, world)
    assert (dst == ex.partner).all()
    torch.distributed.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_pair_exchange_over_gloo(tmp_path, world):
    """rank r swaps payloads with r^1; an unpaired last rank keeps its own (local copy)."""
    mp.spawn(_pair_worker, args=(world, 29620 + world, str(tmp_path)), nprocs=world, join=True)


@pytest.mark.gpu
def test_batching_methods_through_downstream_changes_nothing():
    from splitq import compat
    from splitq.models import QwenPointFiveBModel
    cfg = dict(compat.QWEN2_TINY, hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=256)
    w = QwenPointFiveBModel.from_config("cuda:0", cfg=cfg, seed=0)
    params = {"ratios": [0, 0.5, 1], "layers_of_interest": [0, 2], "methods": ["regular_importance", "last_row", "aggregate_till"],
              "stride": 32}
    ids = torch.randint(0, 512, (1, 200), generator=torch.Generator().manual_seed(5))
    fn = evalloop.make_importance_fn(w, params["methods"], "fused")
    a = evalloop.evaluate(params, w, ids, fn, 128, log=lambda *x: None)
    b = evalloop.evaluate(params, w, ids, fn, 128, log=lambda *x: None, batch_methods=True)
    assert a.n_tokens == b.n_tokens
    assert torch.allclose(torch.tensor(a.total_nll), torch.tensor(b.total_nll), rtol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("family", ["qwen", "pythia"])
def test_batching_all_method_ratio_variants_changes_nothing(family):
    """batch_combos: the payloads of different ratios (different widths) are dequantised one by one
    and their rows run through the downstream half as one batch; NLLs equal the one-by-one loop,
    including the split at the last layer (dequantise fused into the final norm)."""
    from splitq import compat
    from splitq.models import Pythia70Model, QwenPointFiveBModel
    ratios = [0, 0.25, 0.75, 1]
    if family == "qwen":
        cfg = dict(compat.QWEN2_TINY, hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=256)
        w = QwenPointFiveBModel.from_config("cuda:0", cfg=cfg, seed=0)
    else:
        cfg = dict(compat.PYTHIA_TINY, hidden_size=256, num_attention_heads=4, intermediate_size=256)
        w = Pythia70Model.from_config("cuda:0", ratios, cfg=cfg, seed=0)
    last = w.num_layers - 1
    params = {"ratios": ratios, "layers_of_interest": [0, last], "methods": ["regular_importance", "last_row"], "stride": 32}
    ids = torch.randint(0, 512, (1, 200), generator=torch.Generator().manual_seed(6))
    fn = evalloop.make_importance_fn(w, params["methods"], "fused")
    a = evalloop.evaluate(params, w, ids, fn, 128, log=lambda *x: None)
    b = evalloop.evaluate(params, w, ids, fn, 128, log=lambda *x: None, batch_combos=True)
    assert a.n_tokens == b.n_tokens
    assert torch.allclose(torch.tensor(a.total_nll), torch.tensor(b.total_nll), rtol=1e-5)
