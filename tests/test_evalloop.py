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


def free_port() -> int:
    import socket
    with socket.socket() as sock:
        sock.bind(("127.0.0.1", 0))
        return sock.getsockname()[1]


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

    def payload_like(self, rows, S, ratio):
        return torch.empty(())

    def hidden_like(self, rows, S):
        return torch.empty(())

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
    mp.spawn(_worker, args=(2, free_port(), out), nprocs=2, join=True)
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


def _pairs_worker(rank, world, port, out, n_tokens):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    ids = torch.randint(0, 1000, (1, n_tokens), generator=torch.Generator().manual_seed(3))
    res = evalloop.evaluate(PARAMS, StubWrapper(), ids, stub_importance, 256, out_prefix=out, log=lambda *a: None,
                            pairs=True, checkpoint_every=2, file_names=evalloop.FILE_NAMES["pythia"])
    with open(out + f"res{rank}.pkl", "wb") as f:
        pickle.dump((res.total_nll, res.n_tokens, res.windows, res.hop_bytes), f)
    torch.distributed.destroy_process_group()


@pytest.mark.parametrize("world,n_tokens", [(2, 1500), (4, 1500), (3, 700), (4, 256 + 5 * 32)])
def test_gpu_pair_layout_over_gloo_equals_single_process(tmp_path, world, n_tokens):
    """pairs=True: rank 2p = upstream half, rank 2p+1 = downstream half, payloads and ratio-0 hidden
    states cross by send/recv, pair p of P owns windows k = p (mod P); an odd last rank idles but
    still joins the collectives (checkpoint_every=2 makes the periodic all-reduce fire while the
    shards hold different window counts -- the gate depends on the global index only).  Totals
    equal the single-process run bit for bit on every rank."""
    ids = torch.randint(0, 1000, (1, n_tokens), generator=torch.Generator().manual_seed(3))
    single = evalloop.evaluate(PARAMS, StubWrapper(), ids, stub_importance, 256, log=lambda *a: None)
    out = str(tmp_path) + "/"
    mp.spawn(_pairs_worker, args=(world, free_port(), out, n_tokens), nprocs=world, join=True)
    for rank in range(world):
        total, n_tok, windows, hop = pickle.load(open(out + f"res{rank}.pkl", "rb"))
        assert n_tok == single.n_tokens and windows == single.windows
        assert torch.allclose(torch.tensor(total), torch.tensor(single.total_nll), rtol=1e-12, atol=0)
        is_up = rank % 2 == 0 and rank < 2 * (world // 2)
        assert (hop > 0) == is_up
    # the Pythia script's own file names (last_row_exp.py:122-124,142)
    assert pickle.load(open(out + "n_tokens_pythia_70m.pkl", "rb")) == single.n_tokens
    assert os.path.exists(out + "total_nll_pythia_70m.pkl") and os.path.exists(out + "avg_ppl_results_pythia_70m.pkl")


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
    mp.spawn(_pair_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)


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


@pytest.mark.gpu
def test_channel_wise_driver_matches_wrapper_calls(tmp_path):
    """evaluate_channel == the reference's channel_wise.py loop structure (one wrapper.channel_wise call
    per (window, method, layer)), and writes its three pickle files."""
    from splitq import compat
    from splitq.models import QwenPointFiveBModel
    cfg = dict(compat.QWEN2_TINY, hidden_size=256, num_attention_heads=4, num_key_value_heads=2, intermediate_size=256,
               max_position_embeddings=1024)
    w = QwenPointFiveBModel.from_config("cuda:0", cfg=cfg, seed=0)
    methods, layers = ["channel_8", "channel_4", "channel_1_mean", "channel_1_max"], [0, 1]
    ids_all = torch.randint(0, 512, (1, 512 + 2 * 32), generator=torch.Generator().manual_seed(9))
    out = str(tmp_path) + "/"
    res = evalloop.evaluate_channel(layers, methods, w, ids_all, out_prefix=out, log=lambda *a: None)
    total = [[0.0] * len(methods) for _ in layers]
    ntok = 0
    for begin, end, trg in O.window_schedule(ids_all.shape[1], 512, 32):
        ids = ids_all[:, begin:end].to("cuda:0")
        tgt, nlt = O.window_targets(ids, trg)
        for m, method in enumerate(methods):
            for l, loi in enumerate(layers):
                total[l][m] += w.channel_wise(ids, tgt, loi, method).item() * nlt
        ntok += nlt
    assert res.n_tokens == ntok and res.windows == 3
    assert torch.allclose(torch.tensor(res.total_nll), torch.tensor(total), rtol=1e-5)
    assert pickle.load(open(out + "n_tokens_channel.pkl", "rb")) == ntok
    ppl = pickle.load(open(out + "ppl_channel.pkl", "rb"))
    assert ppl[1][2] == pytest.approx(math.exp(total[1][2] / ntok), rel=1e-5)
    assert os.path.exists(out + "total_nll_channel.pkl")


PAIR_WORKER = r'''
import json, os, sys
import torch
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))
from splitq.main import run
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", rank))
params = json.loads(os.environ["SPLITQ_PARAMS"])
res = run("pythia", params, int(os.environ["SPLITQ_TOKENS"]), f"cuda:{rank}", pairs=True, out_prefix=os.environ["SPLITQ_OUT"])
if rank == 0:
    json.dump({"total": res.total_nll, "n": res.n_tokens, "w": res.windows}, open(os.environ["SPLITQ_OUT"] + "pair.json", "w"))
torch.distributed.destroy_process_group()
'''


@pytest.mark.gpu
@pytest.mark.parametrize("batch_combos", [False, True])
def test_gpu_pair_eval_equals_single_gpu_eval(tmp_path, batch_combos):
    """SURVEY 8e on real devices: upstream half on GPU 0 (rank 0), downstream half on GPU 1 (rank 1),
    payloads over NCCL send/recv -- the NLL sums equal the one-GPU run bit for bit (same kernels, same
    inputs; only where the two halves execute differs) and the Pythia file names are written."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import json
    import subprocess
    import sys
    from splitq.main import run
    params = {"ratios": [0, 0.5, 1], "layers_of_interest": [2, 5], "methods": ["regular_importance", "last_row"],
              "max_length": 256, "batch_combos": batch_combos}
    n = 256 + 4 * 32
    single = run("pythia", params, n, "cuda:0")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "pair_worker.py"
    script.write_text(f"ROOT = {root!r}\n" + PAIR_WORKER)
    out = str(tmp_path) + "/"
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), WORLD_SIZE="2",
               SPLITQ_PARAMS=json.dumps(params), SPLITQ_TOKENS=str(n), SPLITQ_OUT=out)
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]
    got = json.load(open(out + "pair.json"))
    assert got["n"] == single.n_tokens and got["w"] == single.windows
    assert torch.equal(torch.tensor(got["total"]), torch.tensor(single.total_nll))
    assert os.path.exists(out + "avg_ppl_results_pythia_70m.pkl")


@pytest.mark.gpu
def test_main_run_reads_optional_params_keys(tmp_path):
    """splitq.main.run with the optional params.json keys of SURVEY section 5: `bits` / `ladder` /
    `scale_mode` reach the wrapper, `head_weights_file` feeds weighted_importance like the reference's
    ../../attention_head_weights.pkl, channel methods route to the channel-wise driver, and the output
    files carry the family's reference names."""
    from splitq import compat
    from splitq.main import run
    L, H = compat.QWEN2_TINY["num_hidden_layers"], compat.QWEN2_TINY["num_attention_heads"]
    base = {"ratios": [0, 0.5], "layers_of_interest": [1], "methods": ["regular_importance"], "max_length": 128, "stride": 32}
    n = 128 + 2 * 32
    # head_dim of the tiny configs is 16: the fused kernel needs 64, so score from materialised attention
    mat = dict(base, importance_impl="materialised")
    a = run("qwen", mat, n, "cuda:0", tiny=True)
    b = run("qwen", dict(mat, bits=8), n, "cuda:0", tiny=True)
    c = run("qwen", dict(mat, scale_mode="token", ladder=[[0.5, 2], [0.5, 8]]), n, "cuda:0", tiny=True)
    assert a.total_nll[0][0][0] == b.total_nll[0][0][0] == c.total_nll[0][0][0]          # ratio 0: no quantiser
    assert len({a.total_nll[0][0][1], b.total_nll[0][0][1], c.total_nll[0][0][1]}) == 3   # ratio 0.5: three quantisers
    hw = (torch.randn(L, H, generator=torch.Generator().manual_seed(1)) * 0.3 + 0.1).tolist()
    path = tmp_path / "attention_head_weights.pkl"
    pickle.dump(hw, open(path, "wb"))
    w1 = run("qwen", dict(mat, methods=["weighted_importance"], head_weights_file=str(path)), n, "cuda:0", tiny=True)
    w2 = run("qwen", dict(mat, methods=["weighted_importance"], head_weights=hw), n, "cuda:0", tiny=True)
    assert w1.total_nll == w2.total_nll
    with pytest.raises(FileNotFoundError):
        run("qwen", dict(mat, methods=["weighted_importance"]), n, "cuda:0", tiny=True)
    out = str(tmp_path) + "/"
    ch = run("qwen", {"layers_of_interest": [1], "methods": ["channel_8", "channel_1_max"], "max_length": 512,
                      "stride": 32, "ratios": [0]}, 512 + 32, "cuda:0", tiny=True, out_prefix=out)
    assert len(ch.total_nll) == 1 and len(ch.total_nll[0]) == 2 and os.path.exists(out + "ppl_channel.pkl")
    run("pythia", dict(mat, experiment="last_row"), n, "cuda:0", tiny=True, out_prefix=out)
    assert os.path.exists(out + "avg_ppl_results_pythia_70m.pkl") and os.path.exists(out + "total_nll_pythia_70m.pkl")
    with pytest.raises(ValueError):
        run("pythia", dict(mat, experiment="initial"), n, "cuda:0", tiny=True)
    with pytest.raises(ValueError):
        run("qwen", dict(mat, pairs=True, split_devices=["cuda:0", "cuda:0"]), n, "cuda:0", tiny=True)
