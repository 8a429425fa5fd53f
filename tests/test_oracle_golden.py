"""The oracle restatement replayed against fixtures produced by the reference's own code
(oracle/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

import splitq_oracle as O


@pytest.fixture(scope="module")
def imp_gold(golden_dir):
    return np.load(os.path.join(golden_dir, "importance.npz"))


@pytest.mark.parametrize("tag", ["a", "b", "c"])
@pytest.mark.parametrize("method", ["regular_importance", "last_row", "aggregate_till", "weighted_importance"])
def test_importance_matches_reference(imp_gold, tag, method):
    A = torch.from_numpy(imp_gold[f"{tag}_A"])
    maps = [A[l] for l in range(A.shape[0])]
    hw = imp_gold[f"{tag}_head_w"].tolist()
    got = torch.stack(O.get_importance_order(method, maps, len(maps), hw)).numpy()
    assert np.array_equal(got, imp_gold[f"{tag}_{method}"])  # same ATen ops -> bit equal


def test_unknown_method_raises_like_reference(imp_gold):
    with pytest.raises(ValueError) as e:
        O.get_importance_order("nope", [torch.zeros(1, 1, 2, 2)], 1)
    assert str(e.value) == str(imp_gold["unknown_method_message"])


def test_fake_quant_block_bit_exact(golden_dir):
    g = np.load(os.path.join(golden_dir, "fake_quant.npz"))
    keys = sorted({k.rsplit("|", 1)[0] for k in g.files})
    assert len(keys) >= 30
    for key in keys:
        ratio = float(key.split("|")[3])
        pre = torch.from_numpy(g[key + "|pre"]).unsqueeze(0)
        imp = torch.from_numpy(g[key + "|imp"])
        post = O.fake_quant_reference(pre, imp, ratio)[0].numpy()
        assert np.array_equal(post, g[key + "|post"]), key


def fullsize_cases(golden_dir):
    """(name, pre (S,D), imp (S,), ratio, rows the reference rewrote, their new values)"""
    g = np.load(os.path.join(golden_dir, "fake_quant_fullsize.npz"))
    for name in sorted({k.split("|")[0] for k in g.files}):
        yield (name, torch.from_numpy(g[name + "|pre"]), torch.from_numpy(g[name + "|imp"]), float(g[name + "|ratio"]),
               g[name + "|rows"], g[name + "|post_rows"])


def test_fullsize_fake_quant_block_bit_exact(golden_dir):
    """The reference wrappers at BASELINE.json's window lengths (Pythia-70M S=512 / S=2048,
    Qwen2-0.5B S=512): the restatement and the payload round trip reproduce the rows the
    reference's in-place int4 block rewrote, and leave every other row untouched."""
    n = 0
    for name, pre, imp, ratio, rows, post_rows in fullsize_cases(golden_dir):
        S, D = pre.shape
        assert (S, D) in ((512, 512), (2048, 512), (512, 896)), name
        post = O.fake_quant_reference(pre.unsqueeze(0), imp, ratio)[0].numpy()
        assert np.array_equal(post[rows], post_rows), name
        keep = np.ones(S, dtype=bool)
        keep[rows] = False
        assert np.array_equal(post[keep], pre.numpy()[keep]), name
        k = O.num_selected(ratio, S)
        _, _, deq = O.boundary_roundtrip(pre, imp, [(k, 4)], O.SCALE_GROUP)
        assert np.array_equal(deq.numpy(), post), name
        n += 1
    assert n == 3


def test_payload_roundtrip_equals_reference_block(golden_dir):
    """pack -> unpack in group/int4 mode reproduces the reference's in-place fake quant bit for bit."""
    g = np.load(os.path.join(golden_dir, "fake_quant.npz"))
    for key in sorted({k.rsplit("|", 1)[0] for k in g.files}):
        ratio = float(key.split("|")[3])
        if ratio == 0:
            continue
        pre = torch.from_numpy(g[key + "|pre"])
        imp = torch.from_numpy(g[key + "|imp"])
        k = O.num_selected(ratio, pre.shape[0])
        bits, payload, deq = O.boundary_roundtrip(pre, imp, [(k, 4)], O.SCALE_GROUP)
        assert payload.size == O.payload_bytes(pre.shape[0], pre.shape[1], [(k, 4)], O.SCALE_GROUP)
        assert np.array_equal(deq.numpy(), g[key + "|post"]), key


def test_aten_facts(golden_dir):
    facts = json.load(open(os.path.join(golden_dir, "aten_facts.json")))
    assert facts["zero_block_dequant_is_nan"]
    z = torch.zeros(4, 8)
    _, _, deq = O.boundary_roundtrip(z, torch.arange(4.0), [(2, 4)], O.SCALE_GROUP)
    assert torch.isnan(deq[:2]).all() and (deq[2:] == 0).all()
    _, _, deq = O.boundary_roundtrip(z, torch.arange(4.0), [(2, 4)], O.SCALE_TOKEN)
    assert (deq == 0).all()
    assert [O.num_selected(r, s) for r, s in [(0.25, 511), (0.75, 2047), (0.1, 30), (1, 77)]] == facts["int_ratio"]
    assert O.least_important(torch.tensor([1.0, 0.0, 1.0, 0.0, 0.0, 2.0]), 6).tolist() == facts["argsort_ties_stable"]


@pytest.mark.parametrize("bits", [2, 4, 8])
def test_pack_unpack_fields(bits):
    rng = np.random.default_rng(bits)
    L = O.levels(bits)
    for D in (1, 5, 16, 33):
        codes = rng.integers(-L, L + 1, size=D).astype(np.int8)
        raw = O._pack_row(codes, bits)
        assert raw.size == (D * bits + 7) // 8
        assert np.array_equal(O._unpack_row(raw, bits, D), codes)


def test_window_schedule_matches_loop():
    # hand-rolled copy of the accounting identities of last_row_exp.py:85-130
    sched = O.window_schedule(1000, 256, 32)
    assert sched[0] == (0, 256, 256)
    assert all(t == 32 for _, _, t in sched[1:-1])
    assert sched[-1][1] == 1000
    assert sum(t for _, _, t in sched) == 1000
    assert O.window_schedule(100, 256, 32) == [(0, 100, 100)]


@pytest.mark.parametrize("method", ["channel_8", "channel_4", "channel_1_mean", "channel_1_max"])
def test_channel_payload_roundtrip_equals_reference_channel_wise(method, golden_dir):
    """Oracle only: packing the channel codes into the payload format and dequantising them reproduces
    what the reference's channel_wise loop left in the hidden states (captured from its own run)."""
    g = np.load(os.path.join(golden_dir, "channel_quant.npz"))
    pre, post = torch.from_numpy(g[f"qwen_tiny|{method}|pre"]), g[f"qwen_tiny|{method}|post"]
    buf = O.pack_payload_channel(pre, method)
    assert buf.size == 32 + O.pad16(pre.shape[0]) + O.pad16(4 * pre.shape[0]) + O.pad16(4 * pre.shape[1]) + \
        pre.shape[0] * O.row_bytes(pre.shape[1], O.CHANNEL_BITS[method])
    got = O.unpack_payload_channel(buf).numpy()
    assert np.array_equal(got, post)
