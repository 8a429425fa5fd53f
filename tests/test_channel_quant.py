"""Channel-wise quantisers: oracle vs fixtures recorded from the reference's
QwenPointFiveBModel.channel_wise, and the CUDA kernels vs both."""
import json
import os

import numpy as np
import pytest
import torch

import splitq
import splitq_oracle as O
from splitq import compat
from splitq.models import QwenPointFiveBModel

METHODS = ["channel_8", "channel_4", "channel_1_mean", "channel_1_max"]


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "channel_quant.npz")), json.load(open(os.path.join(golden_dir, "channel_nll.json")))


def test_oracle_reproduces_reference_channel_loop(gold):
    g, _ = gold
    keys = sorted({k.rsplit("|", 1)[0] for k in g.files})
    assert len(keys) >= 6
    for key in keys:
        method = key.split("|")[1]
        post = O.channel_fake_quant_reference(torch.from_numpy(g[key + "|pre"]).unsqueeze(0), method)[0].numpy()
        assert np.array_equal(post, g[key + "|post"]), key


def check_mean_variant(x, got, u, codes):
    """channel_1_mean divides by a per-channel mean of zero-centred data: the scale is an
    ill-conditioned float sum, so ATen's fp32 summation order and our fp64-accumulated mean agree
    only to the rounding error of the sum.  Checked instead: (1) the scale is within that error of
    ATen's, (2) codes equal ATen's except right at a rounding boundary, (3) output == code * scale."""
    u_ref = x.mean(dim=0) + 1e-8
    tol = 4e-6 * x.abs().mean(dim=0) + 1e-12
    assert ((u - u_ref).abs() <= tol).all()
    ratio = x / u_ref
    want_codes = torch.clamp(torch.round(ratio), -1, 1)
    mismatch = codes.float() != want_codes
    near_boundary = (ratio.abs() - 0.5).abs() < 5e-2
    assert (~mismatch | near_boundary).all()
    assert mismatch.float().mean().item() < 1e-2
    assert np.array_equal(got, (codes.float() * u).numpy())


def test_unknown_channel_method():
    with pytest.raises(ValueError):
        splitq.channel_fake_quant(torch.zeros(1, 2, 4), "channel_3")


@pytest.mark.gpu
def test_kernel_matches_reference_fixtures(gold):
    g, _ = gold
    for key in sorted({k.rsplit("|", 1)[0] for k in g.files}):
        method = key.split("|")[1]
        pre = torch.from_numpy(g[key + "|pre"]).unsqueeze(0).cuda()
        out, stats, codes = splitq.channel_fake_quant(pre, method, want_codes=True)
        got, want = out[0].cpu().numpy(), g[key + "|post"]
        if method == "channel_1_mean":
            check_mean_variant(pre[0].cpu(), got, stats.cpu(), codes[0].cpu())
        else:
            assert np.array_equal(got, want), key       # value-exact (-0.0 == +0.0)
            nz = want != 0
            assert np.array_equal(got.view(np.uint32)[nz], want.view(np.uint32)[nz]), key
        L = {"channel_8": 127, "channel_4": 7}.get(method, 1)
        assert int(codes.abs().max()) <= L


@pytest.mark.gpu
@pytest.mark.parametrize("B,S,D", [(1, 300, 896), (2, 77, 130), (1, 5, 3), (1, 4096, 512)])
@pytest.mark.parametrize("method", METHODS)
def test_kernel_matches_oracle_on_random_shapes(B, S, D, method):
    gen = torch.Generator().manual_seed(S + D)
    h = torch.randn(B, S, D, generator=gen)
    h[:, :, D // 2] = 0.0  # an all-zero channel: reference divides 0/0 (absmax) or by ~1e-8 (mean)
    out, stats, _ = splitq.channel_fake_quant(h.cuda(), method)
    want = O.channel_fake_quant_reference(h, method)
    got = out.cpu()
    assert torch.equal(torch.isnan(got), torch.isnan(want))
    if method == "channel_1_mean":
        _, _, codes = splitq.channel_fake_quant(h.cuda(), method, want_codes=True)
        for b in range(B):
            pass
        check_mean_variant(h.reshape(-1, D), got.reshape(-1, D).numpy(), stats.cpu(), codes.reshape(-1, D).cpu())
    else:
        assert np.array_equal(torch.nan_to_num(got).numpy(), torch.nan_to_num(want).numpy())
    inplace = h.cuda()
    splitq.channel_fake_quant(inplace, method, out=inplace)
    assert torch.equal(torch.nan_to_num(inplace), torch.nan_to_num(out))


@pytest.mark.gpu
@pytest.mark.parametrize("name,cfg", [("qwen_tiny", compat.QWEN2_TINY), ("qwen2_05b", compat.QWEN2_05B)])
def test_wrapper_channel_wise_matches_reference_nll(gold, name, cfg):
    _, book = gold
    w = QwenPointFiveBModel.from_config("cuda:0", cfg=cfg, seed=0)
    ids = torch.tensor([book[name]["ids"]], device="cuda:0")
    for method, want in book[name]["nll"].items():
        got = w.channel_wise(ids, ids.clone(), book[name]["layer"], method).item()
        assert abs(got - want) / abs(want) <= 1e-3, (method, got, want)
        assert abs(np.expm1(got - want)) <= 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("method", ["channel_8", "channel_4", "channel_1_mean", "channel_1_max"])
def test_channel_payload_pack_and_dequant(method, golden_dir):
    """The channel quantisers as an SPQ1 payload (per-channel scales section): bytes equal the oracle's
    packing of the reference formulas, and dequantising the payload equals the fake-quantised floats of
    splitq_channel_fake_quant (hence the reference fixtures) bit for bit, window by window."""
    import numpy as np
    import splitq_oracle as O
    gen = torch.Generator().manual_seed(7)
    B, S, D = 3, 130, 96
    h = torch.randn(B, S, D, generator=gen) * 3
    h[:, :, 5] *= 40
    if method != "channel_1_mean":
        h[1, :, 9] = 0.0     # a dead channel in one window: statistic 0 -> NaN, as the reference's 0/0
    payload = splitq.channel_quant_pack(h.cuda(), method)
    assert payload.shape == (B, splitq.channel_payload_bytes(S, D, method))
    out = splitq.dequant(payload, S, D).cpu()
    for b in range(B):
        want_bytes = O.pack_payload_channel(h[b], method)
        got = payload[b].cpu().numpy()
        if method == "channel_1_mean":   # the statistic is a mean: GPU and CPU summation orders differ in the last bits
            so = 32 + O.pad16(S) + O.pad16(4 * S)
            stat_g = got[so:so + 4 * D].view(np.float32)
            stat_w = want_bytes[so:so + 4 * D].view(np.float32)
            assert np.allclose(stat_g, stat_w, rtol=1e-4, atol=1e-6)
        else:
            assert np.array_equal(got, want_bytes), (method, b)
            want = O.unpack_payload_channel(want_bytes)
            assert np.array_equal(out[b].numpy().view(np.uint32)[~np.isnan(want.numpy())],
                                  want.numpy().view(np.uint32)[~np.isnan(want.numpy())])
            assert np.array_equal(np.isnan(out[b].numpy()), np.isnan(want.numpy()))
        fq, _, _ = splitq.channel_fake_quant(h[b:b + 1].cuda(), method)
        a, c = out[b].numpy(), fq[0].cpu().numpy()
        finite = ~np.isnan(c)
        assert np.array_equal(np.isnan(a), np.isnan(c))
        nz = finite & (c != 0)   # an integer code has no -0
        assert np.array_equal(a.view(np.uint32)[nz], c.view(np.uint32)[nz]) and np.array_equal(a[finite], c[finite])
    # a reference fixture through the payload: qwen_tiny hidden states before / after the reference's own loop
    g = np.load(os.path.join(golden_dir, "channel_quant.npz"))
    key = f"qwen_tiny|{method}"
    pre, post = torch.from_numpy(g[key + "|pre"]), g[key + "|post"]
    got = splitq.dequant(splitq.channel_quant_pack(pre.cuda().unsqueeze(0), method), pre.shape[0], pre.shape[1])[0].cpu().numpy()
    if method == "channel_1_mean":  # the scale is a mean (summation order): values agree to its rounding, a few codes may flip
        assert np.mean(~np.isclose(got, post, rtol=1e-4, atol=1e-9)) < 0.02
    else:
        assert np.array_equal(got, post)
    # fused with the downstream norm, like the token payloads
    gamma = torch.rand(D).cuda()
    _, normed, _ = splitq.dequant_norm(payload, S, D, splitq.NORM_RMSNORM, gamma, None, 1e-6, want_resid=True)
    ok = ~torch.isnan(out).any(dim=2)
    want_n = O.rms_norm(out[ok], gamma.cpu(), 1e-6)
    assert torch.allclose(normed.cpu()[ok], want_n, rtol=1e-5, atol=1e-6)
