"""GPU parity tests: every CUDA kernel, called through the C ABI, against the CPU oracle and the
golden fixtures produced by the reference's own code.

Bars (BASELINE.json north_star): width maps, codes and packed payloads bit-exact; importance
<= 1e-5 relative; dequantised values bit-exact; fused norm <= 1e-5 (floating point reassociation).
"""
import os

import numpy as np
import pytest
import torch

import splitq
import splitq_oracle as O
from splitq import _native as N

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
IMP_RTOL = 1e-5


def synth_attention(gen, B, H, S, tau=1.0):
    mask = torch.full((S, S), float("-inf")).triu(1)
    return torch.softmax(torch.randn(B, H, S, S, generator=gen) * tau + mask, dim=-1, dtype=torch.float32)


def synth_hidden(gen, B, S, D, outliers=True):
    h = torch.randn(B, S, D, generator=gen)
    if outliers and D >= 8:
        ch = torch.randperm(D, generator=gen)[: max(1, D // 100)]
        h[:, :, ch] *= 20.0
    return h


def rel_err(got, want):
    return ((got - want).abs() / want.abs().clamp_min(1e-30)).max().item()


# ------------------------------------------------------------------------------------------
# K1 importance
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag", ["a", "b", "c"])
@pytest.mark.parametrize("method", ["regular_importance", "last_row", "aggregate_till", "weighted_importance"])
def test_importance_against_reference_fixtures(golden_dir, tag, method):
    g = np.load(os.path.join(golden_dir, "importance.npz"))
    A = torch.from_numpy(g[f"{tag}_A"]).to(DEV)
    maps = [A[l] for l in range(A.shape[0])]
    got = splitq.get_importance_order(method, maps, len(maps), g[f"{tag}_head_w"].tolist())
    want = torch.from_numpy(g[f"{tag}_{method}"])
    for l in range(len(maps)):
        assert got[l].shape == want[l].shape
        np.testing.assert_allclose(got[l].cpu().numpy(), want[l].numpy(), rtol=IMP_RTOL, atol=1e-9)


@pytest.mark.parametrize("B,H,S", [(1, 8, 512), (2, 14, 300), (1, 3, 1), (3, 5, 130), (1, 8, 2048)])
def test_importance_colmean_and_last_row_vs_oracle(B, H, S):
    gen = torch.Generator().manual_seed(B * 1000 + S)
    A = synth_attention(gen, B, H, S, tau=2.0)
    col, last = splitq.importance_from_attention(A.to(DEV), want_col=True, want_last=True)
    for b in range(B):
        want_c = O.importance_regular(A[b:b + 1])
        want_l = O.importance_last_row(A[b:b + 1])
        assert rel_err(col[b].cpu(), want_c) <= IMP_RTOL
        assert rel_err(last[b].cpu(), want_l) <= IMP_RTOL
        # ranking parity: the k least important tokens agree wherever the oracle has a clear gap
        assert abs(col[b].sum().item() - 1.0) < 1e-4


def test_importance_causal_hint_and_rectangular():
    gen = torch.Generator().manual_seed(7)
    A = synth_attention(gen, 2, 4, 384)
    full, _ = splitq.importance_from_attention(A.to(DEV))
    hint, _ = splitq.importance_from_attention(A.to(DEV), causal_hint=True)
    assert torch.equal(full, hint)  # skipped tiles hold exact zeros
    R = torch.rand(1, 3, 50, 70, generator=gen)
    col, last = splitq.importance_from_attention(R.to(DEV), want_last=True)
    assert rel_err(col[0].cpu(), R.mean(dim=1).mean(dim=1)[0]) <= IMP_RTOL
    assert rel_err(last[0].cpu(), R[:, :, -1, :].mean(dim=1)[0]) <= IMP_RTOL


def test_importance_unaligned_view_uses_scalar_path():
    gen = torch.Generator().manual_seed(8)
    base = torch.rand(1 * 2 * 64 * 64 + 1, generator=gen).to(DEV)
    A = base[1:].view(1, 2, 64, 64)  # 4-byte aligned only
    col, _ = splitq.importance_from_attention(A)
    assert rel_err(col[0].cpu(), O.importance_regular(A.cpu())) <= IMP_RTOL


# ------------------------------------------------------------------------------------------
# K3a selection
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("S", [1, 7, 1000, 1024, 1025, 4096, 5000, 16384, 32768])
def test_select_bits_matches_stable_argsort(S):
    gen = torch.Generator().manual_seed(S)
    D = 64
    imp = torch.rand(3, S, generator=gen)
    imp[1] = torch.randint(0, 5, (S,), generator=gen).float() / 4.0  # heavy ties
    imp[2, ::3] = 0.0
    imp[2, 1::7] = -0.0  # +0 and -0 are equal; ties by index
    for ladder in ([], [(S // 4, 4)], [(S // 4, 2), (S // 4, 4), (S // 4, 8)], [(0, 4), (S, 8)], [(S, 2)],
                   [(1, 8), (0, 2), (S // 2, 4)]):
        bits, row_off = splitq.select_bits(imp.to(DEV), ladder, D)
        for b in range(3):
            want = O.ladder_bits(imp[b], ladder)
            assert np.array_equal(bits[b].cpu().numpy(), want), (S, ladder, b)
            rb = np.array([O.row_bytes(D, int(x)) for x in want], dtype=np.int64)
            want_off = (np.cumsum(rb) - rb) // 16
            assert np.array_equal(row_off[b].cpu().numpy().astype(np.int64), want_off)


def test_select_rejects_bad_arguments():
    imp = torch.rand(1, 8, device=DEV)
    with pytest.raises(N.SplitqError):
        splitq.select_bits(imp, [(9, 4)], 16)
    with pytest.raises(N.SplitqError):
        splitq.select_bits(imp, [(1, 3)], 16)
    with pytest.raises(N.SplitqError):
        splitq.select_bits(torch.rand(1, N.MAX_SELECT_S + 1, device=DEV), [], 16)


# ------------------------------------------------------------------------------------------
# K3 quantise + pack, K4 dequantise
# ------------------------------------------------------------------------------------------
CASES = [
    (1, 48, 64, [(24, 4)]),
    (2, 100, 512, [(25, 2), (25, 4), (25, 8)]),
    (1, 64, 896, [(16, 2), (16, 4), (16, 8)]),
    (1, 33, 20, [(10, 4), (5, 8)]),
    (2, 19, 5, [(4, 2), (4, 4), (4, 8)]),       # odd D: generic path, sub-byte tails
    (1, 16, 2052, [(8, 4)]),                    # register tile of 32 chunks... (D/4 = 513 -> MAXCH 32)
    (1, 9, 4100, [(3, 2), (3, 8)]),             # D > 4096 -> generic path
    (1, 40, 512, [(40, 4)]),                    # ratio 1: everything quantised
    (1, 40, 512, []),                           # ratio 0 equivalent: all raw
]


@pytest.mark.parametrize("mode", [O.SCALE_GROUP, O.SCALE_TOKEN])
@pytest.mark.parametrize("B,S,D,ladder", CASES)
def test_payload_bit_exact_and_dequant_bit_exact(B, S, D, ladder, mode):
    gen = torch.Generator().manual_seed(S * 7 + D)
    h = synth_hidden(gen, B, S, D)
    imp = torch.rand(B, S, generator=gen)
    nbytes = splitq.payload_bytes(S, D, ladder, mode)
    bits, row_off = splitq.select_bits(imp.to(DEV), ladder, D)
    payload = splitq.quant_pack(h.to(DEV), bits, row_off, mode, nbytes)
    deq = splitq.dequant(payload, S, D)
    for b in range(B):
        want_bits, want_payload, want_deq = O.boundary_roundtrip(h[b], imp[b], ladder, mode)
        assert want_payload.size == nbytes
        got = payload[b].cpu().numpy()
        assert np.array_equal(got, want_payload), f"payload differs at byte {np.nonzero(got != want_payload)[0][:5]}"
        assert np.array_equal(deq[b].cpu().numpy().view(np.uint32), want_deq.numpy().view(np.uint32))


def test_reference_fake_quant_fixtures_bit_exact(golden_dir):
    """select -> pack -> dequant on the GPU == the reference's in-place block (captured from its
    own wrapper run), bit for bit, for every recorded (model, method, layer, ratio)."""
    g = np.load(os.path.join(golden_dir, "fake_quant.npz"))
    keys = sorted({k.rsplit("|", 1)[0] for k in g.files})
    checked = 0
    for key in keys:
        ratio = float(key.split("|")[3])
        pre = torch.from_numpy(g[key + "|pre"])
        imp = torch.from_numpy(g[key + "|imp"])
        S, D = pre.shape
        ladder = splitq.ladder_for_ratio(ratio, S)
        nbytes = splitq.payload_bytes(S, D, ladder, O.SCALE_GROUP)
        bits, row_off = splitq.select_bits(imp.to(DEV), ladder, D)
        payload = splitq.quant_pack(pre.to(DEV), bits, row_off, O.SCALE_GROUP, nbytes)
        post = splitq.dequant(payload, S, D)[0].cpu().numpy()
        # value equality: an integer code has no negative zero, the reference's fp32 fake-quant does
        # (round(-0.3) = -0.0); -0.0 == +0.0 numerically, every other value must match exactly
        want = g[key + "|post"]
        assert np.array_equal(post, want), key
        nz = want != 0
        assert np.array_equal(post.view(np.uint32)[nz], want.view(np.uint32)[nz]), key
        checked += 1
    assert checked >= 30


def test_reference_fullsize_fixtures_bit_exact(golden_dir):
    """Same check at the window lengths BASELINE.json's configs name (Pythia-70M S=512 column-mean,
    S=2048 last-row, Qwen2-0.5B S=512): the GPU select -> pack -> dequantise reproduces every row the
    reference's own wrapper rewrote and leaves all other rows bit-identical."""
    from test_oracle_golden import fullsize_cases
    n = 0
    for name, pre, imp, ratio, rows, post_rows in fullsize_cases(golden_dir):
        S, D = pre.shape
        ladder = splitq.ladder_for_ratio(ratio, S)
        nbytes = splitq.payload_bytes(S, D, ladder, O.SCALE_GROUP)
        bits, row_off = splitq.select_bits(imp.to(DEV), ladder, D)
        payload = splitq.quant_pack(pre.to(DEV), bits, row_off, O.SCALE_GROUP, nbytes)
        post = splitq.dequant(payload, S, D)[0].cpu().numpy()
        assert np.array_equal(post[rows], post_rows), name            # -0.0 == +0.0 (see above)
        nz = post_rows != 0
        assert np.array_equal(post[rows].view(np.uint32)[nz], post_rows.view(np.uint32)[nz]), name
        keep = np.ones(S, dtype=bool)
        keep[rows] = False
        assert np.array_equal(post[keep].view(np.uint32), pre.numpy()[keep].view(np.uint32)), name
        want_payload = O.pack_payload(pre, O.ladder_bits(imp, ladder), O.SCALE_GROUP)
        assert np.array_equal(payload[0].cpu().numpy(), want_payload), name
        n += 1
    assert n == 3


def test_zero_block_reproduces_reference_nan_in_group_mode_only():
    S, D = 8, 16
    h = torch.zeros(1, S, D)
    h[0, 4:] = torch.arange(1, 4 * D + 1).float().view(4, D)
    imp = torch.arange(S).float().unsqueeze(0)
    for mode in (O.SCALE_GROUP, O.SCALE_TOKEN):
        nbytes = splitq.payload_bytes(S, D, [(4, 4)], mode)
        bits, off = splitq.select_bits(imp.to(DEV), [(4, 4)], D)
        deq = splitq.dequant(splitq.quant_pack(h.to(DEV), bits, off, mode, nbytes), S, D)[0].cpu()
        want = O.boundary_roundtrip(h[0], imp[0], [(4, 4)], mode)[2]
        assert torch.equal(torch.isnan(deq), torch.isnan(want))
        assert torch.equal(torch.nan_to_num(deq), torch.nan_to_num(want))
        assert torch.isnan(deq[:4]).all() == (mode == O.SCALE_GROUP)


@pytest.mark.parametrize("kind", ["layernorm", "rmsnorm"])
@pytest.mark.parametrize("S,D", [(50, 512), (33, 896), (7, 20), (5, 6)])
def test_dequant_fused_norm(kind, S, D):
    gen = torch.Generator().manual_seed(D)
    B = 2
    h = synth_hidden(gen, B, S, D)
    imp = torch.rand(B, S, generator=gen)
    ladder = [(S // 4, 2), (S // 4, 4), (S // 4, 8)]
    mode = O.SCALE_TOKEN
    nbytes = splitq.payload_bytes(S, D, ladder, mode)
    bits, off = splitq.select_bits(imp.to(DEV), ladder, D)
    payload = splitq.quant_pack(h.to(DEV), bits, off, mode, nbytes)
    g1, b1 = torch.randn(D, generator=gen), torch.randn(D, generator=gen)
    g2, b2 = torch.randn(D, generator=gen), torch.randn(D, generator=gen)
    if kind == "layernorm":
        resid, n1, n2 = splitq.dequant_norm(payload, S, D, N.NORM_LAYERNORM, g1.to(DEV), b1.to(DEV), 1e-5,
                                            g2.to(DEV), b2.to(DEV))
    else:
        resid, n1, n2 = splitq.dequant_norm(payload, S, D, N.NORM_RMSNORM, g1.to(DEV), None, 1e-6)
    plain = splitq.dequant(payload, S, D)
    assert torch.equal(resid, plain)
    for b in range(B):
        x = O.unpack_payload(payload[b].cpu().numpy())
        if kind == "layernorm":
            np.testing.assert_allclose(n1[b].cpu().numpy(), O.layer_norm(x, g1, b1, 1e-5).numpy(), rtol=1e-5, atol=2e-5)
            np.testing.assert_allclose(n2[b].cpu().numpy(), O.layer_norm(x, g2, b2, 1e-5).numpy(), rtol=1e-5, atol=2e-5)
        else:
            assert n2 is None
            np.testing.assert_allclose(n1[b].cpu().numpy(), O.rms_norm(x, g1, 1e-6).numpy(), rtol=1e-5, atol=2e-5)


def test_corrupt_payload_fails_loudly():
    S, D = 16, 32
    payload = torch.zeros(1, splitq.payload_bytes(S, D, [], 0), dtype=torch.uint8, device=DEV)
    out = splitq.dequant(payload, S, D)
    assert torch.isnan(out).all()


def test_transfer_same_device_copy():
    src = torch.randint(0, 255, (3, 4096), dtype=torch.uint8, device=DEV)
    dst = torch.zeros_like(src)
    splitq.transfer(dst, src)
    torch.cuda.synchronize()
    assert torch.equal(src, dst)


# ------------------------------------------------------------------------------------------
# full size (BASELINE.json config 3 shapes): size-independent properties, then the payload bytes
# ------------------------------------------------------------------------------------------
def test_full_size_roundtrip_properties():
    B, H, S, D = 2, 14, 4096, 896
    gen = torch.Generator(device=DEV).manual_seed(3)
    h = torch.randn(B, S, D, device=DEV, generator=gen)
    logits = torch.randn(B, H, S, S, device=DEV, generator=gen)
    A = torch.softmax(logits + torch.full((S, S), float("-inf"), device=DEV).triu(1), dim=-1)
    del logits
    ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
    plan = splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=O.SCALE_TOKEN)
    res = plan.run(A, h)
    torch.cuda.synchronize()
    ref_imp = A.mean(dim=1).mean(dim=1)
    assert rel_err(res.importance, ref_imp) <= IMP_RTOL
    bits = res.bits
    assert [(bits == w).sum(dim=1).tolist() for w in (2, 4, 8, 32)] == [[1024] * B] * 4
    # the selected sets are exactly the lowest-ranked tokens of the kernel's own scores
    order = torch.argsort(res.importance, dim=1, stable=True)
    for b in range(B):
        assert (bits[b][order[b, :1024]] == 2).all() and (bits[b][order[b, 3072:]] == 32).all()
    raw = bits == 32
    assert torch.equal(res.hidden[raw], h[raw])  # raw rows cross untouched
    m = h.abs().amax(dim=2, keepdim=True)
    for w in (2, 4, 8):
        L = O.levels(w)
        sel = bits == w
        err = (res.hidden - h).abs()[sel]
        mm = m.expand_as(h)[sel]
        assert (err <= mm / L * 0.5 + mm * 1e-6).all()  # half a step + fp32 rounding of q/L*m
    # idempotence: re-quantising the dequantised tensor with the same map reproduces the payload codes
    p2 = splitq.quant_pack(res.hidden, plan.bits, plan.row_off, O.SCALE_TOKEN, plan.nbytes)
    lay = O.payload_layout(S, D, O.SCALE_TOKEN)
    assert torch.equal(p2[:, lay[3]:], res.payload[:, lay[3]:])
    hdr = res.payload[0, :32].cpu().numpy().view(np.uint32)
    assert hdr[0] == O.MAGIC and hdr[2] == S and hdr[3] == D and hdr[6] == plan.nbytes
    # and the bytes themselves, at full size: every window's payload equals the oracle's packing of the
    # same hidden states under the kernel's own scores, and the dequantised states are bit-identical
    for b in range(B):
        want_bits, want_payload, want_h = O.boundary_roundtrip(h[b].cpu(), res.importance[b].cpu(), ladder, O.SCALE_TOKEN)
        assert np.array_equal(bits[b].cpu().numpy(), want_bits)
        assert np.array_equal(res.payload[b].cpu().numpy(), want_payload), b
        assert np.array_equal(res.hidden[b].cpu().numpy().view(np.uint32), want_h.numpy().view(np.uint32)), b


@pytest.mark.parametrize("mode", [O.SCALE_GROUP, O.SCALE_TOKEN])
def test_every_code_dequantises_like_ieee_division(mode):
    """Hand-built payloads holding every code of every width, many scales: the kernel's
    reciprocal + FMA-correction division must equal the oracle's q / L * m bit for bit."""
    import struct
    rng = np.random.default_rng(0)
    D, S = 256, 3
    bits = np.array([8, 4, 2], dtype=np.uint8)
    codes = [np.arange(-128, 128), np.tile(np.arange(-8, 8), 16), np.tile(np.arange(-2, 2), 64)]
    bits_off, rowoff_off, scales_off, codes_off = O.payload_layout(S, D, mode)
    rb = [O.row_bytes(D, int(b)) for b in bits]
    total = codes_off + sum(rb)
    payloads = []
    for trial in range(64):
        scales = np.exp(rng.uniform(-20, 20, size=3)).astype(np.float32)
        buf = np.zeros(total, dtype=np.uint8)
        n_scales = O.GROUP_SLOTS if mode == O.SCALE_GROUP else S
        buf[:32] = np.frombuffer(struct.pack("<8I", O.MAGIC, mode, S, D, n_scales, codes_off, total, 0), dtype=np.uint8)
        buf[bits_off:bits_off + S] = bits
        offs = (np.cumsum(rb) - rb) // 16
        buf[rowoff_off:rowoff_off + 4 * S] = offs.astype(np.uint32).view(np.uint8)
        sc = np.zeros(n_scales, dtype=np.float32)
        sc[:3] = scales  # slots 8/4/2-bit in group mode, tokens 0..2 in token mode
        buf[scales_off:scales_off + 4 * n_scales] = sc.view(np.uint8)
        off = codes_off
        for t in range(S):
            packed = O._pack_row(codes[t].astype(np.int8), int(bits[t]))
            buf[off:off + packed.size] = packed
            off += rb[t]
        payloads.append(buf)
    dev = torch.from_numpy(np.stack(payloads)).to(DEV)
    got = splitq.dequant(dev, S, D).cpu().numpy()
    for i, buf in enumerate(payloads):
        want = O.unpack_payload(buf).numpy()
        assert np.array_equal(got[i].view(np.uint32), want.view(np.uint32)), i


def test_codes_bit_exact_near_rounding_boundaries():
    """Values engineered to sit on and next to the k + 1/2 rounding boundaries, where the fast
    reciprocal path must defer to the exact IEEE sequence."""
    S, D = 64, 512
    gen = torch.Generator().manual_seed(11)
    for bits_w in (8, 4, 2):
        L = O.levels(bits_w)
        m = torch.rand(S, 1, generator=gen) * 10 + 0.1
        k = torch.randint(-L, L, (S, D), generator=gen).float() + 0.5
        h = k / L * m                                   # lands on a boundary up to rounding
        ulp = torch.randint(-3, 4, (S, D), generator=gen, dtype=torch.int32)
        h = (h.view(torch.int32) + ulp).view(torch.float32)
        h[:, 0] = m[:, 0]                               # pin the row absmax
        h = torch.clamp(h, -m, m).unsqueeze(0)
        imp = torch.rand(1, S, generator=gen)
        ladder = [(S, bits_w)]
        for mode in (O.SCALE_GROUP, O.SCALE_TOKEN):
            nbytes = splitq.payload_bytes(S, D, ladder, mode)
            bits, off = splitq.select_bits(imp.to(DEV), ladder, D)
            got = splitq.quant_pack(h.to(DEV), bits, off, mode, nbytes)[0].cpu().numpy()
            want = O.pack_payload(h[0], O.ladder_bits(imp[0], ladder), mode)
            assert np.array_equal(got, want), (bits_w, mode)


def test_overlapped_boundary_free_running_no_host_syncs():
    """The race surface of the software pipeline is slot reuse at depth 2 while the host runs ahead:
    60 steps on distinct inputs are submitted back to back with NO host synchronisation; every step's
    result is copied out by stream-ordered work on the side stream (as a consumer would read it) and
    compared afterwards with a plain, fully synchronised plan."""
    B, H, KVH, S, D, STEPS = 2, 4, 2, 256, 128, 60
    gen = torch.Generator(device=DEV).manual_seed(33)
    ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
    mk = lambda: splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=O.SCALE_TOKEN, KVH=KVH)  # noqa: E731
    q = torch.randn(STEPS, B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(STEPS, B, KVH, S, 64, device=DEV, generator=gen)
    h = torch.randn(STEPS, B, S, D, device=DEV, generator=gen)
    ref_plan = mk()
    want_h = torch.empty(STEPS, B, S, D, device=DEV)
    want_p = torch.empty(STEPS, B, ref_plan.nbytes, dtype=torch.uint8, device=DEV)
    for i in range(STEPS):
        r = ref_plan.run_fused(q[i], k[i], h[i])
        want_h[i].copy_(r.hidden)
        want_p[i].copy_(r.payload)
    torch.cuda.synchronize()
    for kw in ({}, {"pack_on_side": False}, {"main_priority": -3}):
        ob = splitq.OverlappedBoundary(mk, depth=2, fused=True, **kw)
        got_h = torch.zeros(STEPS, B, S, D, device=DEV)
        got_p = torch.zeros(STEPS, B, ref_plan.nbytes, dtype=torch.uint8, device=DEV)
        checksum = torch.zeros((), dtype=torch.int64, device=DEV)
        for i in range(STEPS):
            p = ob.step(q[i], k[i], h[i])
            with torch.cuda.stream(ob.side):          # stream-ordered consumer; the host never waits
                got_h[i].copy_(p.h_out)
                got_p[i].copy_(p.payload_dst)
                checksum += p.h_out.view(torch.int32).sum(dtype=torch.int64)
        ob.join()
        ob.drain()
        assert torch.equal(got_p, want_p), kw
        assert torch.equal(got_h.view(torch.int32), want_h.view(torch.int32)), kw
        assert int(checksum) == int(want_h.view(torch.int32).sum(dtype=torch.int64)), kw


def test_two_plans_on_two_streams_do_not_share_scratch():
    """Two BoundaryPlans driven concurrently on two streams (group mode, so quantise/pack uses its
    class-maximum scratch, and fused scoring, which uses the large K2 scratch): every plan owns its
    workspace, so both results equal the serial ones.  Also the module-level functions, whose default
    workspace is keyed by the current stream."""
    B, H, KVH, S, D = 3, 4, 2, 512, 256
    gen = torch.Generator(device=DEV).manual_seed(41)
    ladder = splitq.ladder_for_ratio(0.5, S)
    mk = lambda: splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=O.SCALE_GROUP, KVH=KVH)  # noqa: E731
    data = [(torch.randn(B, H, S, 64, device=DEV, generator=gen), torch.randn(B, KVH, S, 64, device=DEV, generator=gen),
             torch.randn(B, S, D, device=DEV, generator=gen) * (1 + 10 * i)) for i in range(2)]
    serial = mk()
    want = []
    for q, k, h in data:
        r = serial.run_fused(q, k, h)
        want.append((r.payload.clone(), r.hidden.clone()))
    torch.cuda.synchronize()
    plans, streams = [mk(), mk()], [torch.cuda.Stream(), torch.cuda.Stream()]
    assert plans[0].ws is not plans[1].ws
    for rep in range(20):
        for pl, st, (q, k, h) in zip(plans, streams, data):
            with torch.cuda.stream(st):
                pl.run_fused(q, k, h)
    torch.cuda.synchronize()
    for pl, (wp, wh) in zip(plans, want):
        assert torch.equal(pl.payload_dst, wp) and torch.equal(pl.h_out.view(torch.int32), wh.view(torch.int32))
    # module-level entry points on two streams: default scratch is per (device, tag, stream)
    outs = []
    for st, (q, k, h) in zip(streams, data):
        with torch.cuda.stream(st):
            for rep in range(10):
                col, _ = splitq.importance_fused(q, k, 0.125)
                bits, off = splitq.select_bits(col, ladder, D)
                payload = splitq.quant_pack(h, bits, off, O.SCALE_GROUP, serial.nbytes)
            outs.append(payload)
    torch.cuda.synchronize()
    for payload, (wp, _) in zip(outs, want):
        assert torch.equal(payload, wp)


def test_python_layer_rejects_mismatched_buffers():
    S, D = 64, 32
    h = torch.randn(1, S, D, device=DEV)
    imp = torch.rand(1, S, device=DEV)
    ladder = [(16, 4)]
    bits, off = splitq.select_bits(imp, ladder, D)
    nbytes = splitq.payload_bytes(S, D, ladder, O.SCALE_TOKEN)
    with pytest.raises(N.SplitqError):
        splitq.quant_pack(h, bits.to(torch.int32), off, O.SCALE_TOKEN, nbytes)
    with pytest.raises(N.SplitqError):
        splitq.quant_pack(h, bits, off[:, : S // 2].contiguous(), O.SCALE_TOKEN, nbytes)
    with pytest.raises(N.SplitqError):
        splitq.quant_pack(h, bits, off, O.SCALE_TOKEN, nbytes, payload=torch.empty(1, nbytes - 16, dtype=torch.uint8, device=DEV))
    q = torch.randn(1, 4, S, 64, device=DEV)
    with pytest.raises(N.SplitqError):
        splitq.importance_fused(q, q, 0.125, head_w=[1.0, 2.0])


def test_dequant_refuses_truncated_or_inconsistent_payloads():
    """A payload that crossed a process/device boundary is not trusted: wrong codes_offset, a row
    that ends past the declared size, or a declared size beyond the stride give NaN rows, not an
    out-of-bounds read."""
    S, D = 32, 64
    h = torch.randn(1, S, D, device=DEV)
    imp = torch.rand(1, S, device=DEV)
    ladder = [(8, 2), (8, 4), (8, 8)]
    nbytes = splitq.payload_bytes(S, D, ladder, O.SCALE_TOKEN)
    bits, off = splitq.select_bits(imp, ladder, D)
    good = splitq.quant_pack(h, bits, off, O.SCALE_TOKEN, nbytes)
    assert torch.isfinite(splitq.dequant(good, S, D)).all()
    for word, value in ((5, 16), (6, nbytes - 64), (6, nbytes + 4096)):
        bad = good.clone()
        hdr = bad[0, :32].view(torch.int32)
        hdr[word] = value
        out = splitq.dequant(bad, S, D)
        assert torch.isnan(out).any(), (word, value)
    rows = good.clone()
    lay = O.payload_layout(S, D, O.SCALE_TOKEN)
    rows[0, lay[1]:lay[1] + 4].view(torch.int32)[0] = 1 << 20  # row offset of token 0 far outside
    out = splitq.dequant(rows, S, D)
    assert torch.isnan(out[0, 0]).all() and torch.isfinite(out[0, 1:]).all()


def test_pipelined_front_ends_give_the_same_bytes_as_the_plain_plan():
    """OverlappedBoundary (side-stream hop + dequantise) and HostPipeline (H2D / kernels / D2H on
    three streams) are schedules, not arithmetic: outputs must equal BoundaryPlan.run bit for bit."""
    B, H, KVH, S, D = 2, 4, 2, 256, 128
    gen = torch.Generator(device=DEV).manual_seed(21)
    ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
    mk = lambda: splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=O.SCALE_TOKEN, KVH=KVH)  # noqa: E731
    steps = [(torch.randn(B, H, S, 64, device=DEV, generator=gen), torch.randn(B, KVH, S, 64, device=DEV, generator=gen),
              torch.randn(B, S, D, device=DEV, generator=gen)) for _ in range(5)]
    ref_plan = mk()
    want = []
    for q, k, h in steps:
        r = ref_plan.run_fused(q, k, h)
        want.append((r.payload.clone(), r.hidden.clone()))
    for kw in ({}, {"pack_on_side": False}, {"main_priority": -1}, {"main_priority": -2, "pack_on_side": False}):
        ob = splitq.OverlappedBoundary(mk, depth=2, fused=True, **kw)
        got = []
        for q, k, h in steps:
            p = ob.step(q, k, h)
            ob.side.synchronize()
            got.append((p.payload_dst.clone(), p.h_out.clone()))
        ob.join()
        ob.drain()
        for (wp, wh), (gp, gh) in zip(want, got):
            assert torch.equal(wp, gp) and torch.equal(wh, gh), kw
    pipe = splitq.HostPipeline(mk, depth=2, fused=True)
    outs = [torch.empty(B, S, D, pin_memory=True) for _ in steps]
    for (q, k, h), o in zip(steps, outs):
        pipe.submit([t.cpu().pin_memory() for t in (q, k, h)], o)
    pipe.drain()
    for (wp, wh), o in zip(want, outs):
        assert torch.equal(wh.cpu(), o)
