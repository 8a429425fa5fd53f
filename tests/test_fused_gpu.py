"""K2 (fused tcgen05 importance from Q,K) against the oracle's eager attention + reference
reductions.  Tolerance: importance <= 1e-5 relative in the three-product mode (BASELINE.json north_star)."""
import pytest
import torch

import splitq
import splitq_oracle as O
from splitq import _native as N

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(autouse=True, params=["by shape", "2", "3"], ids=lambda v: f"k2-groups={v}")
def k2_build(request, monkeypatch):
    """The library holds two builds of K2 (two / three epilogue warpgroups, csrc/importance_fused_api.cu) and
    picks one per call by shape; every case of this module also runs pinned to each of them."""
    if request.param == "by shape":
        monkeypatch.delenv("SPLITQ_K2_GROUPS", raising=False)
    else:
        monkeypatch.setenv("SPLITQ_K2_GROUPS", request.param)
    return request.param


def make_qk(gen, B, H, KVH, S, sigma, hf_layout=False):
    hd = 64
    if hf_layout:  # (B,S,H,hd) storage viewed as (B,H,S,hd), like q_proj(...).view(...).transpose(1, 2)
        q = (torch.randn(B, S, H, hd, generator=gen) * sigma).transpose(1, 2)
        k = (torch.randn(B, S, KVH, hd, generator=gen) * sigma).transpose(1, 2)
    else:
        q = torch.randn(B, H, S, hd, generator=gen) * sigma
        k = torch.randn(B, KVH, S, hd, generator=gen) * sigma
    return q, k


def oracle_scores(q, k, causal):
    A = O.attention_probs(q.contiguous(), k.contiguous(), 64 ** -0.5, causal)
    col = torch.stack([O.importance_regular(A[b:b + 1]) for b in range(A.shape[0])])
    last = torch.stack([O.importance_last_row(A[b:b + 1]) for b in range(A.shape[0])])
    return col, last


def max_rel(got, want):
    return ((got - want).abs() / want.abs().clamp_min(1e-30)).max().item()


@pytest.mark.parametrize("B,H,KVH,S", [(1, 2, 1, 128), (1, 4, 2, 256), (2, 8, 8, 384), (1, 14, 2, 512),
                                        (1, 3, 3, 200), (2, 2, 1, 77), (1, 2, 2, 1000), (1, 1, 1, 1)])
@pytest.mark.parametrize("sigma", [0.4, 1.0, 1.5])
def test_fused_matches_eager_attention_importance(B, H, KVH, S, sigma):
    gen = torch.Generator().manual_seed(S * 31 + H)
    q, k = make_qk(gen, B, H, KVH, S, sigma, hf_layout=(S % 2 == 0))
    col, last = splitq.importance_fused(q.to(DEV), k.to(DEV), 64 ** -0.5, causal=True, want_last=True)
    torch.cuda.synchronize()
    want_col, want_last = oracle_scores(q, k, True)
    assert max_rel(col.cpu(), want_col) <= 1e-5
    assert max_rel(last.cpu(), want_last) <= 1e-5


def test_fused_non_causal_and_head_weights():
    gen = torch.Generator().manual_seed(5)
    q, k = make_qk(gen, 1, 4, 2, 300, 1.0)
    w = [0.5, -0.25, 1.5, 0.1]
    col, last = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, causal=False, head_w=w, want_last=True)
    A = O.attention_probs(q, k, 0.125, causal=False)
    want = O.importance_weighted(A, w)
    assert ((col[0].cpu() - want).abs().max() / want.abs().max()).item() <= 1e-5
    want_last = sum(w[h] * A[0, h, -1] for h in range(4))
    assert ((last[0].cpu() - want_last).abs().max() / want_last.abs().max()).item() <= 1e-5


def test_fused_single_product_mode_is_close_but_looser():
    gen = torch.Generator().manual_seed(6)
    q, k = make_qk(gen, 1, 4, 4, 512, 1.0)
    col, _ = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, mode=N.FUSED_F16X1)
    want, _ = oracle_scores(q, k, True)
    assert max_rel(col.cpu(), want) <= 2e-2


def test_fused_equals_materialised_kernel_on_gpu_attention():
    """K2 vs fp64 eager probabilities at a size the CPU oracle would be slow on."""
    gen = torch.Generator(device=DEV).manual_seed(9)
    B, H, KVH, S = 1, 14, 2, 2048
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    kk = k.repeat_interleave(H // KVH, dim=1)
    s = torch.matmul(q.double(), kk.double().transpose(2, 3)) * 0.125
    s = s + torch.full((S, S), float("-inf"), device=DEV, dtype=torch.float64).triu(1)
    A = torch.softmax(s, dim=-1)
    want_col, want_last = A.mean(dim=1).mean(dim=1).float(), A[:, :, -1].mean(dim=1).float()
    col, last = splitq.importance_fused(q, k, 0.125, want_last=True)
    assert max_rel(col, want_col) <= 1e-5
    assert max_rel(last, want_last) <= 1e-5


def test_fused_rejects_other_head_dims():
    q = torch.zeros(1, 1, 8, 32, device=DEV)
    with pytest.raises(N.SplitqError):
        splitq.importance_fused(q, q, 0.1)


@pytest.mark.parametrize("B,H,KVH,S", [(1, 8, 8, 2048), (2, 14, 2, 333), (1, 2, 1, 1)])
def test_last_row_only_uses_the_gemv_path_and_matches(B, H, KVH, S):
    """want_col=False: one query row per head (north star: "single-query GEMV against K")."""
    gen = torch.Generator().manual_seed(S)
    q, k = make_qk(gen, B, H, KVH, S, 1.0, hf_layout=True)
    col, last = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, want_col=False, want_last=True)
    assert col is None
    _, want_last = oracle_scores(q, k, True)
    assert max_rel(last.cpu(), want_last) <= 1e-5
    both_col, both_last = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, want_col=True, want_last=True)
    assert max_rel(both_last.cpu(), want_last) <= 1e-5


GENERAL_PATH = 32 << 4  # undocumented mode bit: never take the bounded-logit (FAST) epilogue


def test_fused_beyond_fp16_range_is_prescaled_not_nan():
    """The tensor-core products run on fp16 operands.  A 128-row block whose largest |x| reaches 2^15 is
    multiplied by a power of two before the conversion and the epilogue multiplies it back per tile,
    so activations beyond 65504 score correctly instead of overflowing to NaN."""
    gen = torch.Generator().manual_seed(1)
    q, k = make_qk(gen, 1, 2, 2, 256, 1.0)
    q[0, 0, 5, 3] = 1.0e5
    k[0, 1, 200, 7] = -3.0e6
    col, last = splitq.importance_fused(q.to(DEV), k.to(DEV), 1e-6, want_last=True)
    A = O.attention_probs(q, k, 1e-6, True)
    want_col, want_last = O.importance_regular(A), O.importance_last_row(A)
    assert torch.isfinite(col).all()
    assert max_rel(col[0].cpu(), want_col) <= 1e-5
    assert max_rel(last[0].cpu(), want_last) <= 1e-5
    col1, _ = splitq.importance_fused(q.to(DEV), k.to(DEV), 1e-6, mode=N.FUSED_F16X1)
    assert torch.isfinite(col1).all() and max_rel(col1[0].cpu(), want_col) <= 2e-2
    # non-finite inputs stay loud
    q[0, 1, 7, 0] = float("inf")
    bad, _ = splitq.importance_fused(q.to(DEV), k.to(DEV), 1e-6)
    assert not torch.isfinite(bad).all()


@pytest.mark.parametrize("B,H,KVH,S,sigma", [(1, 4, 2, 512, 1.0), (2, 6, 3, 333, 1.2), (1, 14, 2, 1024, 2.0)])
def test_fused_general_epilogue_forced(B, H, KVH, S, sigma):
    """Inputs whose norms prove |logit| < 100 take the epilogue without a running maximum; the general
    epilogue (online maximum, per-tile scale) is forced here on the same inputs and must agree with
    the oracle and, to rounding, with the bounded path."""
    gen = torch.Generator().manual_seed(S + H)
    q, k = make_qk(gen, B, H, KVH, S, sigma)
    # against fp64 attention: at sigma = 2 the fp32 CPU oracle is itself 2-4e-6 away from it, by an amount that
    # depends on the host's BLAS threading, which made a 1e-5 bar against that oracle pass on some boxes only
    want_col, want_last = fp64_scores(q.to(DEV), k.to(DEV))
    col_g, last_g = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, want_last=True, mode=N.FUSED_F16X3 | GENERAL_PATH)
    col_f, last_f = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, want_last=True)
    for got_c, got_l in ((col_g, last_g), (col_f, last_f)):
        assert max_rel(got_c, want_col.float()) <= 1e-5
        assert max_rel(got_l, want_last.float()) <= 1e-5
    assert max_rel(col_g, col_f) <= 2e-6


@pytest.mark.parametrize("q_sigma,k_sigma", [(0.1, 20.0), (100.0, 0.02), (3.0, 3.0)])
def test_fused_large_norms_take_the_general_epilogue(q_sigma, k_sigma):
    """Operand norms beyond the caps of the bounded-logit path (|k| = 160, or |q * scaling * log2 e| =
    144, or their product above 100): the general epilogue runs by itself.  Against fp64; the
    (3, 3) case has logits of standard deviation 9 (|s| up to ~45), where an fp32 matmul is itself
    only good to ~1e-5 on the smallest probabilities, hence the looser bar there."""
    B, H, KVH, S = 1, 4, 2, 768
    gen = torch.Generator(device=DEV).manual_seed(int(q_sigma * 10))
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen) * q_sigma
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen) * k_sigma
    want_col, want_last = fp64_scores(q, k)
    col, last = splitq.importance_fused(q, k, 0.125, want_last=True)
    tol = 1e-5 if q_sigma * k_sigma < 4 else 1e-4
    assert max_rel(col, want_col.float()) <= tol
    assert max_rel(last, want_last.float()) <= tol


@pytest.mark.parametrize("S", [127, 129, 255, 257, 640])
@pytest.mark.parametrize("causal", [True, False])
def test_fused_block_boundaries_and_masked_batches(S, causal):
    """Sequence lengths around the 128-row block size: the ragged last block meets the causal
    diagonal in the same tile, which is where the per-batch skip / plain / masked choice of the
    epilogue and the single-column last-row pick can go wrong."""
    gen = torch.Generator().manual_seed(S + int(causal))
    q, k = make_qk(gen, 2, 6, 3, S, 1.2, hf_layout=(S % 2 == 1))
    col, last = splitq.importance_fused(q.to(DEV), k.to(DEV), 64 ** -0.5, causal=causal, want_last=True)
    want_col, want_last = oracle_scores(q, k, causal)
    assert max_rel(col.cpu(), want_col) <= 1e-5
    assert max_rel(last.cpu(), want_last) <= 1e-5


def test_fused_is_bit_reproducible_and_schedule_independent():
    """Partial results meet in fixed-order slots (no atomics): repeated calls give identical bits, and
    so does the same window scored alone or inside a batch, although the batch changes how the
    tile sequence is cut into CTA ranges."""
    gen = torch.Generator(device=DEV).manual_seed(11)
    B, H, KVH, S = 5, 6, 2, 1536
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    col0, last0 = splitq.importance_fused(q, k, 0.125, want_last=True)
    for _ in range(3):
        col, last = splitq.importance_fused(q, k, 0.125, want_last=True)
        assert torch.equal(col, col0) and torch.equal(last, last0)
    for b in (0, 3):
        col1, last1 = splitq.importance_fused(q[b:b + 1], k[b:b + 1], 0.125, want_last=True)
        # the split of an item between two CTAs moves, so the (max, sum) pairs are merged at other
        # points of the row: equal to fp32 rounding of the merge, not to the bit
        assert max_rel(col1[0], col0[b]) <= 2e-6 and max_rel(last1[0], last0[b]) <= 2e-6


@pytest.mark.parametrize("B,H,KVH,S,causal", [(1, 2, 1, 8192, True), (3, 4, 4, 4096, True), (1, 2, 2, 3000, False)])
def test_fused_long_windows_against_fp64_attention(B, H, KVH, S, causal):
    """Long windows (many CTA ranges, items cut between CTAs, 64 blocks per row) against fp64 eager
    attention computed on the GPU head by head."""
    gen = torch.Generator(device=DEV).manual_seed(S + B)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    want_col = torch.zeros(B, S, device=DEV, dtype=torch.float64)
    want_last = torch.zeros(B, S, device=DEV, dtype=torch.float64)
    mask = torch.full((S, S), float("-inf"), device=DEV, dtype=torch.float64).triu(1) if causal else 0.0
    for b in range(B):
        for h in range(H):
            s = q[b, h].double() @ k[b, h // (H // KVH)].double().T * 0.125 + mask
            A = torch.softmax(s, dim=-1)
            want_col[b] += A.mean(dim=0) / H
            want_last[b] += A[-1] / H
    col, last = splitq.importance_fused(q, k, 0.125, causal=causal, want_last=True)
    assert max_rel(col, want_col.float()) <= 1e-5
    assert max_rel(last, want_last.float()) <= 1e-5


def fp64_scores(q, k, causal=True, scaling=0.125):
    """fp64 eager attention on the GPU, head by head -> (column mean, last row), both (B,S) fp64."""
    B, H, S, _ = q.shape
    KVH = k.shape[1]
    want_col = torch.zeros(B, S, device=q.device, dtype=torch.float64)
    want_last = torch.zeros(B, S, device=q.device, dtype=torch.float64)
    mask = torch.full((S, S), float("-inf"), device=q.device, dtype=torch.float64).triu(1) if causal else 0.0
    for b in range(B):
        for h in range(H):
            s = q[b, h].double() @ k[b, h // (H // KVH)].double().T * scaling + mask
            A = torch.softmax(s, dim=-1)
            want_col[b] += A.mean(dim=0) / H
            want_last[b] += A[-1] / H
    return want_col, want_last


@pytest.mark.parametrize("B,H,KVH,S", [(1, 14, 2, 1024), (2, 8, 8, 640), (1, 4, 2, 4096)])
def test_fused_peaky_logits_sigma_2(B, H, KVH, S):
    """SURVEY appendix A.2's failure regime for single-product operands: q,k ~ N(0, 2^2), logits with
    standard deviation 4 and rows dominated by one or two keys.  The three-product split must still
    meet 1e-5 against fp64 (a plain fp32 matmul measures 2.3e-6 / 4.0e-6 there)."""
    gen = torch.Generator(device=DEV).manual_seed(S + H)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen) * 2.0
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen) * 2.0
    want_col, want_last = fp64_scores(q, k)
    col, last = splitq.importance_fused(q, k, 0.125, want_last=True)
    assert max_rel(col, want_col.float()) <= 1e-5
    assert max_rel(last, want_last.float()) <= 1e-5


def test_fused_bench_shape_against_fp64():
    """The exact shape bench.py times (BASELINE.json configs[2]: 8 Qwen2-0.5B windows, H=14, KVH=2,
    S=4096) against fp64 eager attention."""
    B, H, KVH, S = 8, 14, 2, 4096
    gen = torch.Generator(device=DEV).manual_seed(77)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    col, last = splitq.importance_fused(q, k, 0.125, want_last=True)
    want_col, want_last = fp64_scores(q, k)
    assert max_rel(col, want_col.float()) <= 1e-5
    assert max_rel(last, want_last.float()) <= 1e-5


def test_the_two_builds_agree_to_rounding_and_share_one_workspace_size(monkeypatch):
    """Same inputs through both builds: different partial-sum orders (2 x groups slots), so equal to fp32
    rounding, not bitwise; the workspace query covers whichever build a call is handed to."""
    q, k = make_qk(torch.Generator().manual_seed(21), 2, 4, 2, 1000, 1.0)
    got = {}
    for g in ("2", "3"):
        monkeypatch.setenv("SPLITQ_K2_GROUPS", g)
        got[g] = splitq.importance_fused(q.to(DEV), k.to(DEV), 0.125, want_last=True)
    assert max_rel(got["2"][0], got["3"][0]) < 2e-6 and max_rel(got["2"][1], got["3"][1]) < 2e-6
    lib = N.load()
    need = lib.splitq_importance_fused_workspace_bytes(2, 4, 2, 1000, 64)
    ws = torch.empty(need, dtype=torch.uint8, device=DEV)
    col = torch.empty(2, 1000, device=DEV)
    qd, kd = q.to(DEV), k.to(DEV)
    import ctypes
    for g in ("2", "3"):
        monkeypatch.setenv("SPLITQ_K2_GROUPS", g)
        qs = (ctypes.c_longlong * 3)(*qd.stride()[:3]); ks = (ctypes.c_longlong * 3)(*kd.stride()[:3])
        N.check(lib.splitq_importance_fused(qd.data_ptr(), qs, kd.data_ptr(), ks, 2, 4, 2, 1000, 64, 0.125, 1,
                                            N.FUSED_F16X3, None, col.data_ptr(), None, ws.data_ptr(), need,
                                            torch.cuda.current_stream().cuda_stream))
        assert max_rel(col, got[g][0]) == 0.0


@pytest.mark.parametrize("causal", [True, False])
def test_fused_mixed_items_some_bounded_some_not(causal):
    """Pass 1 converts the queries itself and picks its epilogue body per 128-row query block from that
    block's own norm and exponent: one launch here holds ordinary blocks (bounded-logit body), a block
    with rows of large norm (general body, online maximum), a block beyond the fp16 range (pre-scaled and
    converted a second time) and a block of tiny values (scaled up), next to each other and in one head."""
    B, H, KVH, S = 2, 4, 2, 1024
    gen = torch.Generator(device=DEV).manual_seed(77)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    q[0, 0, 3 * 128:4 * 128] *= 12.0          # |q'| ~ 17 x 8: beyond the product cap with |k| ~ 8
    q[0, 1, 5 * 128 + 7, 11] = 9.0e4          # beyond fp16 once scaled: this block gets an exponent
    q[1, 2, 2 * 128:3 * 128] *= 1.0e-6        # far below the fp16 normal range: scaled up
    q[1, 3, 6 * 128:7 * 128] *= 12.0
    want_col, want_last = fp64_scores(q, k, causal=causal)
    col, last = splitq.importance_fused(q, k, 0.125, causal=causal, want_last=True)
    assert torch.isfinite(col).all()
    assert max_rel(col, want_col.float()) <= 2e-5
    assert max_rel(last, want_last.float()) <= 2e-5


@pytest.mark.parametrize("offset", [0, 4, 1])
def test_fused_query_alignments(offset):
    """The loaders read the caller's fp32 queries in 32-byte pieces, as float4 or scalar by scalar,
    whichever the base pointer and the strides allow; same values, same result bits."""
    B, H, KVH, S = 1, 4, 2, 640
    gen = torch.Generator(device=DEV).manual_seed(5)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    ref_col, ref_last = splitq.importance_fused(q, k, 0.125, want_last=True)
    store = torch.empty(q.numel() + 8, device=DEV)
    q_off = store[offset:offset + q.numel()].view_as(q)
    q_off.copy_(q)
    assert q_off.data_ptr() % 32 == (offset * 4) % 32
    col, last = splitq.importance_fused(q_off, k, 0.125, want_last=True)
    assert torch.equal(col, ref_col) and torch.equal(last, ref_last)


@pytest.mark.parametrize("B,H,KVH,S,causal", [(1, 2, 1, 1, True), (3, 2, 2, 100, True), (2, 4, 2, 300, False),
                                              (1, 6, 3, 1100, True), (5, 6, 2, 1536, True), (40, 4, 4, 130, True)])
def test_fused_fills_every_result_slot_it_reads(B, H, KVH, S, causal):
    """The merge kernels read 2 x groups partial-result slots per row; slots that no epilogue group
    fills (short runs, items that are not cut) get their neutral element from the loader warps of the
    CTA that owns the run.  A workspace full of NaN bit patterns must therefore give the same bits as
    a workspace that was used before."""
    import ctypes
    gen = torch.Generator(device=DEV).manual_seed(B * 1000 + S)
    q = torch.randn(B, H, S, 64, device=DEV, generator=gen)
    k = torch.randn(B, KVH, S, 64, device=DEV, generator=gen)
    ref_col, ref_last = splitq.importance_fused(q, k, 0.125, causal=causal, want_last=True)
    lib = N.load()
    need = lib.splitq_importance_fused_workspace_bytes(B, H, KVH, S, 64)
    qs = (ctypes.c_longlong * 3)(*q.stride()[:3]); ks = (ctypes.c_longlong * 3)(*k.stride()[:3])
    for fill in (0xFF, 0x7F):
        ws = torch.full((need,), fill, dtype=torch.uint8, device=DEV)
        col = torch.empty(B, S, device=DEV); last = torch.empty(B, S, device=DEV)
        N.check(lib.splitq_importance_fused(q.data_ptr(), qs, k.data_ptr(), ks, B, H, KVH, S, 64, 0.125, int(causal),
                                            N.FUSED_F16X3, None, col.data_ptr(), last.data_ptr(), ws.data_ptr(), need,
                                            torch.cuda.current_stream().cuda_stream))
        assert torch.equal(col, ref_col) and torch.equal(last, ref_last)
