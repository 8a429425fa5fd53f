"""selection variants of initial_exp and the layer-distance analytics."""
import os

import numpy as np
import pytest
import torch

import splitq_oracle as O
from splitq import analytics, selection


def test_analytics_match_notebook_formulas():
    rng = np.random.default_rng(0)
    d = rng.random((4, 5, 64))
    d[0, 0, :7] = 0.0  # exact zeros exercise the p != 0 guard
    d /= d.sum(-1, keepdims=True)
    want = O.layer_distance_matrix_np(d)
    got = analytics.layer_distance_matrix(torch.from_numpy(d)).numpy()
    assert np.allclose(got, want, rtol=1e-12, equal_nan=True)
    p, q = torch.from_numpy(d[0, 0]), torch.from_numpy(d[1, 0])
    assert analytics.kl_divergence(p, q).item() == pytest.approx(O.kl_divergence_np(d[0, 0], d[1, 0]), rel=1e-12)
    assert analytics.jensen_shannon(p, p).item() == pytest.approx(0.0, abs=1e-15)


def test_spearman():
    from scipy.stats import spearmanr
    gen = torch.Generator().manual_seed(1)
    a, b = torch.rand(3, 200, generator=gen), torch.rand(3, 200, generator=gen)
    got = analytics.spearman(a, b)
    for i in range(3):
        assert got[i].item() == pytest.approx(spearmanr(a[i].numpy(), b[i].numpy())[0], abs=1e-12)
    assert analytics.spearman(a, a).min().item() == pytest.approx(1.0)


def test_fixed_point_top_rho_matches_reference_loop():
    """The oracle's integer-mass restatement (what the kernel computes) against the reference's
    float loop (pythia_model.py:93-106) on ordinary distributions, plus the corner cases of the rule."""
    gen = torch.Generator().manual_seed(2)
    for S in (16, 257, 4096):
        dist = torch.softmax(torch.randn(S, generator=gen) * 2, dim=0).numpy()
        for ratio in range(0, 12):
            assert O.top_rho_keep_count_fixed(dist, 1 - 0.1 * ratio) == O.top_rho_keep_count(dist, ratio)
    quarter = np.full(4, 0.25, dtype=np.float32)
    assert O.top_rho_keep_count_fixed(quarter, 0.5) == 2        # equal scores: taken one by one
    assert O.top_rho_keep_count_fixed(quarter, 0.0) == 0        # threshold already met
    assert O.top_rho_keep_count_fixed(quarter, 5.0) == 4        # never met: everything is kept
    assert O.top_rho_keep_count_fixed(np.array([0.5, 0.5, 0, 0], dtype=np.float32), 1.0) == 2
    sets = selection.quantization_sets_from_ordering(np.arange(50), [0, 3, 10])
    assert [len(s) for s in sets] == [0, 15, 50]


@pytest.mark.gpu
def test_top_rho_kernel_mode_matches_oracle():
    """splitq_select_bits_per_window with mass thresholds: the kept count equals the oracle's integer
    rule for every ratio, and the width map is exactly "all but the kept most-important tokens" with
    ties by index; then the per-window-count mode."""
    import splitq
    from splitq import boundary
    gen = torch.Generator().manual_seed(7)
    D = 64
    for S in (16, 257, 1024, 4096, 5000):
        dists = [torch.softmax(torch.randn(S, generator=gen) * 2, dim=0),
                 torch.full((S,), 1.0 / S),                                   # all equal: pure tie-breaking
                 torch.cat([torch.zeros(S // 2), torch.softmax(torch.randn(S - S // 2, generator=gen), dim=0)])]
        dists.append(torch.round(dists[0] * 64) / 64 + 1e-4)                 # few distinct values
        for dist in dists:
            ratios = list(range(0, 12))
            bits, row_off, k = selection.top_rho_select(dist.cuda(), ratios, D, bits=4, want_counts=True)
            k = k.cpu().tolist()
            for b, ratio in enumerate(ratios):
                keep = O.top_rho_keep_count_fixed(dist.numpy(), 1 - 0.1 * ratio)
                assert S - k[b] == keep, (S, ratio)
                want = O.ladder_bits(dist, [(k[b], 4)])
                assert np.array_equal(bits[b].cpu().numpy(), want), (S, ratio)
                rb = np.array([O.row_bytes(D, int(w)) for w in want], dtype=np.int64)
                assert np.array_equal(row_off[b].cpu().numpy().astype(np.int64), (np.cumsum(rb) - rb) // 16)
            assert selection.top_rho_keep_count(dist.cuda(), 3) == O.top_rho_keep_count_fixed(dist.numpy(), 0.7)
    imp = torch.rand(40, 777, generator=gen)   # more windows than one launch takes (32)
    counts = [int(x) for x in torch.randint(0, 778, (40,), generator=gen)]
    bits, _ = boundary.select_bits_per_window(imp.cuda(), 8, D, counts=counts)
    for b in range(40):
        assert np.array_equal(bits[b].cpu().numpy(), O.ladder_bits(imp[b], [(counts[b], 8)])), b
    with pytest.raises(splitq.SplitqError):
        boundary.select_bits_per_window(imp.cuda(), 8, D, counts=counts[:3])


@pytest.mark.gpu
def test_batched_entry_points_of_the_first_experiment():
    """Pythia70Model.batched_quantization_helper / batched_top_rho_quantization (pythia_model.py:71-145):
    row `ratio` of the batch equals a single-window pass whose ladder quantises the same token set."""
    from splitq import boundary, compat
    from splitq.models import Pythia70Model
    ratios = [0, 1, 2, 3, 4, 5]
    cfg = dict(compat.PYTHIA_TINY, hidden_size=256, num_attention_heads=4, intermediate_size=256)
    w = Pythia70Model.from_config("cuda:0", ratios, cfg=cfg, seed=0)
    S = 96
    ids1 = torch.randint(0, 512, (1, S), generator=torch.Generator().manual_seed(4))
    ids = ids1.expand(len(ratios), -1).contiguous()
    dist = torch.softmax(torch.randn(S, generator=torch.Generator().manual_seed(5)) * 2, dim=0)
    ordering = torch.argsort(dist, stable=True).numpy()
    for layer in (0, 2):
        got_o = w.batched_quantization_helper(ids, ordering, ids1, quant_layer=layer)
        got_t = w.batched_top_rho_quantization(ids, dist.numpy(), ids1, quant_layer=layer)
        base = w.batched_quantization_helper(ids, ordering, ids1, quant_layer=layer, do_quant=False)
        assert len(got_o) == len(got_t) == len(ratios)
        for r in ratios:
            for got, k in ((got_o, int(0.1 * r * S)), (got_t, S - O.top_rho_keep_count_fixed(dist.numpy(), 1 - 0.1 * r))):
                with torch.no_grad():
                    hidden = w.upstream(ids1, layer)
                    if k == 0:
                        want = w.downstream((1, S), ids1.cuda(), layer, hidden=hidden)
                    else:
                        bits, off = boundary.select_bits(dist.cuda(), [(k, 4)], 256)
                        payload = boundary.quant_pack(hidden, bits, off, w.scale_mode, boundary.payload_bytes(S, 256, [(k, 4)], w.scale_mode))
                        want = w.downstream((1, S), ids1.cuda(), layer, payload=payload)
                assert got[r].item() == pytest.approx(want.item(), rel=2e-5), (layer, r, k)
            assert base[r].item() == pytest.approx(base[0].item(), rel=1e-6)
    assert got_o[0].item() == pytest.approx(base[0].item(), rel=2e-5)   # ratio 0 quantises nothing


@pytest.mark.gpu
def test_layer_distance_kernel_matches_notebook_formulas():
    """analytics on CUDA tensors (the (L, N, S) importance the scoring kernels leave on the device)
    against the numpy restatement of the notebook cells; JS through splitq_layer_js."""
    import splitq
    gen = torch.Generator().manual_seed(3)
    L, N, S = 5, 3, 700
    d = torch.rand(L, N, S, generator=gen)
    d[0, 0, :9] = 0.0
    d = (d / d.sum(-1, keepdim=True)).float()
    want = O.layer_distance_matrix_np(d.double().numpy())
    got = analytics.layer_distance_matrix(d.cuda())
    assert got.is_cuda and np.allclose(got.cpu().numpy(), want, rtol=1e-10, equal_nan=True)
    # and on distributions produced by the fused scoring kernel itself
    q = torch.randn(N, 4, 512, 64, generator=gen).cuda()
    k = torch.randn(N, 2, 512, 64, generator=gen).cuda()
    layers = torch.stack([splitq.importance_fused(q * (1 + 0.3 * l), k, 0.125)[0] for l in range(4)])  # (L, N, S)
    got = analytics.layer_distance_matrix(layers)
    want = O.layer_distance_matrix_np(layers.double().cpu().numpy())
    assert np.allclose(got.cpu().numpy(), want, rtol=1e-10, equal_nan=True)
    from scipy.stats import spearmanr
    rho = analytics.spearman(layers[0], layers[3])
    for n in range(N):
        assert rho[n].item() == pytest.approx(spearmanr(layers[0, n].cpu().numpy(), layers[3, n].cpu().numpy())[0], abs=1e-9)


@pytest.mark.gpu
def test_extract_attentions_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "importance.npz"))
    want = np.load(os.path.join(golden_dir, "initial_orderings.npz"))
    A = torch.from_numpy(g["a_A"]).cuda()
    maps = [A[l] for l in range(A.shape[0])]
    keys = ["aggregate upto 2", "maximum aggregation", "upto ratio", 0, 2]
    got = selection.extract_attentions(maps, keys)
    for k in keys:
        if k == "upto ratio":
            np.testing.assert_allclose(got[k], want[str(k)], rtol=1e-5)
        else:
            # orderings agree wherever the importance gap exceeds the 1e-5 scoring tolerance
            diff = got[k] != want[str(k)]
            assert diff.mean() < 0.1
            if diff.any():
                imp = torch.from_numpy(g["a_regular_importance"])
                ref = imp[:3].mean(0) if k == "aggregate upto 2" else (imp[:3].max(0).values if k == "maximum aggregation" else imp[k])
                a, b = ref[got[k][diff]], ref[want[str(k)][diff]]
                assert ((a - b).abs() / b.abs()).max().item() < 1e-4
