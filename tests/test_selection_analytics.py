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


def test_top_rho_count_matches_reference_loop():
    gen = torch.Generator().manual_seed(2)
    for S in (16, 257):
        dist = torch.softmax(torch.randn(S, generator=gen) * 2, dim=0)
        for ratio in range(0, 11):
            assert selection.top_rho_keep_count(dist, ratio) == O.top_rho_keep_count(dist.numpy(), ratio)
    assert selection.top_rho_ladder(dist, 0) in ([], [(0, 4)])
    k = selection.top_rho_ladder(dist, 5)[0][0]
    assert k == 257 - O.top_rho_keep_count(dist.numpy(), 5)
    sets = selection.quantization_sets_from_ordering(np.arange(50), [0, 3, 10])
    assert [len(s) for s in sets] == [0, 15, 50]


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
