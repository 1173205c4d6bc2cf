"""The numpy oracle of the f3/f4 rows against the vectors minted from the reference's own loss modules
(tests/golden/make_golden_losses.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import oracle_losses as ol

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gl():
    return np.load(os.path.join(HERE, "golden", "reference_golden_losses.npz"))


def test_margin_loss_matches_reference(gl):
    emb, labels = gl["emb"], gl["labels"]
    trip = gl["mg_dist_trip"]
    loss, g, gb = ol.margin_given(emb, trip, gl["mg_dist_beta"][labels[trip[:, 0]]], 0.2)
    assert abs(loss - float(gl["mg_dist_loss"])) < 1e-6
    np.testing.assert_allclose(g, gl["mg_dist_grad"], rtol=0, atol=1e-6)
    gbeta = np.zeros(12)
    np.add.at(gbeta, labels[trip[:, 0]], gb)
    np.testing.assert_allclose(gbeta, gl["mg_dist_gbeta"], rtol=0, atol=1e-6)
    loss, g, _ = ol.margin_given(emb, gl["mg_semi_trip"], 0.9, 0.2)
    assert abs(loss - float(gl["mg_semi_loss"])) < 1e-6
    np.testing.assert_allclose(g, gl["mg_semi_grad"], rtol=0, atol=1e-6)


def test_npair_loss_matches_reference(gl):
    loss, g = ol.npair_given(gl["emb"] * np.float32(gl["np_scale"]), gl["np_tuples"], 0.02)
    assert abs(loss - float(gl["np_loss"])) < 2e-6
    np.testing.assert_allclose(g, gl["np_grad"], rtol=0, atol=1e-6)


def test_proxynca_matches_reference(gl):
    loss, g, gp = ol.proxynca(gl["emb"] * np.float32(gl["px_scale"]), gl["px_proxies"], gl["labels"])
    assert abs(loss - float(gl["px_loss"])) < 2e-6
    np.testing.assert_allclose(g, gl["px_grad"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(gp, gl["px_gproxies"], rtol=0, atol=1e-6)


def test_memory_logits_match_restated_lines(gl):
    logits, gx = ol.memory_logits(gl["ml_x"], gl["ml_mem"], float(gl["ml_temp"]), gl["ml_w"])
    np.testing.assert_allclose(logits, gl["ml_logits"], rtol=0, atol=2e-5)   # logits ~ 20..30 in fp32
    assert np.isfinite(gl["ml_gx"]).all()
    np.testing.assert_allclose(gx, gl["ml_gx"], rtol=0, atol=2e-5)


def test_sphere_weights_match_reference(gl):
    import torch  # only to restate pdist's clamp exactly as the sampler feeds it
    emb, labels = gl["emb"].astype(np.float64), gl["labels"]
    prod = emb @ emb.T
    n2 = np.diag(prod)
    res = np.clip(n2[:, None] + n2[None, :] - 2 * prod, 0, None)
    dist = np.clip(np.sqrt(np.clip(res, 1e-4, None)), 0.5, None)
    for row, i in enumerate((0, 17)):
        q = ol.inverse_sphere_distances(dist[i], labels, labels[i], emb.shape[1])
        np.testing.assert_allclose(q, gl["sphere_q"][row], rtol=2e-3, atol=1e-9)
    del torch
