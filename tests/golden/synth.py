"""Seeded synthetic features shared by the golden generators (authoring container) and the tests (everywhere):
the fixtures store only metadata and the reference's outputs, the features are regenerated from the seed.
np.random.RandomState is bit-stable across machines and numpy versions."""
import numpy as np


def class_feats(seed, label_ids, dim, spread=0.9):
    """Pooled-feature look-alikes: randn + spread * centre[label], row-normalised, scaled by
    U(0.6, 1) (mean over T of unit vectors has norm <= 1)."""
    rng = np.random.RandomState(seed)
    label_ids = np.asarray(label_ids)
    centre = rng.randn(int(label_ids.max()) + 2, dim).astype(np.float32)
    x = rng.randn(len(label_ids), dim).astype(np.float32) + np.float32(spread) * centre[label_ids]
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    x *= rng.uniform(0.6, 1.0, (len(label_ids), 1)).astype(np.float32)
    return np.ascontiguousarray(x, np.float32)
