"""TEST INFRASTRUCTURE -- small seeded synthetic datasets (numpy) for the oracle,
golden-vector generation and CPU/GPU parity tests.  SURVEY.md 8(d): SIFT-shape =
uint8-valued Gaussian-mixture blobs; DEEP-shape = unit-norm float blobs."""
import numpy as np


def blobs(kind, n, d, nblob, seed, means=None):
    rng = np.random.default_rng(seed)
    if means is None:
        mrng = np.random.default_rng(seed * 7919 + 13)
        if kind == "sift":
            means = mrng.gamma(2.0, 15.0, size=(nblob, d)).astype(np.float32)
        else:
            means = mrng.standard_normal((nblob, d)).astype(np.float32)
    which = rng.integers(0, means.shape[0], size=n)
    if kind == "sift":
        x = means[which] + rng.standard_normal((n, d)).astype(np.float32) * 12.0
        x = np.clip(np.rint(x), 0, 255).astype(np.float32)
    else:
        x = means[which] + rng.standard_normal((n, d)).astype(np.float32) * 0.35
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        x = x.astype(np.float32)
    return x, means


def lloyd(x, nc, iters, seed):
    """A few Lloyd iterations; any centroid file works (k-means is not on the path)."""
    rng = np.random.default_rng(seed)
    c = x[rng.choice(x.shape[0], nc, replace=False)].copy()
    for _ in range(iters):
        d2 = (x * x).sum(1)[:, None] - 2 * x @ c.T + (c * c).sum(1)[None]
        a = d2.argmin(1)
        for j in range(nc):
            m = a == j
            if m.any():
                c[j] = x[m].mean(0)
    return np.ascontiguousarray(c, np.float32)


def dataset(kind, nb, nq, nt, nc, d, seed):
    nblob = max(nc // 4, 4)
    base, means = blobs(kind, nb, d, nblob, seed)
    queries, _ = blobs(kind, nq, d, nblob, seed + 1, means)
    learn, _ = blobs(kind, nt, d, nblob, seed + 2, means)
    centroids = lloyd(learn, nc, 4, seed + 3)
    return base, queries, learn, centroids
