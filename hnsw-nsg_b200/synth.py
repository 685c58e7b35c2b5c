"""Seeded synthetic datasets in the shapes BASELINE.json names. Pure numpy; no CUDA.
Used by bench.py and by the tests so that every arm sees identical inputs."""
import numpy as np


def sift_shape(n, dim=128, seed=1, latent=16, basis_seed=777, chunk=1 << 18):
    """SIFT-shaped rows: integer-valued fp32 in [0,218], non-negative and heavy-tailed like SIFT
    histograms, low intrinsic dimension (`latent` Gaussian factors mixed by a fixed basis + small
    isotropic noise) so that the kNN graph is connected the way real descriptor sets are.
    Every value is an integer <= 218: exactly representable in bf16, and every L2 partial sum is an
    integer < 2^24, so fp32 distances are exact under ANY summation order (CPU SIMD, GPU, tensor core)."""
    W = np.random.default_rng(basis_seed).normal(0, 1, (latent, dim)).astype(np.float32) / np.sqrt(latent)
    rng = np.random.default_rng(seed)
    out = np.empty((n, dim), np.float32)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        z = rng.standard_normal((b - a, latent), dtype=np.float32)
        y = np.abs(z @ W) * 45.0 + rng.standard_normal((b - a, dim), dtype=np.float32) * 6.0
        out[a:b] = np.floor(np.clip(np.abs(y), 0, 218))
    return out


def _factor_rows(n, dim, seed, latent, basis_seed, noise, chunk=1 << 18):
    W = np.random.default_rng(basis_seed).normal(0, 1, (latent, dim)).astype(np.float32) / np.sqrt(latent)
    rng = np.random.default_rng(seed)
    out = np.empty((n, dim), np.float32)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        z = rng.standard_normal((b - a, latent), dtype=np.float32)
        out[a:b] = z @ W + rng.standard_normal((b - a, dim), dtype=np.float32) * noise
    return out


def glove_shape(n, dim=100, seed=1, latent=24):
    """GloVe-shaped rows: real-valued, L2-normalised (metric: inner product, 1 - <a,b>)."""
    x = _factor_rows(n, dim, seed, latent, 778, 0.25)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x


def gist_shape(n, dim=960, seed=1, latent=32):
    """GIST-shaped rows: real-valued in [0,1), high dimensional (bandwidth-bound gathers)."""
    x = _factor_rows(n, dim, seed, latent, 779, 0.15)
    return np.clip(0.25 + 0.12 * x, 0.0, 0.999).astype(np.float32)


def deep_shape(n, dim=96, seed=1, latent=20):
    """Deep1B-shaped rows: real-valued, L2-normalised, metric L2."""
    x = _factor_rows(n, dim, seed, latent, 780, 0.2)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x
