"""Seeded parity cases shared by the golden generator, the CPU tests and the GPU tests."""
import numpy as np

from hgpumap import synth


def blobs(n, d, k, seed=20261106, first=0):
    """The benchmark workload at a small size: (points[n,d], centroids[k,d])."""
    centres = synth.centres(seed, k, d)
    cent = synth.perturb(seed + 1, centres)
    return synth.points(seed, first, n, centres), cent


def adversarial_2d(seed=7):
    """d=2 edge cases the domain has: duplicated centroids, points
    exactly midway between two centroids, +-0, denormals, NaN/Inf rows, points far outside
    the centroids' box, k not a multiple of 32, n not a multiple of any tile."""
    rng = np.random.default_rng(seed)
    k = 37
    cent = rng.uniform(-50, 50, size=(k, 2)).astype(np.float32)
    cent[5] = cent[3]            # duplicate: lowest index must win
    cent[20] = cent[3]
    cent[9] = (0.0, 0.0)
    cent[10] = (-0.0, 0.0)       # -0 == +0: index 9 wins
    cent[11] = (2.0, 0.0)
    cent[12] = (4.0, 0.0)
    rows = [rng.uniform(-60, 60, size=(5000, 2)).astype(np.float32)]
    mid = ((cent[:-1].astype(np.float64) + cent[1:].astype(np.float64)) / 2).astype(np.float32)
    rows.append(mid)                                   # (nearly) equidistant pairs
    rows.append(np.array([[3.0, 0.0], [1.0, 0.0], [3.0, 1e-30], [1.0, -1e-30]], np.float32))  # exact ties
    rows.append(np.repeat(cent, 3, axis=0))            # points on centroids
    rows.append(np.array([[0.0, -0.0], [-0.0, -0.0], [1e-40, -1e-40], [1e-45, 0], [-1e-38, 1e-38]], np.float32))
    rows.append(np.array([[np.nan, 1.0], [1.0, np.nan], [np.nan, np.nan], [np.inf, 0.0], [-np.inf, np.inf],
                          [0.0, np.inf], [3.0e38, 3.0e38], [-3.0e38, 1.0], [1e20, -1e20], [1e10, 1e10],
                          [1e5, -1e5], [500.0, 499.0], [-61.0, 0.0]], np.float32))
    rows.append(rng.normal(0, 1e-3, size=(2000, 2)).astype(np.float32) + cent[rng.integers(0, k, 2000)])
    pts = np.concatenate(rows).astype(np.float32)
    rng.shuffle(pts, axis=0)
    return pts, cent


def tiny_scale_2d(seed=11):
    """All coordinates ~1e-25: every FP32 square underflows to 0, so the scan returns index 0
    for every point although the exact distances differ -- pruning must not 'improve' on it."""
    rng = np.random.default_rng(seed)
    cent = (rng.uniform(-1, 1, size=(16, 2)) * 1e-25).astype(np.float32)
    pts = (rng.uniform(-1, 1, size=(3000, 2)) * 1e-25).astype(np.float32)
    return pts, cent


def close_centroids_2d(seed=13):
    """Many centroids packed far below the cell size: every cell lists > 8 candidates."""
    rng = np.random.default_rng(seed)
    cent = np.concatenate([rng.normal(0, 1e-3, size=(40, 2)), rng.uniform(-100, 100, size=(24, 2))]).astype(np.float32)
    pts = np.concatenate([rng.normal(0, 2e-3, size=(4000, 2)), rng.uniform(-120, 120, size=(4000, 2))]).astype(np.float32)
    return pts, cent


def degenerate_axis_2d():
    """All centroids share one x: zero extent on an axis."""
    cent = np.stack([np.full(9, 3.5, np.float32), np.linspace(-4, 4, 9, dtype=np.float32)], axis=1)
    rng = np.random.default_rng(3)
    pts = rng.uniform(-6, 6, size=(3001, 2)).astype(np.float32)
    pts[:500, 0] = 3.5
    return pts, cent


def skewed_2d(n, k=128, share=0.9, seed=17):
    """One cluster holds `share` of the points (the rest are spread over the others): its partial sums take
    far more adds than any word of the fixed-point combiner may hold between two flushes, and its column
    overflows the counting-sort combiner's striped rows."""
    rng = np.random.default_rng(seed)
    cent = rng.uniform(-100, 100, size=(k, 2)).astype(np.float32)
    which = np.where(rng.random(n) < share, 5 % k, rng.integers(0, k, n))
    pts = (cent[which] + rng.normal(0, 0.05, size=(n, 2))).astype(np.float32)
    return pts, cent


def tiny_values_2d(seed=19):
    """Coordinates far below the grid's extent next to ordinary ones: values the fixed-point combiner's two limbs
    cannot hold (non-zero, under 2^-20 of the extent) must take its FP64 path; exact zeros must not need to."""
    rng = np.random.default_rng(seed)
    cent = rng.uniform(-100, 100, size=(64, 2)).astype(np.float32)
    cent[0] = (0.0, 0.0)
    cent[1] = (0.0, 40.0)
    cent[2] = (-35.0, 0.0)
    n = 40000
    pts = rng.uniform(-110, 110, size=(n, 2)).astype(np.float32)
    scale = 10.0 ** rng.uniform(-44, -3, size=n)
    pts[: n // 4, 0] = (rng.normal(0, 1, n // 4) * scale[: n // 4]).astype(np.float32)       # tiny x, any y
    pts[n // 4 : n // 2, 1] = (rng.normal(0, 1, n // 4) * scale[n // 4 : n // 2]).astype(np.float32)  # tiny y
    pts[n // 2 : n // 2 + 3000] = 0.0                                                          # exact zeros
    pts[n // 2 + 3000 : n // 2 + 6000, 0] = 0.0
    pts[n // 2 + 6000 : n // 2 + 9000, 1] = -0.0
    rng.shuffle(pts, axis=0)
    return pts, cent


def general_d(n, d, k, seed):
    rng = np.random.default_rng(seed)
    cent = rng.normal(0, 4, size=(k, d)).astype(np.float32)
    which = rng.integers(0, k, n)
    pts = (cent[which] + rng.normal(0, 1.5, size=(n, d))).astype(np.float32)
    return pts, cent
