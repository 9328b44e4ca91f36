"""The benchmark's synthetic workload restated in numpy.

Same integer hashing and the same single-rounded FP32 operations as csrc/synth.cu, so
the three implementations (numpy here, C in the library, CUDA on the device) agree bit
for bit; tests pin them against each other and against tests/golden/.
"""
from __future__ import annotations

import numpy as np

SEED_POINTS = 20261106
SEED_CENTROIDS = 20261107

_GOLD = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
IH8_SCALE = np.float32(float.fromhex("0x1.3988e2p-16"))
CENTRE_STEP = np.float32(float.fromhex("0x1.9p-17"))


def mix64(seed: int, ctr: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + (ctr.astype(np.uint64) + np.uint64(1)) * _GOLD
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _ih8(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    s = np.zeros(a.shape, dtype=np.int64)
    for i in range(4):
        sh = np.uint64(16 * i)
        s += ((a >> sh) & np.uint64(0xFFFF)).astype(np.int64)
        s += ((b >> sh) & np.uint64(0xFFFF)).astype(np.int64)
    return (s - 262140).astype(np.int32)


def centres(seed: int, k: int, d: int) -> np.ndarray:
    u = (mix64(seed, np.arange(k * d, dtype=np.uint64)) >> np.uint64(40)).astype(np.uint32)
    f = u.astype(np.float32) * CENTRE_STEP
    return (f - np.float32(100.0)).astype(np.float32).reshape(k, d)


def perturb(seed: int, centres_: np.ndarray) -> np.ndarray:
    c = np.ascontiguousarray(centres_, dtype=np.float32)
    i = np.arange(c.size, dtype=np.uint64)
    v = _ih8(mix64(seed, np.uint64(2) * i), mix64(seed, np.uint64(2) * i + np.uint64(1)))
    noise = (v.astype(np.float32) * IH8_SCALE) * np.float32(0.1)
    return (c.reshape(-1) + noise).astype(np.float32).reshape(c.shape)


def points(seed: int, first: int, n: int, centres_: np.ndarray) -> np.ndarray:
    c = np.ascontiguousarray(centres_, dtype=np.float32)
    k, d = c.shape
    stride = np.uint64(2 * d + 1)
    i = np.arange(first, first + n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        base = i * stride
        h = mix64(seed, base)
        comp = (((h >> np.uint64(32)) * np.uint64(k)) >> np.uint64(32)).astype(np.int64)
        out = np.empty((n, d), dtype=np.float32)
        for j in range(d):
            v = _ih8(mix64(seed, base + np.uint64(1 + 2 * j)), mix64(seed, base + np.uint64(2 + 2 * j)))
            out[:, j] = c[comp, j] + v.astype(np.float32) * IH8_SCALE
    return out
