"""Synthetic MNIST-shaped workload and the file formats either side of the path.

The reference ships no data (its hard-coded mnist3_*.csv files are absent, code/main3.cpp:311-314),
so every workload here is generated: uint8 [N x 784] images built from 10 classes of stroke
prototypes with shifts, cross-class blending and label noise, labels 0..9, binary target
"digit == 1 -> +1 else -1" (code/main3.cpp:49-52).  The same bytes can be written as
idx3/idx1-ubyte (product input), as the reference's CSV (header line, last column = label,
code/main3.cpp:13-54) or kept in memory.
"""
from __future__ import annotations

import struct

import numpy as np

N_FEATURES = 784
SIDE = 28


def _prototypes(rng: np.random.Generator, n_classes: int, n_sub: int, sub_jitter: float,
                stroke: float) -> np.ndarray:
    yy, xx = np.mgrid[0:SIDE, 0:SIDE].astype(np.float64)
    protos = np.zeros((n_classes, n_sub, SIDE, SIDE))
    for c in range(n_classes):
        base = rng.uniform(6, 22, size=(4, 2))  # class skeleton: 4 anchor points
        for s in range(n_sub):
            pts = base + rng.normal(0, sub_jitter, size=base.shape)
            img = np.zeros((SIDE, SIDE))
            order = [(0, 1), (1, 2), (2, 3)] if (c + s) % 2 else [(0, 2), (2, 1), (1, 3)]
            for a, b in order:
                p, q = pts[a], pts[b]
                v = q - p
                L2 = float(v @ v) + 1e-9
                t = np.clip(((xx - p[0]) * v[0] + (yy - p[1]) * v[1]) / L2, 0, 1)
                d2 = (xx - (p[0] + t * v[0])) ** 2 + (yy - (p[1] + t * v[1])) ** 2
                img = np.maximum(img, np.exp(-d2 / (2 * stroke ** 2)))
            protos[c, s] = img
    return protos


def make_dataset(n: int, seed: int = 587, n_sub: int = 4, blend: float = 0.8,
                 label_noise: float = 0.02, max_shift: int = 2, sub_jitter: float = 1.2,
                 stroke: float = 1.8, split: int = 0):
    """Return (X uint8 [n,784], digits uint8 [n]).  Deterministic in (n, seed, knobs): row i does
    not depend on n, so the first k rows of a bigger set are the k-row set.  split=0 is the training
    stream, split=1 the test stream (same class prototypes, independent samples)."""
    proto_rng = np.random.default_rng([seed, 0])
    protos = _prototypes(proto_rng, 10, n_sub, sub_jitter, stroke)
    X = np.zeros((n, N_FEATURES), dtype=np.uint8)
    digits = np.zeros(n, dtype=np.uint8)
    CH = 4096
    for start in range(0, n, CH):
        m = min(CH, n - start)
        rng = np.random.default_rng([seed, 1 + split, start // CH])
        c = rng.integers(0, 10, size=CH)
        s = rng.integers(0, n_sub, size=CH)
        c2 = (c + rng.integers(1, 10, size=CH)) % 10
        s2 = rng.integers(0, n_sub, size=CH)
        w = rng.uniform(0, blend, size=CH)
        gain = rng.uniform(0.55, 1.0, size=CH)
        dx = rng.integers(-2, 3, size=CH).clip(-max_shift, max_shift)
        dy = rng.integers(-2, 3, size=CH).clip(-max_shift, max_shift)
        noise = rng.normal(0, 0.10, size=(CH, SIDE, SIDE))
        flip = rng.uniform(size=CH) < label_noise
        newlab = rng.integers(0, 10, size=CH)
        img = protos[c, s] + w[:, None, None] * protos[c2, s2]
        out = np.zeros_like(img)
        for ddx in range(-2, 3):
            for ddy in range(-2, 3):
                sel = np.nonzero((dx == ddx) & (dy == ddy))[0]
                if sel.size:
                    out[sel] = np.roll(img[sel], shift=(ddy, ddx), axis=(1, 2))
        out = out * gain[:, None, None] + noise * (out > 0.05)
        out = np.clip(out, 0, 1)
        q = np.rint(out * 255.0)
        q[q < 48] = 0
        X[start:start + m] = q.reshape(CH, -1)[:m].astype(np.uint8)
        lab = np.where(flip, newlab, c)
        digits[start:start + m] = lab[:m].astype(np.uint8)
    return X, digits


def binary_labels(digits: np.ndarray, positive_digit: int = 1) -> np.ndarray:
    """code/main3.cpp:49-52: label != 1 -> -1."""
    return np.where(digits == positive_digit, 1, -1).astype(np.int32)


# ---- file formats -------------------------------------------------------------------------

def write_idx(images_path: str, labels_path: str, X: np.ndarray, digits: np.ndarray) -> None:
    """MNIST idx3-ubyte / idx1-ubyte (big-endian magic 0x00000803 / 0x00000801)."""
    n = X.shape[0]
    with open(images_path, "wb") as f:
        f.write(struct.pack(">IIII", 0x00000803, n, SIDE, SIDE))
        f.write(np.ascontiguousarray(X, dtype=np.uint8).tobytes())
    with open(labels_path, "wb") as f:
        f.write(struct.pack(">II", 0x00000801, n))
        f.write(np.ascontiguousarray(digits, dtype=np.uint8).tobytes())


def read_idx(images_path: str, labels_path: str):
    with open(images_path, "rb") as f:
        magic, n, r, c = struct.unpack(">IIII", f.read(16))
        if magic != 0x00000803:
            raise ValueError(f"{images_path}: bad idx3 magic {magic:#x}")
        X = np.frombuffer(f.read(n * r * c), dtype=np.uint8).reshape(n, r * c)
    with open(labels_path, "rb") as f:
        magic, n2 = struct.unpack(">II", f.read(8))
        if magic != 0x00000801 or n2 != n:
            raise ValueError(f"{labels_path}: bad idx1 header")
        digits = np.frombuffer(f.read(n), dtype=np.uint8)
    return X.copy(), digits.copy()


def write_csv(path: str, X: np.ndarray, digits: np.ndarray) -> None:
    """The reference's input format (code/main3.cpp:13-54): one header line, then
    feature,...,feature,label with integer pixels (so std::stod is exact)."""
    d = X.shape[1]
    with open(path, "w") as f:
        f.write(",".join([f"p{j}" for j in range(d)] + ["label"]) + "\n")
        for row, lab in zip(X, digits):
            f.write(",".join(map(str, row.tolist())) + f",{int(lab)}\n")
