"""Synthetic inputs of the named shapes.  Every value is a pure function of (seed, column, GLOBAL row
index) through a counter hash, so any row sharding over ranks yields the same global dataset.

Shapes: ``airbnb`` (C1/C2: the ML 01 / ML 06 feature table — 7 StringIndexer-ed categoricals + 26 numerics, label price,
`ML 01 - Data Cleansing.py:49-74,220-234`, `ML 06 - Decision Trees.py:44,60-62`), ``c3`` (10M x 64 continuous),
``c4`` (100M x 128 mixed).  The generator runs on whatever device it is given (torch ops only)."""
from __future__ import annotations

import math

import numpy as np
import torch

_M64 = (1 << 64) - 1


def _s64(x: int) -> int:
    x &= _M64
    return x - (1 << 64) if x >= (1 << 63) else x


_K1, _K2, _K3 = _s64(0x9E3779B97F4A7C15), _s64(0xBF58476D1CE4E5B9), _s64(0x94D049BB133111EB)


def _lsr(z: torch.Tensor, k: int) -> torch.Tensor:
    return (z >> k) & ((1 << (64 - k)) - 1)


def _mix(z: torch.Tensor) -> torch.Tensor:
    z = (z ^ _lsr(z, 30)) * _K2
    z = (z ^ _lsr(z, 27)) * _K3
    return z ^ _lsr(z, 31)


def uniform(seed: int, stream: int, idx: torch.Tensor) -> torch.Tensor:
    """u in (0,1), f64, for int64 global indices idx."""
    base = _s64(seed * 0x632BE59BD9B4E019 + (stream + 1) * 0xD6E8FEB86659FD93)
    z = _mix(idx * _K1 + base)
    return (_lsr(z, 11).to(torch.float64) + 0.5) * (2.0 ** -53)


def normal(seed: int, stream: int, idx: torch.Tensor) -> torch.Tensor:
    u1 = uniform(seed, 2 * stream + 1000, idx)
    u2 = uniform(seed, 2 * stream + 1001, idx)
    return torch.sqrt(-2.0 * torch.log(u1)) * torch.cos(2.0 * math.pi * u2)


def _zipf_codes(u: torch.Tensor, arity: int, s: float = 1.2) -> torch.Tensor:
    w = np.arange(1, arity + 1, dtype=np.float64) ** (-s)
    cdf = torch.as_tensor(np.cumsum(w / w.sum())[:-1], dtype=torch.float64, device=u.device)
    return torch.bucketize(u, cdf).to(torch.float64)


def _poisson_codes(u: torch.Tensor, lam: float, kmax: int = 24) -> torch.Tensor:
    p = [math.exp(-lam) * lam ** k / math.factorial(k) for k in range(kmax)]
    cdf = torch.as_tensor(np.cumsum(p)[:-1], dtype=torch.float64, device=u.device)
    return torch.bucketize(u, cdf).to(torch.float64)


AIRBNB_ARITIES = [2, 2, 3, 5, 5, 26, 36]   # 7 string columns of ML 01:49-74 after StringIndexer (largest in (32, 40], ML 06:110-118)


def airbnb(n_rows: int, row0: int = 0, seed: int = 42, device="cpu"):
    """Returns X f64 [n, 33], y f64 [n], arity int32 [33].  Column order = StringIndexer outputs first, then numerics
    (`ML 06 - Decision Trees.py:60-62`)."""
    idx = torch.arange(row0, row0 + n_rows, dtype=torch.int64, device=device)
    cols = []
    for j, a in enumerate(AIRBNB_ARITIES):
        cols.append(_zipf_codes(uniform(seed, j, idx), a))
    lat = 37.76 + 0.02 * normal(seed, 10, idx)
    lon = -122.43 + 0.03 * normal(seed, 11, idx)
    cols += [lat, lon]
    counts = []
    for j, (mu, sg) in enumerate([(1.0, 0.5), (0.4, 0.5), (0.2, 0.5), (0.1, 0.4), (0.8, 1.0), (2.0, 1.6), (0.3, 1.2)]):
        v = torch.floor(torch.exp(mu + sg * normal(seed, 20 + j, idx)))
        if j == 3:
            v = torch.floor(2 * torch.exp(mu + sg * normal(seed, 20 + j, idx))) / 2.0   # bathrooms in {k/2}
        counts.append(v)
    cols += counts
    for j in range(7):
        hi = 100.0 if j == 0 else 10.0
        lo = 20.0 if j == 0 else 2.0
        v = hi - torch.floor(torch.abs(normal(seed, 40 + j, idx)) * (hi - lo) / 4.0)
        cols.append(torch.clamp(v, lo, hi))
    for j in range(10):
        cols.append((uniform(seed, 60 + j, idx) < 0.2).to(torch.float64))
    X = torch.stack(cols, 1).contiguous()
    price = torch.exp(5.0 + 0.7 * normal(seed, 90, idx))
    price = price + 40.0 * X[:, 9] + 25.0 * X[:, 11] + 30.0 * (X[:, 6] == 0).to(torch.float64) * (X[:, 5] < 5).to(torch.float64) \
        + 60.0 * (X[:, 3] == 1).to(torch.float64) - 15.0 * (X[:, 7] < 37.75).to(torch.float64) + 2.0 * (X[:, 16] - 90.0).clamp(min=-40)
    y = torch.clamp(torch.round(price), 10.0, 10000.0)
    arity = np.array(AIRBNB_ARITIES + [0] * 26, np.int32)
    return X, y, arity


def c3(n_rows: int, row0: int = 0, seed: int = 42, device="cuda", F: int = 64):
    """C3: 10M x 64 continuous — 32 N(0,1) rounded to fp32, 16 lognormal, 16 integer Poisson(3) (exercises the zero filter)."""
    idx = torch.arange(row0, row0 + n_rows, dtype=torch.int64, device=device)
    X = torch.empty((n_rows, F), dtype=torch.float64, device=device)
    n_norm, n_logn = F // 2, F // 4
    for f in range(F):
        if f < n_norm:
            X[:, f] = normal(seed, f, idx).to(torch.float32).to(torch.float64)
        elif f < n_norm + n_logn:
            X[:, f] = torch.exp(0.5 * normal(seed, f, idx)).to(torch.float32).to(torch.float64)
        else:
            X[:, f] = _poisson_codes(uniform(seed, f, idx), 3.0)
    y = normal(seed, 5000, idx)
    for k in range(8):   # fixed step thresholds (no data-dependent constants: shards must agree)
        f = (k * 7) % F
        thr = (0.1 * k - 0.3) if f < n_norm else ((1.0 + 0.1 * k) if f < n_norm + n_logn else 2.5)
        y = y + (1.0 + 0.5 * k) * (X[:, f] > thr).to(torch.float64)
    return X, y, np.zeros(F, np.int32)


C4_ARITIES = [2, 4, 8, 16, 36, 64, 128, 256]


def c4(n_rows: int, row0: int = 0, seed: int = 42, device="cuda", F: int = 128):
    """C4: 100M x 128 — 96 continuous as C3 + 32 categorical with arities cycling {2,...,256} (Zipf codes)."""
    n_cat = F // 4
    n_cont = F - n_cat
    Xc, y, _ = c3(n_rows, row0, seed, device, F=n_cont)
    idx = torch.arange(row0, row0 + n_rows, dtype=torch.int64, device=device)
    X = torch.empty((n_rows, F), dtype=torch.float64, device=device)
    X[:, :n_cont] = Xc
    del Xc
    arity = np.zeros(F, np.int32)
    for j in range(n_cat):
        a = C4_ARITIES[j % len(C4_ARITIES)]
        arity[n_cont + j] = a
        X[:, n_cont + j] = _zipf_codes(uniform(seed, 7000 + j, idx), a)
    for k in range(4):
        col = n_cont + 3 + 8 * k
        y = y + 0.8 * (X[:, col] % 3 == 1).to(torch.float64)
    return X, y, arity
