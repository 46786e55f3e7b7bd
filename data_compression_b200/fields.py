"""Synthetic float32 fields for the parity tests and the bench.

All three are >= 0 after the per-chunk min-subtract, and keep CT=7 inside the reference decoder's
defined domain (type=2, mask exponent e=2) like the reference's own float_rand_* datasets
(impl/dataset/gen_nums_float.py:28).  Host (numpy) and device (torch) generators are bit-identical:
`noisy` is integer arithmetic, `smooth` is one host-generated period tiled on the device,
`constant` is a literal.
"""
import numpy as np

SMOOTH_PERIOD = 65536
NOISY_SEED = 20240607
CONSTANT_VALUE = np.float32(0.123456789)
KINDS = ("smooth", "noisy", "constant")


def _smooth_period():
    i = np.arange(SMOOTH_PERIOD, dtype=np.float64)
    return (5.0 + 4.5 * np.sin(2.0 * np.pi * i / SMOOTH_PERIOD)).astype(np.float32)


def _fmix32_np(h):
    h = h.astype(np.uint64)
    m = np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x85EBCA6B)) & m
    h ^= h >> np.uint64(13)
    h = (h * np.uint64(0xC2B2AE35)) & m
    h ^= h >> np.uint64(16)
    return h


def make_field(kind, n, start=0):
    """numpy float32 field of n elements whose first element has global index `start`."""
    if kind == "smooth":
        p = _smooth_period()
        idx = (np.arange(n, dtype=np.int64) + start) & (SMOOTH_PERIOD - 1)
        return p[idx]
    if kind == "noisy":
        i = (np.arange(n, dtype=np.uint64) + np.uint64(start)) & np.uint64(0xFFFFFFFF)
        h = _fmix32_np((i * np.uint64(0x9E3779B9) + np.uint64(NOISY_SEED)) & np.uint64(0xFFFFFFFF))
        return ((h >> np.uint64(8)).astype(np.float32) * np.float32(10.0 / 16777216.0)).astype(np.float32)
    if kind == "constant":
        return np.full(n, CONSTANT_VALUE, dtype=np.float32)
    raise ValueError(kind)


def make_field_torch(kind, n, device, start=0, out=None):
    """Same field as make_field, generated on `device` (slice-wise so 2^30 elements fit comfortably)."""
    import torch
    x = out if out is not None else torch.empty(n, dtype=torch.float32, device=device)
    if kind == "constant":
        x.fill_(float(CONSTANT_VALUE))
        return x
    if kind == "smooth":
        p = torch.from_numpy(_smooth_period()).to(device)
    step = 1 << 26
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        i = torch.arange(lo + start, hi + start, dtype=torch.int64, device=device)
        if kind == "smooth":
            x[lo:hi] = p[i & (SMOOTH_PERIOD - 1)]
        elif kind == "noisy":
            m = 0xFFFFFFFF
            h = ((i & m) * 0x9E3779B9 + NOISY_SEED) & m
            h = h ^ (h >> 16)
            h = (h * 0x85EBCA6B) & m
            h = h ^ (h >> 13)
            h = (h * 0xC2B2AE35) & m
            h = h ^ (h >> 16)
            x[lo:hi] = (h >> 8).to(torch.float32) * float(np.float32(10.0 / 16777216.0))
        else:
            raise ValueError(kind)
    return x
