"""CPU definitions of the mock-model forwards (TEST INFRASTRUCTURE).

* ``echo``   -- the reference's actual model: ``output = inputs`` (fake_model.py:52). Bit-exact.
* ``linear`` -- BASELINE.json configs[1] "mock linear model".  NOT in the reference
  ("parity unpinned"); defined here as ``Y = X @ W.T + b`` with
  fp32 storage, evaluated in float64.
* ``mlp``    -- BASELINE.json configs[4] "4096-hidden bf16 MLP".  NOT in the reference;
  defined as ``Y = bf16(relu(bf16(X) @ bf16(W1).T + b1)) @ bf16(W2).T + b2`` with fp32
  accumulation emulated in float64 on bf16-rounded operands, hidden rounded to bf16.

Tolerances the GPU kernels are held to (written again in the tests that use them):
  linear, tf32 tensor-core inputs :  |d| <= 2**-10 * sum_k |x_k w_k| + 1e-6
  linear, fp32 CUDA-core path     :  |d| <= 1e-5 * sum_k |x_k w_k| + 1e-6
  mlp, bf16 out                   :  |d| <= 2**-7 * |y| + 1e-2 * max_j |y_j|  per row
"""
from __future__ import annotations

import numpy as np


def echo(rows):
    return rows


def linear_f64(x: np.ndarray, w: np.ndarray, b: np.ndarray) -> np.ndarray:
    """x [B,K] fp32, w [N,K] fp32, b [N] fp32 -> float64 [B,N]."""
    return x.astype(np.float64) @ w.astype(np.float64).T + b.astype(np.float64)


def linear_abs_bound(x: np.ndarray, w: np.ndarray) -> np.ndarray:
    """sum_k |x_k w_k| per output element, the scale of the rounding-error bound."""
    return np.abs(x).astype(np.float64) @ np.abs(w).astype(np.float64).T


def to_bf16(a: np.ndarray) -> np.ndarray:
    """Round fp32 -> bf16 (round-to-nearest-even), returned as fp32 holding bf16 values."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    u = a.view(np.uint32).astype(np.uint64)
    rounding = ((u >> 16) & 1) + 0x7FFF
    u = ((u + rounding) >> 16) << 16
    out = (u & 0xFFFFFFFF).astype(np.uint32).view(np.float32)
    return np.where(np.isnan(a), a, out).astype(np.float32)


def bf16_bits(a: np.ndarray) -> np.ndarray:
    """fp32 array already holding bf16-representable values -> uint16 bit patterns."""
    return (np.ascontiguousarray(a, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def bf16_from_bits(bits: np.ndarray) -> np.ndarray:
    return (bits.astype(np.uint32) << 16).view(np.float32)


def mlp_bf16(x, w1, b1, w2, b2) -> np.ndarray:
    """Operands fp32 arrays holding bf16 values; biases fp32. Returns float64 [B, D_out]
    (the un-rounded value the bf16 output is compared with)."""
    h = x.astype(np.float64) @ w1.astype(np.float64).T + b1.astype(np.float64)
    h = to_bf16(np.maximum(h, 0.0).astype(np.float32)).astype(np.float64)
    return h @ w2.astype(np.float64).T + b2.astype(np.float64)
