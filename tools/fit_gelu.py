#!/usr/bin/env python
"""Fits the degree-5 polynomial q(a) ~ -log2(erfc(a / sqrt 2) / 2) used by gelu_erf_fast (csrc/gemm_tc.cuh) and reports the
fp32 GELU error of the fitted form over [-30, 30]."""
import numpy as np
from scipy.special import erfc

a = np.linspace(0, 7, 200001)
h = 0.5 * erfc(a / np.sqrt(2))
q = -np.log2(h)
w = a * h + 1e-4                      # weight = sensitivity of the GELU value to an error in q
c = np.polynomial.polynomial.polyfit(a, q, 5, w=w)
for _ in range(60):                   # iterative re-weighting towards the minimax fit of the GELU error
    err = np.abs(a * (np.exp2(-np.polynomial.polynomial.polyval(a, c)) - h))
    w = w * (1 + 2 * err / err.max())
    c = np.polynomial.polynomial.polyfit(a, q, 5, w=w)
c32 = np.float32(c)
print("coefficients (c0..c5):", [repr(float(v)) for v in c32])
x = np.float32(np.linspace(-30, 30, 2000001))
aa = np.abs(x)
qq = c32[5]
for k in (4, 3, 2, 1, 0):
    qq = np.float32(qq * aa + c32[k])
t = x * np.exp2(-qq).astype(np.float32)
g = np.where(x > 0, x - t, t)
ref = x.astype(np.float64) * 0.5 * erfc(-x.astype(np.float64) / np.sqrt(2))
print("max |gelu error| fp32:", np.abs(g - ref).max())
