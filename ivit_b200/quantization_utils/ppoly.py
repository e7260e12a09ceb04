"""The `ppoly` registry entries (layer_selection.py:116-131): piecewise-polynomial integer GELU and softmax.

Reference: ppoly_backend.py (segment fit :80-216, integer coefficients :219-290, integer Horner :293-351) and
ppoly_modules.py (PPolyIntGELU :14-136, PPolyIntSoftmax :137-342).  The coefficients are fitted ONCE, on the host, when
a fixed module meets its first batch (the reference does the same: numpy polyfit), and live in the state_dict as the
buffers `fixed_bounds` / `fixed_coeffs` (/ `fixed_scaling_factor_out`).  On the frozen integer path a ppoly function of
an int8 code is a table:

  * GELU + mlp.qact1       -> the [code] line of the GELU table kernel (every row-max row identical),
  * softmax exp term       -> the [row-max code][code] table of the attention kernel (kind 2: no Shiftmax clamp,
                              denominator clamped to >= 1, 2^32 numerator, ppoly_modules.py:315-323),

built here with numpy in the reference's operation order from the stored integer coefficients; the forward itself
runs on the device kernels only.
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32


# ------------------------------------------------------------------------------------------ integer evaluation
def horner_int32(x_int, bounds, coeffs) -> np.ndarray:
    """Integer Horner over the segment of each x (ppoly_backend.py:293-351): segment 0 is x < b[0], segment i is
    b[i-1] <= x < b[i], the last one x >= b[-1]; int64 accumulation, result wrapped to int32 (`.to(torch.int32)`)."""
    x = np.asarray(x_int).astype(np.int64)
    b = np.asarray(bounds).reshape(-1).astype(np.int64)
    c = np.asarray(coeffs).astype(np.int64).reshape(b.size + 1, -1)
    seg = np.searchsorted(b, x, side="right")
    acc = c[seg, 0]
    for d in range(1, c.shape[1]):
        acc = acc * x + c[seg, d]
    return acc.astype(np.int32)


def gelu_code_values(s_in, bounds, coeffs, s_out, scale_bits: int):
    """PPolyIntGELU.forward (ppoly_modules.py:110-136) on the 256 float32 carriers c * s_in of an int8 code c = -128..127.
    Returns the float32 outputs (multiples of s_out)."""
    s_in = f32(np.asarray(s_in).reshape(-1)[0])
    s_out = f32(np.asarray(s_out).reshape(-1)[0])
    x = (np.arange(-128, 128).astype(f32) * s_in).astype(f32)
    x_int = np.floor((x / s_in).astype(f32))
    y_int = horner_int32(x_int, bounds, coeffs)
    y = (y_int.astype(f32) / f32(2 ** scale_bits)).astype(f32)
    return (s_out * np.floor((y / s_out).astype(f32))).astype(f32)


def softmax_term_table(s_attn, bounds, coeffs, exp_bits: int) -> np.ndarray:
    """float32 [256][256]: the exp term of PPolyIntSoftmax (ppoly_modules.py:302-312) for (row-max code, code) of
    qact_attn1's int8 output; entries with code > row max are 0 (never read)."""
    s = f32(np.asarray(s_attn).reshape(-1)[0])
    c = np.arange(-128, 128).astype(f32)
    xi = np.floor((((c * s).astype(f32)) / s).astype(f32)).astype(np.int64)          # floor(x / s).to(int32)
    off = xi[None, :] - xi[:, None] + 127                                             # [row max][code]
    e = np.maximum(horner_int32(off, bounds, coeffs), 0)
    t = np.floor((e.astype(f32) / f32(2 ** (30 - exp_bits + 1))).astype(f32)).astype(f32)
    t[np.triu_indices(256, k=1)] = 0.0
    return t


# ------------------------------------------------------------------------------------------ segment fit (host, once)
def _segment_sse(xs, ys, lo, hi, degree):
    m = (xs >= lo) & (xs <= hi)
    if not m.any():
        return 0.0
    xq, yq = xs[m], ys[m]
    co = np.polyfit(xq, yq, degree)
    return np.sum((yq - np.polyval(co, xq)) ** 2)


def refine_bounds(xs, ys, x_lo, x_hi, segments, degree, rounds: int = 10):
    """Coordinate descent over the interior segment boundaries (ppoly_backend.py:15-77): each boundary in turn tries
    10 positions within 0.3 of its neighbours' span (never closer than a quarter of the uniform width) and keeps the
    one with the lowest total squared error of per-segment least-squares fits.  Only the two segments next to a moved
    boundary are refitted; the total is still accumulated over all segments in order, as the reference sums it."""
    b = np.linspace(x_lo, x_hi, segments + 1, dtype=np.float32)
    min_w = (x_hi - x_lo) / (segments * 4)
    for _ in range(rounds):
        for i in range(1, segments):
            span = b[i + 1] - b[i - 1]
            lo_s = max(b[i - 1] + min_w, b[i] - 0.3 * span)
            hi_s = min(b[i + 1] - min_w, b[i] + 0.3 * span)
            if lo_s >= hi_s:
                continue
            base = [_segment_sse(xs, ys, b[j], b[j + 1], degree) for j in range(segments)]
            best_pos, best_err = b[i], float("inf")
            for pos in np.linspace(lo_s, hi_s, 10):
                trial = b.copy()
                trial[i] = pos
                total = 0
                for j in range(segments):
                    total += (_segment_sse(xs, ys, trial[j], trial[j + 1], degree) if j in (i - 1, i) else base[j])
                if total < best_err:
                    best_err, best_pos = total, pos
            b[i] = best_pos
    return b


def fit_pieces(xs, ys, x_lo, x_hi, segments, degree, alpha=0.0, optim_bounds=True):
    """[((lo, hi), float32 coefficients, highest power first)] per segment (ppoly_backend.py:80-216): least squares on
    the segment's samples in normalised coordinates t = (x - centre) / half-width, expanded back to powers of x."""
    xs, ys = np.asarray(xs).astype(np.float64), np.asarray(ys).astype(np.float64)
    x_lo, x_hi = float(x_lo), float(x_hi)
    b = refine_bounds(xs, ys, x_lo, x_hi, segments, degree) if optim_bounds else \
        np.linspace(x_lo, x_hi, segments + 1, dtype=np.float32)
    pad = alpha * ((x_hi - x_lo) / segments)
    out = []
    for i, (lo, hi) in enumerate(zip(b[:-1], b[1:])):
        f_lo = lo - pad if i > 0 else lo
        f_hi = hi + pad if i < segments - 1 else hi
        m = (xs >= f_lo) & (xs <= f_hi)
        xq, yq = xs[m], ys[m]
        if len(xq) > degree:
            centre, half = (f_lo + f_hi) / 2.0, (f_hi - f_lo) / 2.0
            if abs(half) < 1e-10:
                half, t = 1.0, xq - centre
            else:
                t = (xq - centre) / half
            cn = np.polyfit(t, yq, degree)
            co = np.zeros(degree + 1, dtype=np.float64)
            for j in range(degree + 1):                      # c_j * ((x - centre) / half) ** (degree - j), binomially
                p = degree - j
                for k in range(p + 1):
                    binom = math.factorial(p) / (math.factorial(k) * math.factorial(p - k))
                    co[degree - k] += cn[j] * binom * ((-centre / half) ** (p - k)) / (half ** k)
            co = co.astype(np.float32)
        else:
            co = np.zeros(degree + 1, dtype=np.float32)
            if len(yq) > 0:
                co[-1] = np.mean(yq)
        out.append(((lo, hi), co))
    return out


def integer_coefficients(pieces, scaling_factor, scale_bits: int, device):
    """(bounds int32 [S-1, ...], coeffs int32 [S, D+1, ...]) in torch (ppoly_backend.py:219-290): boundary
    floor(lo / s); coefficient of x^p: floor(c * s^p * 2^N).  float32 tensor arithmetic, as the reference."""
    import torch
    bounds, coeffs = [], []
    for idx, ((lo, _hi), co) in enumerate(pieces):
        if idx > 0:
            bounds.append(torch.floor(torch.tensor(lo, device=device) / scaling_factor))
        deg = len(co) - 1
        coeffs.append(torch.stack([torch.floor((c * (scaling_factor ** (deg - i)) * (2 ** scale_bits)).clone().detach())
                                   for i, c in enumerate(co)]))
    b = torch.stack(bounds).to(torch.int32) if bounds else torch.tensor([], dtype=torch.int32, device=device)
    return b, torch.stack(coeffs).to(torch.int32)
