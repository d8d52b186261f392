"""Sampled float64 / integer re-computation of single GEMMs of the BinaryNet step on the CPU --
TEST INFRASTRUCTURE ONLY (imported by tests/ only; see oracle/binarynet_oracle.py for the rules).

At BASELINE.json's headline size (batch 4096) the full CPU oracle needs tens of GB for its
autograd tape, so the full-size parity test re-computes SAMPLED entries of every contraction
instead: the test hands these functions the engine's own input maps (gathered on the GPU, so each
GEMM is checked in isolation) and compares the engine's outputs at the same entries.

Semantics restated (reference call sites examples/larq_experiment.py:71-99):
  forward  y[p, co]        = sum_{kh,kw,ci} s[p + (kh-pad, kw-pad), ci] * sign(W)[kh,kw,ci,co]   (zero 'same' padding)
  dgrad    ds[q, ci]       = sum_{kh,kw,co} dy[q - (kh-pad, kw-pad), co] * sign(W)[kh,kw,ci,co]
  wgrad    dW[kh,kw,ci,co] = sum_p s[p + (kh-pad, kw-pad), ci] * dy[p, co]
with sign(0) = +1 (larq ste_sign).  PARITY UNPINNED, like the rest of the oracle.
"""
from __future__ import annotations

import numpy as np


def sign_pm1(w: np.ndarray) -> np.ndarray:
    return np.where(np.asarray(w) >= 0, 1, -1).astype(np.int64)


def conv_forward_at(s_patches: np.ndarray, w_hwio: np.ndarray) -> np.ndarray:
    """s_patches int [n, kh, kw, cin] (zeros outside the image) -> exact integers [n, cout]."""
    # float64 BLAS is exact here (|sum| <= 9 * 512 << 2^53) and much faster than an int64 einsum
    n = s_patches.shape[0]
    y = s_patches.reshape(n, -1).astype(np.float64) @ sign_pm1(w_hwio).reshape(-1, w_hwio.shape[-1]).astype(np.float64)
    return np.rint(y).astype(np.int64)


def conv_dgrad_at(dy_patches: np.ndarray, w_hwio: np.ndarray) -> np.ndarray:
    """dy_patches f64 [n, kh, kw, cout] with patch[kh,kw] = dy[q - (kh-pad, kw-pad)] (zeros outside)
    -> ds f64 [n, cin]."""
    n = dy_patches.shape[0]
    q = sign_pm1(w_hwio).astype(np.float64)                       # [kh, kw, ci, co]
    qt = q.transpose(0, 1, 3, 2).reshape(-1, q.shape[2])          # [(kh, kw, co), ci]
    return dy_patches.astype(np.float64).reshape(n, -1) @ qt


def conv_wgrad_entries(s_shifted: np.ndarray, dy: np.ndarray) -> np.ndarray:
    """s_shifted int8 [taps, M, nci] (tap t: s[p + offset_t, ci], zeros outside), dy f64 [M, nco]
    -> dW f64 [taps, nci, nco]."""
    dy = np.ascontiguousarray(dy, dtype=np.float64)
    return np.stack([s.astype(np.float64).T @ dy for s in s_shifted])


def dense_forward(s_rows: np.ndarray, w_kn: np.ndarray) -> np.ndarray:
    return np.rint(s_rows.astype(np.float64) @ sign_pm1(w_kn).astype(np.float64)).astype(np.int64)


def dense_dgrad(dy_rows: np.ndarray, w_kn: np.ndarray) -> np.ndarray:
    return dy_rows.astype(np.float64) @ sign_pm1(w_kn).astype(np.float64).T


def dense_wgrad_entries(s_cols: np.ndarray, dy_cols: np.ndarray) -> np.ndarray:
    """s_cols int8 [B, nk], dy_cols f64 [B, nn] -> dW f64 [nk, nn]."""
    return s_cols.astype(np.float64).T @ dy_cols.astype(np.float64)


BORDER_TOL = 4e-6   # |(|a| - 1)| below this: the ste mask may legitimately differ between two float evaluations


def bn_backward_columns(ds: np.ndarray, x: np.ndarray, mean: np.ndarray, rstd: np.ndarray, beta: np.ndarray,
                        ste: bool = True):
    """BatchNormalization(scale=False) backward over ALL rows of some channels, float64
    (SURVEY.md Appendix A.1 / A.4).  ds, x: [M, nc]; mean, rstd, beta: [nc] (the forward's values).
    Returns a dict: sum_da, sum_dax [nc]; dy [M, nc]; border (bool [M, nc]: elements whose ste mask
    |a| <= 1 is decided within float32 rounding of the boundary -- another correct implementation
    may decide them the other way); slack_da, slack_dax [nc]: how far the two sums can move if every
    border element flips (the sums are linear in the masked gradient)."""
    ds, x = ds.astype(np.float64), x.astype(np.float64)
    mean, rstd, beta = (np.asarray(v, dtype=np.float64) for v in (mean, rstd, beta))
    xh = (x - mean) * rstd
    a = xh + beta
    if ste:
        da = np.where(np.abs(a) <= 1.0, ds, 0.0)
        border = np.abs(np.abs(a) - 1.0) < BORDER_TOL
    else:
        da = ds
        border = np.zeros(ds.shape, dtype=bool)
    s0, s1 = da.sum(0), (da * xh).sum(0)
    m = float(ds.shape[0])
    dy = rstd * (da - s0 / m - xh * (s1 / m))
    bd = np.where(border, np.abs(ds), 0.0)
    return {"sum_da": s0, "sum_dax": s1, "dy": dy, "border": border, "xhat": xh,
            "slack_da": bd.sum(0), "slack_dax": (bd * np.abs(xh)).sum(0)}


def first_layer_backward_columns(taps: np.ndarray, w_kc: np.ndarray, ds0: np.ndarray, beta: np.ndarray,
                                 eps: float = 1e-3):
    """Whole backward of the un-quantised first conv + BatchNorm + ste_sign for some output
    channels, float64 (examples/larq_experiment.py:62-70 under model.fit).
    taps [M, K] (input window of every output pixel), w_kc = sign(W0)[:, channels] [K, nc],
    ds0 [M, nc] gradient w.r.t. the sign output.  Returns dbeta [nc], dW [K, nc] and the slack of
    each (see bn_backward_columns): dW is linear in the masked gradient, so a border element b at
    pixel p moves dW[k, c] by at most |rstd_c ds_b (x_k(p) - mean x_k - xhat_b mean(x_k xhat_c))|."""
    taps = taps.astype(np.float64)
    y = taps @ w_kc.astype(np.float64)
    mean, var = y.mean(0), y.var(0)
    rstd = 1.0 / np.sqrt(var + eps)
    r = bn_backward_columns(ds0, y, mean, rstd, beta.astype(np.float64))
    dw = taps.T @ r["dy"]
    mx = taps.mean(0)                                   # [K]
    mxxh = taps.T @ r["xhat"] / float(taps.shape[0])    # [K, nc]
    slack_w = np.zeros_like(dw)
    pb, cb = np.nonzero(r["border"])
    for p_, c_ in zip(pb, cb):
        slack_w[:, c_] += np.abs(rstd[c_] * ds0[p_, c_] * (taps[p_] - mx - r["xhat"][p_, c_] * mxxh[:, c_]))
    return {"dbeta": r["sum_da"], "dw": dw, "slack_dbeta": r["slack_da"], "slack_dw": slack_w,
            "mean": mean, "rstd": rstd}


def excess(err: np.ndarray, slack: np.ndarray) -> float:
    """Frobenius norm of the part of |err| that the slack does not explain."""
    return float(np.linalg.norm(np.maximum(np.abs(err) - slack, 0.0)))
