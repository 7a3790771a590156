"""Shared helpers for the test-suite: fixture loaders and seeded generators.

The generators follow the *distributions* of photon-test-utils' SparkTestUtils
(photon-test-utils/src/main/scala/com/linkedin/photon/ml/test/SparkTestUtils.scala:314-316,774-822:
"benign" = uniform +-1 features, "outlier" = 90% N(0,1e-3) / 10% +-1) with numpy's RNG; the exact
commons-math3 Well19937a draws are not reproducible here and no reference test depends on them.
"""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def kats():
    with open(os.path.join(GOLDEN, "reference_kats.json")) as f:
        return json.load(f)


def read_libsvm(path, d=None, add_intercept=False, zero_based=False, max_rows=None):
    """LibSVM text -> CSR with labels mapped >0 -> 1 else 0
    (photon-client/.../io/deprecated/LibSVMInputDataFormat.scala:51-82)."""
    row_ptr, col, val, label = [0], [], [], []
    with open(path) as f:
        for i, line in enumerate(f):
            if max_rows is not None and i >= max_rows:
                break
            ts = line.split()
            if not ts:
                continue
            label.append(1.0 if float(ts[0]) > 0 else 0.0)
            for t in ts[1:]:
                j, v = t.split(":")
                col.append(int(j) - (0 if zero_based else 1))
                val.append(float(v))
            row_ptr.append(len(col))
    col = np.asarray(col, dtype=np.int32)
    val = np.asarray(val, dtype=np.float64)
    row_ptr = np.asarray(row_ptr, dtype=np.int64)
    label = np.asarray(label, dtype=np.float64)
    if d is None:
        d = int(col.max()) + 1
    if add_intercept:
        row_ptr, col, val = append_intercept(row_ptr, col, val, d)
        d += 1
    return len(label), d, row_ptr, col, val, label


def append_intercept(row_ptr, col, val, d):
    n = len(row_ptr) - 1
    lens = np.diff(row_ptr)
    new_ptr = np.concatenate([[0], np.cumsum(lens + 1)]).astype(np.int64)
    new_col = np.empty(new_ptr[-1], dtype=np.int32)
    new_val = np.empty(new_ptr[-1], dtype=np.float64)
    for i in range(n):
        a, b = row_ptr[i], row_ptr[i + 1]
        na = new_ptr[i]
        new_col[na:na + (b - a)] = col[a:b]
        new_val[na:na + (b - a)] = val[a:b]
        new_col[new_ptr[i + 1] - 1] = d
        new_val[new_ptr[i + 1] - 1] = 1.0
    return new_ptr, new_col, new_val


def dense_to_csr(X, keep_zeros=False):
    X = np.asarray(X, dtype=np.float64)
    n, d = X.shape
    row_ptr, col, val = [0], [], []
    for i in range(n):
        for j in range(d):
            if keep_zeros or X[i, j] != 0.0:
                col.append(j)
                val.append(X[i, j])
        row_ptr.append(len(col))
    return (np.asarray(row_ptr, dtype=np.int64), np.asarray(col, dtype=np.int32), np.asarray(val, dtype=np.float64))


def csr_to_dense(n, d, row_ptr, col, val):
    X = np.zeros((n, d))
    for i in range(n):
        for k in range(row_ptr[i], row_ptr[i + 1]):
            X[i, col[k]] += val[k]
    return X


def random_csr(rng, n, d, nnz_per_row, fp32_exact=True, skew=None, distinct=True, empty_rows=0.0):
    """Seeded ragged CSR.  skew=None: uniform columns; skew=a: Zipf(a)-ish column popularity."""
    if skew is None:
        p = None
    else:
        p = 1.0 / np.arange(1, d + 1) ** skew
        p /= p.sum()
    lens = rng.poisson(nnz_per_row, size=n).astype(np.int64)
    lens = np.minimum(lens, d)
    if empty_rows > 0:
        lens[rng.random(n) < empty_rows] = 0
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = np.empty(row_ptr[-1], dtype=np.int32)
    for i in range(n):
        k = lens[i]
        if k == 0:
            continue
        if distinct:
            c = rng.choice(d, size=k, replace=False, p=p)
        else:
            c = rng.choice(d, size=k, replace=True, p=p)
        col[row_ptr[i]:row_ptr[i + 1]] = np.sort(c)
    val = rng.uniform(-1.0, 1.0, size=row_ptr[-1])
    if fp32_exact:
        val = val.astype(np.float32).astype(np.float64)
    return row_ptr, col, val


def fast_random_csr(rng, n, d, k, skew=None):
    """Vectorised generator for big cases: exactly k draws per row (duplicates allowed, kept)."""
    if skew is None:
        col = rng.integers(0, d, size=(n, k), dtype=np.int64)
    else:
        p = 1.0 / np.arange(1, d + 1) ** skew
        cdf = np.cumsum(p / p.sum())
        col = np.searchsorted(cdf, rng.random((n, k))).clip(0, d - 1)
    col = np.sort(col, axis=1).astype(np.int32).ravel()
    row_ptr = (np.arange(n + 1, dtype=np.int64) * k)
    val = (rng.integers(1, 5, size=n * k) / 4.0).astype(np.float64)
    return row_ptr, col, val


def labels_for(rng, loss, n, margins=None):
    if loss == 0 or loss == 3:  # logistic / smoothed hinge: {0,1}
        return (rng.random(n) < 0.5).astype(np.float64)
    if loss == 1:  # poisson counts
        lam = np.exp(np.clip(margins if margins is not None else rng.normal(size=n), -3, 3))
        return rng.poisson(lam).astype(np.float64)
    return rng.normal(size=n)


def column_summary(n, d, row_ptr, col, val):
    """mean / unbiased variance / max / min per column, zeros included -- what MLlib colStats
    returns and lib/stat/BasicStatisticalSummary.scala:60-117 wraps."""
    X = csr_to_dense(n, d, row_ptr, col, val)
    mean = X.mean(axis=0)
    var = X.var(axis=0, ddof=1) if n > 1 else np.zeros(d)
    return mean, var, X.max(axis=0), X.min(axis=0)


def normalization_context(kind, n, d, row_ptr, col, val, intercept):
    """lib/normalization/NormalizationContext.scala:124-162 -> (factors, shifts, intercept)."""
    mean, var, mx, mn = column_summary(n, d, row_ptr, col, val)
    if kind == "NONE":
        return None, None, intercept
    if kind == "SCALE_WITH_MAX_MAGNITUDE":
        mag = np.maximum(np.abs(mx), np.abs(mn))
        return np.where(mag == 0, 1.0, 1.0 / np.where(mag == 0, 1.0, mag)), None, intercept
    std = np.sqrt(var)
    factors = np.where(std == 0, 1.0, 1.0 / np.where(std == 0, 1.0, std))
    if kind == "SCALE_WITH_STANDARD_DEVIATION":
        return factors, None, intercept
    if kind == "STANDARDIZATION":
        shifts = mean.copy()
        if intercept is not None:
            shifts[intercept] = 0.0
            factors[intercept] = 1.0
        return factors, shifts, intercept
    raise ValueError(kind)


def model_to_transformed_space(coef, factors, shifts, intercept):
    """NormalizationContext.modelToTransformedSpace, :97-111."""
    out = np.array(coef, dtype=np.float64)
    if shifts is not None:
        out[intercept] += out.dot(shifts)
    if factors is not None:
        out = out / factors
    return out


def model_to_original_space(coef, factors, shifts, intercept):
    """NormalizationContext.modelToOriginalSpace, :64-80."""
    out = np.array(coef, dtype=np.float64)
    if factors is not None:
        out = out * factors
    if shifts is not None:
        out[intercept] -= out.dot(shifts)
    return out


def rel_err(a, b):
    """norm-wise relative error: max|a-b| / max|b| (abs when b == 0)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.max(np.abs(b)) if b.size else 0.0
    num = np.max(np.abs(a - b)) if b.size else 0.0
    return num / den if den > 0 else num
