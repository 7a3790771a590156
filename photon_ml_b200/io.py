"""Packer front-ends: the formats either side of the hot path.

  read_libsvm              client/io/deprecated/LibSVMInputDataFormat.scala:51-82 -> CSR + labels
  BasicStatisticalSummary  lib/stat/BasicStatisticalSummary.scala:60-117 (MLlib colStats: zeros included,
                           unbiased variance) computed sparsely from the CSR arrays
  normalization_context    lib/normalization/NormalizationContext.scala:124-162 (the factory)
  IndexMapProjection       api/projector/IndexMapProjectorRDD.scala:166-206 + IndexMapProjector.scala:45-91: per-entity
                           projection of the shard's feature space onto the features the entity's rows use
(paths under photon-client / photon-lib src/main/scala/com/linkedin/photon/ml).  Host-side, numpy; they run once
per data set, before packing.
"""
import numpy as np

from .function import NormalizationContext


def read_libsvm(path, feature_dimension=None, add_intercept=True, zero_based=False, max_rows=None):
    """LibSVM text -> (n, d, row_ptr, col, val, label).  Labels > 0 become 1, the rest 0; indices are one-based
    unless zero_based; with add_intercept a column of ones is appended as the last feature."""
    labels, cols, vals, lens = [], [], [], []
    with open(path) as f:
        for i, line in enumerate(f):
            if max_rows is not None and i >= max_rows:
                break
            ts = line.split()
            if not ts:
                continue
            labels.append(1.0 if float(ts[0]) > 0 else 0.0)
            pairs = [t.split(":") for t in ts[1:]]
            cols.extend(int(p[0]) for p in pairs)
            vals.extend(float(p[1]) for p in pairs)
            lens.append(len(pairs))
    col = np.asarray(cols, dtype=np.int64) - (0 if zero_based else 1)
    val = np.asarray(vals, dtype=np.float64)
    lens = np.asarray(lens, dtype=np.int64)
    n = len(labels)
    d = int(feature_dimension) if feature_dimension is not None else (int(col.max()) + 1 if col.size else 0)
    if col.size and (col.min() < 0 or col.max() >= d):
        raise ValueError("feature index outside [0, %d)" % d)
    row_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    if add_intercept:
        # one extra entry (d, 1.0) at the end of every row
        new_ptr = row_ptr + np.arange(n + 1)
        new_col = np.empty(col.size + n, dtype=np.int64)
        new_val = np.empty(val.size + n, dtype=np.float64)
        keep = np.ones(new_col.size, dtype=bool)
        keep[new_ptr[1:] - 1] = False
        new_col[keep], new_val[keep] = col, val
        new_col[~keep], new_val[~keep] = d, 1.0
        row_ptr, col, val, d = new_ptr, new_col, new_val, d + 1
    return n, d, row_ptr, col.astype(np.int32), val, np.asarray(labels, dtype=np.float64)


class BasicStatisticalSummary:
    """mean, unbiased variance, max, min, number of non-zeros per column -- zeros count as observations."""

    def __init__(self, n, d, row_ptr, col, val):
        col = np.asarray(col, dtype=np.int64)
        val = np.asarray(val, dtype=np.float64)
        self.count = int(n)
        s1 = np.bincount(col, weights=val, minlength=d)
        s2 = np.bincount(col, weights=val * val, minlength=d)
        self.num_nonzeros = np.bincount(col[val != 0], minlength=d).astype(np.int64)
        stored = np.bincount(col, minlength=d)
        self.mean = s1 / n if n else np.zeros(d)
        # sum (x - mean)^2 over all n observations, the implicit zeros included
        ss = s2 - n * self.mean * self.mean
        self.variance = np.maximum(ss, 0.0) / (n - 1) if n > 1 else np.zeros(d)
        mx = np.full(d, -np.inf)
        mn = np.full(d, np.inf)
        np.maximum.at(mx, col, val)
        np.minimum.at(mn, col, val)
        has_implicit_zero = stored < n
        self.max = np.where(has_implicit_zero, np.maximum(mx, 0.0), mx)
        self.min = np.where(has_implicit_zero, np.minimum(mn, 0.0), mn)
        self.max = np.where(np.isinf(self.max), 0.0, self.max)
        self.min = np.where(np.isinf(self.min), 0.0, self.min)


NONE, SCALE_WITH_MAX_MAGNITUDE, SCALE_WITH_STANDARD_DEVIATION, STANDARDIZATION = (
    "NONE", "SCALE_WITH_MAX_MAGNITUDE", "SCALE_WITH_STANDARD_DEVIATION", "STANDARDIZATION")


def normalization_context(normalization_type, summary, intercept_id):
    """NormalizationContext.apply(normalizationType, summary, interceptId)."""
    if normalization_type == NONE:
        return NormalizationContext(None, None, intercept_id)
    if normalization_type == SCALE_WITH_MAX_MAGNITUDE:
        magnitude = np.maximum(np.abs(summary.max), np.abs(summary.min))
        factors = np.where(magnitude == 0, 1.0, 1.0 / np.where(magnitude == 0, 1.0, magnitude))
        return NormalizationContext(factors, None, intercept_id)
    std = np.sqrt(summary.variance)
    factors = np.where(std == 0, 1.0, 1.0 / np.where(std == 0, 1.0, std))
    if normalization_type == SCALE_WITH_STANDARD_DEVIATION:
        return NormalizationContext(factors, None, intercept_id)
    if normalization_type == STANDARDIZATION:
        shifts = summary.mean.copy()
        if intercept_id is not None:     # do not transform the intercept
            shifts[intercept_id] = 0.0
            factors[intercept_id] = 1.0
        return NormalizationContext(factors, shifts, intercept_id)
    raise ValueError("NormalizationType %s not recognized." % normalization_type)


class IndexMapProjection:
    """Per-entity index-map projectors for a random-effect data set whose rows are grouped by entity: entity e's
    projected space is the set of original feature indices active in its rows (IndexMapProjectorRDD.
    buildIndexMapProjector: activeIndices ... zipWithIndex), here numbered in ascending original index so the result
    is deterministic (the reference numbers them in Set iteration order; the model in the original space is the same).

      ent_dim[e]                        projected dimension of entity e
      coef_ptr[e] .. coef_ptr[e + 1]    its slice of the concatenated projected coefficient vector (pml_batch layout)
      col_local[k]                      projected index of stored entry k (what pml_batch_create_csr takes)
      original_index[j]                 original feature index of concatenated projected coefficient j
    """

    def __init__(self, ent_row_ptr, row_ptr, col, original_dimension):
        ent_row_ptr = np.asarray(ent_row_ptr, dtype=np.int64)
        row_ptr = np.asarray(row_ptr, dtype=np.int64)
        col = np.asarray(col, dtype=np.int64)
        n_entities = len(ent_row_ptr) - 1
        self.original_dimension = int(original_dimension)
        if col.size and (col.min() < 0 or col.max() >= self.original_dimension):
            raise ValueError("feature index outside [0, %d)" % self.original_dimension)
        entries_per_entity = row_ptr[ent_row_ptr[1:]] - row_ptr[ent_row_ptr[:-1]]
        ent_of_entry = np.repeat(np.arange(n_entities, dtype=np.int64), entries_per_entity)
        keys, inverse = np.unique(ent_of_entry * self.original_dimension + col, return_inverse=True)
        ent_of_key = keys // self.original_dimension
        self.original_index = (keys % self.original_dimension).astype(np.int64)
        self.ent_dim = np.bincount(ent_of_key, minlength=n_entities).astype(np.int32)
        self.coef_ptr = np.concatenate([[0], np.cumsum(self.ent_dim)]).astype(np.int64)
        self.col_local = (inverse - self.coef_ptr[ent_of_entry]).astype(np.int32)

    def project_coefficients(self, coefficients_original, e):
        """Projector.projectCoefficients for entity e: original space -> its projected space."""
        a, b = int(self.coef_ptr[e]), int(self.coef_ptr[e + 1])
        return np.asarray(coefficients_original, dtype=np.float64)[self.original_index[a:b]]

    def to_original_space(self, coefficients_projected_all, e):
        """IndexMapProjector.projectCoefficients back (projectedToOriginalSpaceMap): entity e's model as a dense vector
        of the original dimension, from the concatenated projected coefficients of all entities."""
        a, b = int(self.coef_ptr[e]), int(self.coef_ptr[e + 1])
        out = np.zeros(self.original_dimension)
        out[self.original_index[a:b]] = np.asarray(coefficients_projected_all, dtype=np.float64)[a:b]
        return out
