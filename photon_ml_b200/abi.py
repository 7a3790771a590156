"""ctypes binding of include/photon_ml_b200.h -- the same symbols the JNI shim binds.

There is no CPU fallback: loading fails loudly if the CUDA library has not been built, and
Context() fails loudly if no CUDA device is visible.
"""
import ctypes
import os
import weakref

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# PML_LIB points at another build of the same library (kernel A/B runs: tools/); the default is the in-tree build
LIB_PATH = os.environ.get("PML_LIB") or os.path.join(HERE, "lib", "libphoton_ml_b200.so")

OK, EINVAL, ECUDA, ENOMEM, ESTATE = 0, 1, 2, 3, 4
LOGISTIC, POISSON, SQUARED, SMOOTHED_HINGE = 0, 1, 2, 3
F32, F64 = 0, 1
STORE_F32, STORE_F64, STORE_AUTO, STORE_ONES = 0, 1, 2, 3

# every symbol include/photon_ml_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "pml_last_error", "pml_ctx_create", "pml_ctx_destroy", "pml_ctx_synchronize", "pml_ctx_stream",
    "pml_host_alloc", "pml_host_free", "pml_dataset_create_csr", "pml_dataset_create_csr_dev",
    "pml_dataset_destroy", "pml_dataset_get_info", "pml_dataset_set_offsets", "pml_dataset_set_weights",
    "pml_dataset_set_labels", "pml_dataset_set_normalization", "pml_value_and_gradient", "pml_hessian_vector",
    "pml_hessian_diagonal", "pml_scores", "pml_value_and_gradient_raw_dev", "pml_hessian_vector_raw_dev",
    "pml_finalize_dev", "pml_value_and_gradient_dev", "pml_hessian_vector_dev", "pml_ctx_launch_count",
    "pml_dataset_set_margin_variant", "pml_ctx_set_profiling", "pml_ctx_read_profile",
    "pml_value_noalloc_dev", "pml_batch_create_csr", "pml_batch_destroy", "pml_batch_dataset", "pml_batch_coef_ptr",
    "pml_batch_value_and_gradient", "pml_batch_value_and_gradient_dev", "pml_scores_dev", "pml_dataset_set_offsets_dev", "pml_offsets_from_scores_dev",
    "pml_batch_hessian_vector", "pml_batch_hessian_vector_dev", "pml_batch_set_offsets", "pml_batch_set_active",
    "pml_peer_create", "pml_peer_handle", "pml_peer_connect", "pml_peer_destroy", "pml_peer_buffers",
    "pml_value_and_gradient_allreduce_dev", "pml_hessian_vector_allreduce_dev",
    "pml_peer_multicast_export", "pml_peer_multicast_join", "pml_peer_multicast_bind", "pml_peer_multicast_release",
    "pml_peer_transport", "pml_ctx_pci_bus_id", "pml_peer_merge_dev",
    "pml_batch_set_active_dev", "pml_batch_segment_dot_dev", "pml_batch_segment_axpby_dev", "pml_loss_from_scores_dev",
    "pml_batch_set_normalization",
    "pml_peer_slice", "pml_value_and_gradient_allreduce", "pml_hessian_vector_allreduce", "pml_peer_status",
    "pml_peer_emulate", "pml_hessian_diagonal_allreduce", "pml_batch_cg_step_dev", "pml_batch_lbfgs_direction_dev", "pml_batch_lbfgs_push_dev",
]


class PhotonError(RuntimeError):
    """PML_ECUDA / PML_ENOMEM / PML_ESTATE (a RuntimeException on the Scala side)."""

    def __init__(self, code, message):
        super().__init__("photon_ml_b200 error %d: %s" % (code, message))
        self.code = code


class IllegalArgumentError(ValueError):
    """PML_EINVAL: a `require` of the reference failed (IllegalArgumentException in Scala)."""

    code = EINVAL


class DatasetInfo(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int64), ("nnz", ctypes.c_int64), ("d", ctypes.c_int32),
                ("value_bytes", ctypes.c_int32), ("margin_variant", ctypes.c_int32), ("n_chunks", ctypes.c_int32),
                ("nonempty_cols", ctypes.c_int64), ("device_bytes", ctypes.c_int64),
                ("vg_algorithmic_bytes", ctypes.c_int64), ("hv_algorithmic_bytes", ctypes.c_int64),
                ("tiled_hot_cols", ctypes.c_int32), ("tiled_tiles", ctypes.c_int32), ("tiled_tile_rows", ctypes.c_int32),
                ("sliced_ell_padding_permille", ctypes.c_int32)]


_lib = None


def load():
    """dlopen the CUDA library (no compute happens here; safe without a GPU)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, u64, dbl = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_uint64, ctypes.c_double
    pp = ctypes.POINTER(vp)
    lib.pml_last_error.restype = ctypes.c_char_p
    lib.pml_last_error.argtypes = []
    sigs = {
        "pml_ctx_create": [ctypes.c_int, pp],
        "pml_ctx_destroy": [vp],
        "pml_ctx_synchronize": [vp],
        "pml_ctx_stream": [vp, pp],
        "pml_host_alloc": [pp, i64],
        "pml_host_free": [vp],
        "pml_dataset_create_csr": [vp, i64, i32, vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp, pp],
        "pml_dataset_create_csr_dev": [vp, i64, i32, vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp, pp],
        "pml_dataset_destroy": [vp],
        "pml_dataset_get_info": [vp, ctypes.POINTER(DatasetInfo)],
        "pml_dataset_set_offsets": [vp, vp],
        "pml_dataset_set_weights": [vp, vp],
        "pml_dataset_set_labels": [vp, vp],
        "pml_dataset_set_normalization": [vp, vp, vp, i32],
        "pml_value_and_gradient": [vp, ctypes.c_int, vp, ctypes.POINTER(dbl), vp],
        "pml_hessian_vector": [vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_hessian_diagonal": [vp, ctypes.c_int, vp, vp],
        "pml_scores": [vp, vp, vp],
        "pml_value_and_gradient_raw_dev": [vp, ctypes.c_int, vp, vp],
        "pml_hessian_vector_raw_dev": [vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_finalize_dev": [vp, ctypes.c_int, vp, vp, vp],
        "pml_value_and_gradient_dev": [vp, ctypes.c_int, vp, vp, vp],
        "pml_hessian_vector_dev": [vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_ctx_launch_count": [vp, ctypes.POINTER(i64)],
        "pml_dataset_set_margin_variant": [vp, ctypes.c_int],
        "pml_ctx_set_profiling": [vp, ctypes.c_int],
        "pml_ctx_read_profile": [vp, ctypes.POINTER(dbl), ctypes.POINTER(i64)],
        "pml_value_noalloc_dev": [ctypes.c_int, vp, i64, i32, vp, vp, vp, ctypes.c_int, vp, vp, vp, ctypes.c_int, vp, vp],
    }
    sigs.update({
        "pml_batch_create_csr": [vp, i64, vp, vp, vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp, pp],
        "pml_batch_destroy": [vp],
        "pml_batch_dataset": [vp, pp],
        "pml_batch_coef_ptr": [vp, vp],
        "pml_batch_value_and_gradient": [vp, ctypes.c_int, vp, vp, vp],
        "pml_batch_value_and_gradient_dev": [vp, ctypes.c_int, vp, vp, vp],
        "pml_scores_dev": [vp, vp, vp],
        "pml_dataset_set_offsets_dev": [vp, vp],
        "pml_offsets_from_scores_dev": [vp, i64, vp, vp, vp, vp],
        "pml_batch_hessian_vector": [vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_batch_hessian_vector_dev": [vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_batch_set_offsets": [vp, vp],
        "pml_batch_set_active": [vp, vp],
        "pml_peer_create": [vp, ctypes.c_int, ctypes.c_int, i32, pp],
        "pml_peer_handle": [vp, vp],
        "pml_peer_connect": [vp, vp],
        "pml_peer_destroy": [vp],
        "pml_peer_buffers": [vp, pp, pp],
        "pml_peer_multicast_export": [vp, ctypes.POINTER(ctypes.c_int)],
        "pml_peer_multicast_join": [vp, ctypes.c_int],
        "pml_peer_multicast_bind": [vp],
        "pml_peer_multicast_release": [vp],
        "pml_peer_transport": [vp, ctypes.POINTER(ctypes.c_int)],
        "pml_ctx_pci_bus_id": [vp, ctypes.c_char_p, ctypes.c_int],
        "pml_peer_merge_dev": [vp, vp, ctypes.c_int],
        "pml_loss_from_scores_dev": [vp, ctypes.c_int, vp, vp],
        "pml_batch_set_active_dev": [vp, vp],
        "pml_batch_set_normalization": [vp, vp, vp, vp],
        "pml_batch_segment_dot_dev": [vp, vp, vp, vp],
        "pml_batch_segment_axpby_dev": [vp, vp, vp, vp, vp, vp, vp],
        "pml_batch_cg_step_dev": [vp, vp, dbl, vp, vp, vp, vp, vp, vp, vp],
        "pml_batch_lbfgs_direction_dev": [vp, vp, ctypes.c_int, vp, vp, vp, vp, vp, vp, vp],
        "pml_batch_lbfgs_push_dev": [vp, vp, ctypes.c_int, vp, vp, vp, vp, vp, vp, vp, vp],
        "pml_value_and_gradient_allreduce_dev": [vp, vp, ctypes.c_int, vp, vp, vp],
        "pml_hessian_vector_allreduce_dev": [vp, vp, ctypes.c_int, vp, u64, vp, vp],
        "pml_peer_slice": [vp, ctypes.POINTER(i64), ctypes.POINTER(i64)],
        "pml_value_and_gradient_allreduce": [vp, vp, ctypes.c_int, vp, ctypes.POINTER(dbl), vp, ctypes.c_int],
        "pml_hessian_vector_allreduce": [vp, vp, ctypes.c_int, vp, u64, vp, vp, ctypes.c_int],
        "pml_peer_status": [vp, ctypes.POINTER(ctypes.c_int)],
        "pml_hessian_diagonal_allreduce": [vp, vp, ctypes.c_int, vp, vp, ctypes.c_int],
        "pml_peer_emulate": [vp, ctypes.c_int, i32, ctypes.c_int, vp, vp, vp, vp, vp, vp],
    })
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc):
    if rc == OK:
        return
    msg = load().pml_last_error().decode("utf-8", "replace")
    if rc == EINVAL:
        raise IllegalArgumentError(msg)
    raise PhotonError(rc, msg)


def _np(a, dtype, size=None, name="array"):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=dtype)
    if size is not None and a.shape != (size,):
        raise IllegalArgumentError("Size mismatch. %s size: %d != %d." % (name, a.size, size))
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def value_noalloc_dev(device, stream, n, d, row_ptr_dev, col_dev, val_dev, values_dtype, label_dev, offset_dev,
                      weight_dev, loss, coef_dev, out_value_dev):
    """pml_value_noalloc_dev: objective value from caller-owned device CSR arrays, for callers that
    must not allocate (all arguments after `stream` are sizes or integer device addresses).
    Asynchronous on `stream`; at most 16,777,216 rows; calls on one device must not overlap."""
    check(load().pml_value_noalloc_dev(int(device), stream, n, d, row_ptr_dev, col_dev, val_dev, values_dtype, label_dev,
                                       offset_dev, weight_dev, loss, coef_dev, out_value_dev))


def peer_emulate(ctx, raw, factors=None, shifts=None, vec_in=None, repeats=1):
    """pml_peer_emulate: the exchange of len(raw) ranks on one device.  raw: (world, d+2).  Returns the (world, d+1)
    results, and with vec_in (world, d) also the (world, d) assembled input vectors."""
    raw = np.ascontiguousarray(raw, dtype=np.float64)
    world, d = raw.shape[0], raw.shape[1] - 2
    out = np.empty((world, d + 1))
    f = None if factors is None else _np(factors, np.float64, d, "factors")
    s = None if shifts is None else _np(shifts, np.float64, d, "shifts")
    vin = None if vec_in is None else np.ascontiguousarray(vec_in, dtype=np.float64)
    vout = None if vec_in is None else np.empty((world, d))
    check(load().pml_peer_emulate(ctx._h, world, d, repeats, _ptr(raw), None if f is None else _ptr(f),
                                  None if s is None else _ptr(s), _ptr(out), None if vin is None else _ptr(vin),
                                  None if vout is None else _ptr(vout)))
    return out if vec_in is None else (out, vout)


def offsets_from_scores_dev(ctx, n, base_offsets_dev, scores_dev, row_map_dev, out_offsets_dev):
    """out[i] = base[i] + scores[row_map[i]] on the device (integer device addresses; None = absent)."""
    check(load().pml_offsets_from_scores_dev(ctx._h, n, base_offsets_dev, scores_dev, row_map_dev, out_offsets_dev))


class Context:
    """One GPU, one stream (pml_ctx)."""

    def __init__(self, device=0):
        self._h = ctypes.c_void_p()
        check(load().pml_ctx_create(int(device), ctypes.byref(self._h)))
        self.device = int(device)

    def synchronize(self):
        check(load().pml_ctx_synchronize(self._h))

    @property
    def stream(self):
        s = ctypes.c_void_p()
        check(load().pml_ctx_stream(self._h, ctypes.byref(s)))
        return s.value or 0

    @property
    def launch_count(self):
        v = ctypes.c_int64()
        check(load().pml_ctx_launch_count(self._h, ctypes.byref(v)))
        return v.value

    KERNEL_KINDS = ["prepare_vectors", "row_pass", "col_pass", "fixup", "finalize", "row_epilogue", "col_pass_tiled",
                    "entity_values", "peer_merge"]

    def set_profiling(self, on):
        check(load().pml_ctx_set_profiling(self._h, 1 if on else 0))

    def read_profile(self):
        """{kernel class: (total ms, launches)} since the last read (synchronises)."""
        ms = (ctypes.c_double * len(self.KERNEL_KINDS))()
        cnt = (ctypes.c_int64 * len(self.KERNEL_KINDS))()
        check(load().pml_ctx_read_profile(self._h, ms, cnt))
        return {k: (ms[i], cnt[i]) for i, k in enumerate(self.KERNEL_KINDS)}

    def _adopt(self, child):
        """Handles created on this context are closed before it (whatever order the interpreter drops them in)."""
        if not hasattr(self, "_children"):
            self._children = weakref.WeakSet()
        self._children.add(child)

    def close(self):
        if self._h:
            for child in list(getattr(self, "_children", ())):
                try:
                    child.close()
                except Exception:
                    pass
            load().pml_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Dataset:
    """LabeledPoint rows packed once into device CSR + CSC (pml_dataset)."""

    def __init__(self, ctx, handle, keepalive=None):
        self.ctx = ctx
        self._h = handle
        self._keep = keepalive
        ctx._adopt(self)
        info = self.info()
        self.n, self.d, self.nnz = info.n, info.d, info.nnz

    @classmethod
    def from_csr(cls, ctx, n, d, row_ptr, col, val, label, offset=None, weight=None, storage=STORE_AUTO):
        row_ptr = _np(row_ptr, np.int64, n + 1, "row_ptr")
        col = _np(col, np.int32)
        val = np.ascontiguousarray(val)
        if val.dtype == np.float32:
            dtype = F32
        else:
            val = val.astype(np.float64, copy=False)
            dtype = F64
        label = _np(label, np.float64, n, "labels")
        offset = _np(offset, np.float64, n, "offsets")
        weight = _np(weight, np.float64, n, "weights")
        h = ctypes.c_void_p()
        check(load().pml_dataset_create_csr(ctx._h, n, d, _ptr(row_ptr), _ptr(col), _ptr(val), dtype, storage,
                                            _ptr(label), _ptr(offset), _ptr(weight), ctypes.byref(h)))
        return cls(ctx, h)

    @classmethod
    def from_device_pointers(cls, ctx, n, d, row_ptr_dev, col_dev, val_dev, values_dtype, label_dev, offset_dev=None,
                             weight_dev=None, storage=STORE_AUTO):
        """All arguments are integer device addresses (e.g. torch.Tensor.data_ptr())."""
        h = ctypes.c_void_p()
        check(load().pml_dataset_create_csr_dev(ctx._h, n, d, row_ptr_dev, col_dev, val_dev, values_dtype, storage,
                                                label_dev, offset_dev, weight_dev, ctypes.byref(h)))
        return cls(ctx, h)

    def info(self):
        out = DatasetInfo()
        check(load().pml_dataset_get_info(self._h, ctypes.byref(out)))
        return out

    def set_offsets(self, offsets):
        check(load().pml_dataset_set_offsets(self._h, _ptr(_np(offsets, np.float64, self.n, "offsets"))))

    def set_weights(self, weights):
        check(load().pml_dataset_set_weights(self._h, _ptr(_np(weights, np.float64, self.n, "weights"))))

    def set_labels(self, labels):
        check(load().pml_dataset_set_labels(self._h, _ptr(_np(labels, np.float64, self.n, "labels"))))

    def set_normalization(self, factors=None, shifts=None, intercept=None):
        f = _np(factors, np.float64, self.d, "Factors vector")
        s = _np(shifts, np.float64, self.d, "Shifts vector")
        check(load().pml_dataset_set_normalization(self._h, _ptr(f), _ptr(s), -1 if intercept is None else int(intercept)))
        # what is set on the device right now: None, or a token the objective classes (function.py) replace by
        # the NormalizationContext object they applied
        self._applied_norm = None if (f is None and s is None) else object()

    def set_margin_variant(self, lanes_per_row):
        check(load().pml_dataset_set_margin_variant(self._h, int(lanes_per_row)))

    # ---- host-buffer evaluations -------------------------------------------------------------
    def value_and_gradient(self, loss, coef, want_gradient=True, out=None):
        coef = _np(coef, np.float64, self.d, "Coefficient")
        value = ctypes.c_double()
        grad = None
        if want_gradient:
            grad = out if out is not None else np.empty(self.d, dtype=np.float64)
        check(load().pml_value_and_gradient(self._h, loss, _ptr(coef), ctypes.byref(value), _ptr(grad)))
        return value.value, grad

    def hessian_vector(self, loss, coef, v, tag=0, out=None):
        coef = _np(coef, np.float64, self.d, "Coefficient")
        v = _np(v, np.float64, self.d, "Multiply vector")
        hv = out if out is not None else np.empty(self.d, dtype=np.float64)
        check(load().pml_hessian_vector(self._h, loss, _ptr(coef), int(tag), _ptr(v), _ptr(hv)))
        return hv

    def hessian_diagonal(self, loss, coef):
        coef = _np(coef, np.float64, self.d, "Coefficient")
        diag = np.empty(self.d, dtype=np.float64)
        check(load().pml_hessian_diagonal(self._h, loss, _ptr(coef), _ptr(diag)))
        return diag

    def scores(self, coef):
        coef = _np(coef, np.float64, self.d, "Coefficient")
        out = np.empty(self.n, dtype=np.float64)
        check(load().pml_scores(self._h, _ptr(coef), _ptr(out)))
        return out

    # ---- device-buffer evaluations (integer device addresses) --------------------------------
    def scores_dev(self, coef_dev, out_scores_dev):
        check(load().pml_scores_dev(self._h, coef_dev, out_scores_dev))

    def set_offsets_dev(self, offsets_dev):
        check(load().pml_dataset_set_offsets_dev(self._h, offsets_dev))

    def value_and_gradient_raw_dev(self, loss, coef_dev, out_raw_dev):
        check(load().pml_value_and_gradient_raw_dev(self._h, loss, coef_dev, out_raw_dev))

    def hessian_vector_raw_dev(self, loss, coef_dev, v_dev, out_raw_dev, tag=0):
        check(load().pml_hessian_vector_raw_dev(self._h, loss, coef_dev, int(tag), v_dev, out_raw_dev))

    def finalize_dev(self, loss, raw_dev, out_value_dev, out_vec_dev):
        check(load().pml_finalize_dev(self._h, loss, raw_dev, out_value_dev, out_vec_dev))

    def value_and_gradient_dev(self, loss, coef_dev, out_value_dev, out_grad_dev):
        check(load().pml_value_and_gradient_dev(self._h, loss, coef_dev, out_value_dev, out_grad_dev))

    def hessian_vector_dev(self, loss, coef_dev, v_dev, out_hv_dev, tag=0):
        check(load().pml_hessian_vector_dev(self._h, loss, coef_dev, int(tag), v_dev, out_hv_dev))

    def loss_from_scores_dev(self, loss, scores_dev, out_value_dev):
        """sum_i wt_i * l(scores_i + offset_i, y_i) for scores already on the device (the training-loss evaluator)."""
        check(load().pml_loss_from_scores_dev(self._h, loss, scores_dev, out_value_dev))

    def close(self):
        if self._h and not getattr(self, "_borrowed", False):
            load().pml_dataset_destroy(self._h)
        self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Batch:
    """Many per-entity local problems evaluated together (pml_batch): the random-effect path."""

    def __init__(self, ctx, ent_row_ptr, ent_dim, row_ptr, col_local, val, label, offset=None, weight=None,
                 storage=STORE_AUTO):
        self.ctx = ctx
        ent_row_ptr = _np(ent_row_ptr, np.int64)
        self.n_entities = len(ent_row_ptr) - 1
        ent_dim = _np(ent_dim, np.int32, self.n_entities, "ent_dim")
        n = int(ent_row_ptr[-1]) if self.n_entities > 0 else 0
        row_ptr = _np(row_ptr, np.int64, n + 1, "row_ptr")
        col_local = _np(col_local, np.int32)
        val = np.ascontiguousarray(val)
        if val.dtype == np.float32:
            dtype = F32
        else:
            val = val.astype(np.float64, copy=False)
            dtype = F64
        label = _np(label, np.float64, n, "labels")
        offset = _np(offset, np.float64, n, "offsets")
        weight = _np(weight, np.float64, n, "weights")
        self._h = ctypes.c_void_p()
        check(load().pml_batch_create_csr(ctx._h, self.n_entities, _ptr(ent_row_ptr), _ptr(ent_dim), _ptr(row_ptr),
                                          _ptr(col_local), _ptr(val), dtype, storage, _ptr(label), _ptr(offset),
                                          _ptr(weight), ctypes.byref(self._h)))
        self.coef_ptr = np.empty(self.n_entities + 1, dtype=np.int64)
        check(load().pml_batch_coef_ptr(self._h, _ptr(self.coef_ptr)))
        dsh = ctypes.c_void_p()
        check(load().pml_batch_dataset(self._h, ctypes.byref(dsh)))
        self.dataset = Dataset(ctx, dsh)   # block-diagonal view: set_offsets, hessian_vector, ... for all entities
        self.dataset._borrowed = True
        self.d_total = self.dataset.d
        ctx._adopt(self)

    def value_and_gradient(self, loss, coefs, want_gradient=True):
        coefs = _np(coefs, np.float64, self.d_total, "Coefficient")
        values = np.empty(self.n_entities, dtype=np.float64)
        grads = np.empty(self.d_total, dtype=np.float64) if want_gradient else None
        check(load().pml_batch_value_and_gradient(self._h, loss, _ptr(coefs), _ptr(values), _ptr(grads)))
        return values, grads

    def value_and_gradient_dev(self, loss, coefs_dev, out_values_dev, out_grads_dev):
        """Device buffers (integer addresses); asynchronous on the context's stream."""
        check(load().pml_batch_value_and_gradient_dev(self._h, loss, coefs_dev, out_values_dev, out_grads_dev))

    def hessian_vector(self, loss, coefs, vs, tag=0):
        """Concatenated per-entity Hessian-vector products (SingleNodeGLMLossFunction.hessianVector for every entity)."""
        coefs = _np(coefs, np.float64, self.d_total, "Coefficient")
        vs = _np(vs, np.float64, self.d_total, "Multiply vector")
        out = np.empty(self.d_total, dtype=np.float64)
        check(load().pml_batch_hessian_vector(self._h, loss, _ptr(coefs), tag, _ptr(vs), _ptr(out)))
        return out

    def hessian_vector_dev(self, loss, coefs_dev, vs_dev, out_hvs_dev, tag=0):
        check(load().pml_batch_hessian_vector_dev(self._h, loss, coefs_dev, tag, vs_dev, out_hvs_dev))

    def set_offsets(self, offsets):
        """Residual offsets of all rows, concatenated entity order (LocalDataSet.addScoresToOffsets)."""
        n = self.dataset.n
        check(load().pml_batch_set_offsets(self._h, _ptr(_np(offsets, np.float64, n, "offsets"))))

    def set_active(self, active):
        """active[e] falsy: entity e's value / gradient / Hessian-vector entries come back as 0.  None = all."""
        a = None if active is None else np.ascontiguousarray(np.asarray(active) != 0, dtype=np.uint8)
        if a is not None and a.shape != (self.n_entities,):
            raise IllegalArgumentError("Size mismatch. active size: %d != %d." % (a.size, self.n_entities))
        check(load().pml_batch_set_active(self._h, _ptr(a)))

    def set_normalization(self, factors=None, shifts=None, intercept_local=None):
        """Per-entity NormalizationContext: factors / shifts concatenated like the coefficients, intercept_local[e] =
        entity e's intercept index in its own space (-1 = none).  All None removes it."""
        f = _np(factors, np.float64, self.d_total, "factors")
        s = _np(shifts, np.float64, self.d_total, "shifts")
        ic = _np(intercept_local, np.int32, self.n_entities, "intercept ids")
        check(load().pml_batch_set_normalization(self._h, _ptr(f), _ptr(s), _ptr(ic)))

    def set_active_dev(self, active_dev):
        """active_dev: device address of E bytes (copied), or None for all active."""
        check(load().pml_batch_set_active_dev(self._h, active_dev))

    def segment_dot_dev(self, x_dev, y_dev, out_dev):
        """out[e] = sum_j x_j y_j over entity e's coefficients (device addresses; E doubles out)."""
        check(load().pml_batch_segment_dot_dev(self._h, x_dev, y_dev, out_dev))

    def lbfgs_direction_dev(self, go_dev, m, S_dev, Y_dev, rho_dev, head_dev, len_dev, grad_dev, out_dev):
        check(load().pml_batch_lbfgs_direction_dev(self._h, go_dev, int(m), S_dev, Y_dev, rho_dev, head_dev, len_dev, grad_dev,
                                                   out_dev))

    def lbfgs_push_dev(self, mask_dev, m, rho_new_dev, s_dev, y_dev, S_dev, Y_dev, rho_dev, head_dev, len_dev):
        check(load().pml_batch_lbfgs_push_dev(self._h, mask_dev, int(m), rho_new_dev, s_dev, y_dev, S_dev, Y_dev, rho_dev,
                                              head_dev, len_dev))

    def cg_step_dev(self, run_dev, l2, direction_dev, hd_dev, step_dev, residual_dev, rtr_dev, delta_dev, out_over_dev):
        check(load().pml_batch_cg_step_dev(self._h, run_dev, float(l2), direction_dev, hd_dev, step_dev, residual_dev, rtr_dev,
                                           delta_dev, out_over_dev))

    def segment_axpby_dev(self, a_dev, x_dev, c_dev, y_dev, out_dev, active_dev=None):
        """out_j = a[e(j)] x_j + c[e(j)] y_j; a / c None = 1, y None = no second term; inactive entities keep out_j."""
        check(load().pml_batch_segment_axpby_dev(self._h, a_dev, x_dev, c_dev, y_dev, active_dev, out_dev))

    def close(self):
        if self._h:
            self.dataset.close()
            load().pml_batch_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


PEER_HANDLE_BYTES = 64
MAX_PEERS = 8


class _DeviceArray:
    """A view of library-owned device memory that torch.as_tensor(..., device='cuda') can wrap without a copy."""

    def __init__(self, ptr, count, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 2}


class Peer:
    """This rank's end of the merge over NVLink peer memory (pml_peer): one process per GPU.

    `exchange(handle_bytes) -> [handle_bytes of every rank, in rank order]` is the only host-side communication;
    the default uses torch.distributed.all_gather_object on the default process group."""

    def __init__(self, ctx, d, rank=None, world=None, exchange=None, multicast=None):
        if rank is None or world is None:
            import torch.distributed as dist
            rank, world = dist.get_rank(), dist.get_world_size()
        self.ctx, self.d, self.rank, self.world = ctx, int(d), int(rank), int(world)
        self._h = ctypes.c_void_p()
        check(load().pml_peer_create(ctx._h, self.rank, self.world, self.d, ctypes.byref(self._h)))
        mine = ctypes.create_string_buffer(PEER_HANDLE_BYTES)
        check(load().pml_peer_handle(self._h, mine))
        if exchange is None:
            def exchange(h):
                import torch.distributed as dist
                got = [None] * self.world
                dist.all_gather_object(got, h)
                return got
        handles = exchange(mine.raw) if self.world > 1 else [mine.raw]
        blob = b"".join(handles)
        if len(blob) != PEER_HANDLE_BYTES * self.world:
            raise IllegalArgumentError("expected %d handles of %d bytes" % (self.world, PEER_HANDLE_BYTES))
        check(load().pml_peer_connect(self._h, blob))
        self.transport = "p2p"
        # The multicast transport is opt-in: fp64 multimem requests are 8 bytes each (no vector form), and measured
        # on B200 the switch reduces them slower than peer loads move 16-byte pairs once d exceeds ~10^5.
        if multicast is None:
            multicast = os.environ.get("PML_MULTICAST") == "1"
        if multicast and self.world > 1:
            self.multicast_error = self._try_multicast(exchange)
            if self.multicast_error is None:
                self.transport = "multicast"
        raw, res = ctypes.c_void_p(), ctypes.c_void_p()
        check(load().pml_peer_buffers(self._h, ctypes.byref(raw), ctypes.byref(res)))
        self.raw_ptr, self.result_ptr = raw.value, res.value
        ctx._adopt(self)

    def _try_multicast(self, exchange):
        """Binds the ranks' slabs to one NVSwitch multicast object; returns None on success, else why not (every
        rank returns the same verdict, and the peer-to-peer transport stays in place)."""
        import socket
        import tempfile
        lib = load()
        bus = ctypes.create_string_buffer(64)
        check(lib.pml_ctx_pci_bus_id(self.ctx._h, bus, 64))
        buses = exchange(bus.value)
        if len(set(buses)) != self.world:
            return "ranks share a GPU (%s)" % ", ".join(b.decode() for b in buses)
        fd = ctypes.c_int(-1)
        ok, why, path, srv = True, None, None, None
        if self.rank == 0:
            rc = lib.pml_peer_multicast_export(self._h, ctypes.byref(fd))
            ok = rc == OK
            if not ok:
                why = lib.pml_last_error().decode("utf-8", "replace")
            else:
                path = os.path.join(tempfile.gettempdir(), "pml_mc_%d_%x.sock" % (os.getpid(), id(self)))
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(path)
                srv.listen(self.world)
                srv.settimeout(60)
        ok, why, path = exchange((ok, why, path))[0]
        if not ok:
            return why
        mine, err = -1, None
        try:
            if self.rank == 0:
                for _ in range(self.world - 1):
                    conn, _ = srv.accept()
                    socket.send_fds(conn, [b"m"], [fd.value])
                    conn.close()
            else:
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                c.settimeout(60)
                c.connect(path)
                _, fds, _, _ = socket.recv_fds(c, 16, 1)
                c.close()
                mine = fds[0]
        except OSError as e:
            err = "passing the multicast file descriptor failed: %s" % e
        finally:
            if srv is not None:
                srv.close()
                try:
                    os.unlink(path)
                except OSError:
                    pass
        if err is None and lib.pml_peer_multicast_join(self._h, mine) != OK:
            err = lib.pml_last_error().decode("utf-8", "replace")
        for f in (mine, fd.value):
            if f >= 0:
                os.close(f)
        errs = [e for e in exchange(err) if e is not None]
        if not errs and lib.pml_peer_multicast_bind(self._h) != OK:
            err = lib.pml_last_error().decode("utf-8", "replace")
        errs = errs or [e for e in exchange(err) if e is not None]
        if errs:
            lib.pml_peer_multicast_release(self._h)
            return errs[0]
        return None

    def result_array(self):
        """[vector(d), value] of the last *_allreduce_dev call, as a __cuda_array_interface__ object (no copy)."""
        return _DeviceArray(self.result_ptr, self.d + 1, self)

    def value_and_gradient(self, ds, loss, coef_dev, out_value_dev=None, out_grad_dev=None):
        check(load().pml_value_and_gradient_allreduce_dev(ds._h, self._h, loss, coef_dev, out_value_dev, out_grad_dev))

    def hessian_vector(self, ds, loss, coef_dev, v_dev, out_hv_dev=None, tag=0):
        check(load().pml_hessian_vector_allreduce_dev(ds._h, self._h, loss, coef_dev, tag, v_dev, out_hv_dev))

    def slice(self):
        """[lo, hi): the part of the coefficient vector this rank uploads, and of the result it downloads, in the
        host-buffer calls below."""
        lo, hi = ctypes.c_int64(), ctypes.c_int64()
        check(load().pml_peer_slice(self._h, ctypes.byref(lo), ctypes.byref(hi)))
        return lo.value, hi.value

    def value_and_gradient_host(self, ds, loss, coef, out_grad, gather_output=False):
        """pml_value_and_gradient_allreduce: host vectors in and out; returns the objective value.  With
        gather_output=False only out_grad[lo:hi] (this rank's slice) is written."""
        value = ctypes.c_double()
        check(load().pml_value_and_gradient_allreduce(ds._h, self._h, loss, _ptr(_np(coef, np.float64, self.d, "coef")),
                                                      ctypes.byref(value), _ptr(out_grad), 1 if gather_output else 0))
        return value.value

    def hessian_vector_host(self, ds, loss, coef, v, out_hv, tag=0, gather_output=False):
        check(load().pml_hessian_vector_allreduce(ds._h, self._h, loss,
                                                  None if coef is None else _ptr(_np(coef, np.float64, self.d, "coef")), tag,
                                                  _ptr(_np(v, np.float64, self.d, "v")), _ptr(out_hv),
                                                  1 if gather_output else 0))

    def hessian_diagonal_host(self, ds, loss, coef, out_diag, gather_output=False):
        check(load().pml_hessian_diagonal_allreduce(ds._h, self._h, loss, _ptr(_np(coef, np.float64, self.d, "coef")),
                                                    _ptr(out_diag), 1 if gather_output else 0))

    def status(self):
        out = ctypes.c_int()
        check(load().pml_peer_status(self._h, ctypes.byref(out)))
        return out.value

    def merge(self, ds, loss):
        """Merge + getVector of whatever the ranks left in their raw buffers (raw_ptr)."""
        check(load().pml_peer_merge_dev(ds._h, self._h, loss))

    def close(self):
        if self._h:
            load().pml_peer_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
