"""CPU-side checks of the boundary: the library loads, exports every symbol the header declares,
and refuses to run without a device (no CPU fallback)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "photon_ml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pml_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import photon_ml_b200 as pml
    lib = pml.abi.load()
    names = _declared()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), name
    assert sorted(pml.abi.SYMBOLS) == names


def test_no_cpu_fallback():
    import torch
    import photon_ml_b200 as pml
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(pml.PhotonError) as e:
        pml.Context(0)
    assert e.value.code == pml.abi.ECUDA


def test_product_never_touches_the_oracle():
    # no import, include, dlopen or path reference to oracle/ anywhere in the shipped package
    pkg = os.path.join(ROOT, "photon_ml_b200")
    bad = re.compile(r"^\s*(import|from)\s+oracle\b|libglm_oracle|glm_oracle|[\"'/]oracle[\"'/]|#include.*oracle", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp")):
                assert not bad.search(open(os.path.join(dirpath, f)).read()), os.path.join(dirpath, f)


def test_null_handles_are_rejected_not_dereferenced():
    """Argument marshalling of every entry point that takes a handle: NULL comes back as PML_EINVAL (no GPU needed;
    also pins the ctypes signatures in abi.py against the header's argument order)."""
    import ctypes
    import photon_ml_b200 as pml
    lib = pml.abi.load()
    out = ctypes.c_void_p()
    buf = ctypes.create_string_buffer(64)
    calls = [
        lambda: lib.pml_dataset_set_offsets(None, None),
        lambda: lib.pml_value_and_gradient_dev(None, 0, None, None, None),
        lambda: lib.pml_hessian_vector_dev(None, 0, None, 0, None, None),
        lambda: lib.pml_batch_value_and_gradient(None, 0, None, None, None),
        lambda: lib.pml_batch_hessian_vector(None, 0, None, 0, None, None),
        lambda: lib.pml_batch_hessian_vector_dev(None, 0, None, 0, None, None),
        lambda: lib.pml_batch_set_offsets(None, None),
        lambda: lib.pml_batch_set_active(None, None),
        lambda: lib.pml_peer_create(None, 0, 1, 4, ctypes.byref(out)),
        lambda: lib.pml_peer_handle(None, buf),
        lambda: lib.pml_peer_connect(None, buf),
        lambda: lib.pml_peer_buffers(None, ctypes.byref(out), ctypes.byref(out)),
        lambda: lib.pml_value_and_gradient_allreduce_dev(None, None, 0, None, None, None),
        lambda: lib.pml_hessian_vector_allreduce_dev(None, None, 0, None, 0, None, None),
    ]
    for i, call in enumerate(calls):
        assert call() == pml.abi.EINVAL, i
        assert lib.pml_last_error()
    assert lib.pml_peer_destroy(None) == pml.abi.OK
