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
