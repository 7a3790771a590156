"""The C++ host-side mirror (include/photon_ml_b200.hpp): it must compile and link against the CUDA
library (CPU check), and on a GPU its TRON run must agree with the Python driver on the same data."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "test_host_mirror.cpp")
LIB_DIR = os.path.join(ROOT, "photon_ml_b200", "lib")


def _build(tmp_path):
    from photon_ml_b200 import _build
    _build.build_library()
    exe = str(tmp_path / "test_host_mirror")
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), SRC,
                           "-o", exe, "-L", LIB_DIR, "-lphoton_ml_b200", "-Wl,-rpath," + LIB_DIR])
    return exe


def test_cpp_host_mirror_compiles_and_links(tmp_path):
    assert os.path.exists(_build(tmp_path))


@pytest.mark.gpu
def test_cpp_tron_matches_python_driver(tmp_path):
    import photon_ml_b200 as pml
    from photon_ml_b200.function import DistributedGLMLossFunction, NoNormalization
    from photon_ml_b200.optimization import TRON
    from tests import util
    exe = _build(tmp_path)
    path = os.path.join(util.GOLDEN, "a9a_1605.libsvm")
    out = subprocess.check_output([exe, path, "123", "10.0"], text=True)
    got = dict(line.split() for line in out.strip().splitlines())
    assert got["require_size_throws"] == "1"
    assert float(got["batch_worst_diff"]) <= 1e-12 and got["batch_require_throws"] == "1" and got["peer_created"] == "1"
    assert float(got["peer_host_diff"]) == 0.0 and got["peer_slice_ok"] == "1"
    n, d, row_ptr, col, val, label = util.read_libsvm(path, d=123, add_intercept=True)
    ctx = pml.Context(0)
    ds = pml.Dataset.from_csr(ctx, n, d, row_ptr, col, val, label)
    obj = DistributedGLMLossFunction.create(pml.LOGISTIC, l2RegWeight=10.0)
    tron = TRON(NoNormalization())
    w, f = tron.optimize(obj, np.zeros(d), ds)
    assert int(got["iterations"]) == tron.currentState.iter
    assert int(got["evaluations"]) == tron.evaluations
    assert int(got["hv_evaluations"]) == tron.hessian_vector_evaluations
    assert abs(float(got["value_at_zero"]) - n * np.log(2.0)) <= 1e-9 * n
    assert abs(float(got["value"]) - f) <= 1e-9 * abs(f)
    checksum = float(np.dot(np.arange(1, d + 1), w))
    assert abs(float(got["checksum"]) - checksum) <= 1e-6 * max(1.0, abs(checksum))
    diag = obj.hessianDiagonal(ds, w)
    assert abs(float(got["diag0"]) - diag[0]) <= 1e-9 * abs(diag[0])
    ds.close()
    ctx.close()
