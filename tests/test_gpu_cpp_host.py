"""GPU: the C++ host layer linked with libmimgpu.so (the product path: C++ solve() loops above
the C ABI, CUDA kernels below), compared with the reference invariants and the Python mirror."""
import pytest

import cpp_host

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def binary():
    return cpp_host.build_gpu()


@pytest.mark.parametrize("check", cpp_host.CHECKS, ids=lambda f: f.__name__)
def test_cpp_host_gpu(binary, tmp_path, check):
    check(binary, tmp_path, None)
