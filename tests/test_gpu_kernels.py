"""GPU numerics: every sm_100a kernel against the fp32 PyTorch evaluation of the same op
(benchmarks/gpu_selftest.py holds the case tables; here each group is one pytest test)."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "benchmarks"))

pytestmark = pytest.mark.gpu


def test_library_is_loaded():
    from picd_b200.ops import lib_available, lib_path
    assert lib_available(), f"native library missing: {lib_path()}"


def test_memory_bound_kernels():
    import gpu_selftest
    assert gpu_selftest.group_basic()


def test_tcgen05_gemm():
    import gpu_selftest
    assert gpu_selftest.group_gemm()


def test_tcgen05_attention():
    import gpu_selftest
    assert gpu_selftest.group_attn()


def test_tcgen05_conv_fprop_dgrad_wgrad():
    import gpu_selftest
    assert gpu_selftest.group_conv()


def test_fused_paths_compose_like_the_fp32_oracle():
    import gpu_selftest
    assert gpu_selftest.group_compose()


def test_native_path_trains_to_accuracy():
    import gpu_selftest
    assert gpu_selftest.group_converge()


def test_models_match_torch_reference():
    import gpu_selftest
    assert gpu_selftest.group_model()
