"""Host logic that needs no GPU: which FP8 GEMM kernel `ops.cutlass_scaled_mm` picks for the shapes of SURVEY.md
section 8(a) row a5 on a 148-SM B200 (the rule is a pure function so that the table in DESIGN.md section 3.4 is testable),
and the programmatic-dependent-launch level plumbing of the C ABI."""
import pytest

from b200ops import _lib, ops

SMS = 148
# (M, N, K) -> kernel; C2 (Llama-3.1-8B, TP1), C3 (Llama-3.1-70B at TP8 / TP2), large-M twins
CASES = [
    ((256, 6144, 4096), "v5"), ((256, 4096, 4096), "v5"), ((256, 28672, 4096), "v3"), ((256, 4096, 14336), "v5"),
    ((128, 1280, 8192), "v5"), ((128, 8192, 1024), "v5"), ((128, 7168, 8192), "v5"), ((128, 8192, 3584), "v5"),
    ((128, 28672, 8192), "v3"), ((128, 8192, 14336), "v5"),
    ((16, 4096, 4096), "v4"), ((64, 4096, 14336), "v4"), ((64, 6144, 4096), "v4"), ((64, 28672, 4096), "v3"),
    ((32, 14336, 4096), "v5"),           # small batch but too many weight tiles for one wave of split-K clusters
    ((4096, 4096, 4096), "v3"), ((257, 256, 128), "v3"), ((1, 256, 128), "v4"),
]


@pytest.mark.parametrize("mnk,expect", CASES)
def test_pick_gemm_impl(mnk, expect):
    assert ops.pick_gemm_impl(*mnk, True, SMS) == expect


def test_pick_gemm_impl_unaligned_output_goes_to_the_one_cta_kernel():
    # v4 / v5 write their tiles with TMA stores: an output whose base or row stride is not 16-byte aligned cannot
    # use them; v3 has its own plain-store path
    assert ops.pick_gemm_impl(256, 4096, 4096, False, SMS) == "v2"
    assert ops.pick_gemm_impl(16, 4096, 4096, False, SMS) == "v2"
    assert ops.pick_gemm_impl(4096, 4096, 4096, False, SMS) == "v3"


def test_pick_gemm_impl_scales_with_the_sm_count():
    # fewer SMs -> fewer CTA pairs -> the persistent kernel takes over earlier
    assert ops.pick_gemm_impl(256, 6144, 4096, True, 148) == "v5"
    assert ops.pick_gemm_impl(256, 6144, 4096, True, 40) == "v3"


def test_set_pdl_round_trips_without_a_gpu():
    """b200_set_pdl is host state only: levels clamp to 0..2 and the previous level comes back."""
    lib = _lib.load()
    first = lib.b200_set_pdl(0)
    assert first in (0, 1, 2)
    assert lib.b200_set_pdl(2) == 0
    assert lib.b200_set_pdl(7) == 2          # clamps
    assert lib.b200_set_pdl(-3) == 2
    assert lib.b200_set_pdl(first) == 0
    assert ops.set_pdl(first) == first
