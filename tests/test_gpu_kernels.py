"""GPU unit tests of individual device routines."""
import pytest

pytestmark = pytest.mark.gpu


def test_division_and_widening_shortcuts_are_exact(built):
    """div_exact == IEEE float division and widen_positive == (double) on 2^28 operands
    spanning the sweep's ranges (voigt.cuh): bit-exact, zero mismatches allowed."""
    assert built.selftest_arithmetic(1 << 28, seed=12345) == 0


def test_fma_peak_is_plausible(built):
    fp64, fp32 = built.measure_fma_peak(0)
    assert 10. < fp64 < 60. and 30. < fp32 < 120.


def test_per_line_voigt_matches_the_restatement_bit_for_bit(built):
    """voigt_line_eval (hoisted constants, fp32x2 CPF12, slow-path-free division) ==
    voigt_core (RFM_voigt.c:109-277 sample by sample) on 2^28 (x, Y) pairs over all regions."""
    assert built.selftest_voigt(1 << 28, seed=4321) == 0
