"""Host logic of the fitting step: the position learning-rate schedule (gaussian_model/gaussian_utils.py:223-257)."""
import pytest

from nlos_gaussian_renderer_b200 import fit


def test_expon_lr_endpoints_and_delay():
    assert fit.expon_lr(0, 1.6e-4, 1.6e-6, max_steps=100) == pytest.approx(1.6e-4)
    assert fit.expon_lr(100, 1.6e-4, 1.6e-6, max_steps=100) == pytest.approx(1.6e-6)
    assert fit.expon_lr(50, 1.6e-4, 1.6e-6, max_steps=100) == pytest.approx(1.6e-5)
    assert fit.expon_lr(0, 1e-2, 1e-3, lr_delay_steps=10, lr_delay_mult=0.01, max_steps=100) == pytest.approx(1e-4)
    assert fit.expon_lr(-1, 1e-2, 1e-3) == 0.0 and fit.expon_lr(5, 0.0, 0.0) == 0.0
