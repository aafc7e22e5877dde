"""The FP32 path's model of the reference's Normalize() error against the reference's arithmetic (CPU),
and the device function against the model (GPU)."""
import numpy as np
import pytest

from acne_ref import model_eps, reference_eps


def test_model_matches_reference_arithmetic_distribution():
    rng = np.random.default_rng(7)
    ref = reference_eps(400000, rng)
    mod = model_eps(rng.integers(0, 2 ** 32, 400000, dtype=np.uint64).astype(np.uint32)).astype(np.float64)
    # the quantity that decides a self-hit is the sign and size of eps: compare the whole distribution
    qs = [0.01, 0.05, 0.1, 0.216, 0.25, 0.5, 0.75, 0.9, 0.95, 0.99]
    assert np.allclose(np.quantile(ref, qs), np.quantile(mod, qs), atol=2.5e-9), (np.quantile(ref, qs), np.quantile(mod, qs))
    assert abs((ref < 0).mean() - (mod < 0).mean()) < 0.01          # ~0.216: share of hit points inside the sphere
    assert abs(ref.mean() - mod.mean()) < 1e-9 and abs(ref.std() - mod.std()) < 1e-9


def test_reference_eps_depends_little_on_radius():
    rng = np.random.default_rng(8)
    a = reference_eps(200000, rng, 0.05, 0.3)
    b = reference_eps(200000, rng, 0.3, 2.0)
    assert abs((a < 0).mean() - (b < 0).mean()) < 0.02
    assert abs(a.std() - b.std()) < 3e-9


@pytest.mark.gpu
def test_device_eps_is_the_model():
    from raytracingmin_b200 import api
    rng = np.random.default_rng(9)
    bits = rng.integers(0, 2 ** 32, 100000, dtype=np.uint64).astype(np.uint32)
    bits[:4] = [0, 0xFFFFFFFF, 0x80808080, 0x01020304]
    dev, mod = api.acne_eps_device(bits).astype(np.float64), model_eps(bits).astype(np.float64)
    # nvcc contracts eta = S*c into the final FMA: agreement to 1 ulp of 6e-8 (the model's spread is 3.7e-8)
    assert np.abs(dev - mod).max() < 2e-14
    assert np.array_equal(dev < 0, mod < 0)
