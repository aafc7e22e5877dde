"""Next-event estimation, CPU side: the oracle's definition (rtmo_render_nee) has the
same expectation as the reference's estimator (rtmo_render, L-ABC, itself pinned bit-for-bit by the
reference's objects) and a smaller variance.  PARITY UNPINNED by construction: the reference's own
light-sampling code (SurfaeSample, src/Renderer.cpp:119-198) is dead and unfinished."""
import numpy as np
import pytest

import oracle_lib as O
import raytracingmin_b200 as rtm


def _means(sc, n_seeds, fn):
    return np.array([fn(sc, seed).mean() for seed in range(100, 100 + n_seeds)])


@pytest.mark.parametrize("name,dims,tol", [("cornellBox", (32, 18, 256, 1), 0.004), ("settingData", (32, 18, 256, 1), 0.004),
                                           ("simple2", (24, 12, 2048, 1), 0.03)])
def test_nee_expectation_equals_the_reference_estimators(scene_path, name, dims, tol):
    sc = rtm.load_scene(scene_path(name), "abc").with_dims(*dims)
    plain = _means(sc, 6, lambda s, seed: O.oracle_render(s, O.LEVEL_ABC, seed)[0])
    nee = _means(sc, 6, lambda s, seed: O.oracle_render_nee(s, seed + 50)[0])
    se = np.sqrt(plain.var(ddof=1) / 6 + nee.var(ddof=1) / 6)
    assert abs(plain.mean() - nee.mean()) < max(4 * se, tol * plain.mean()), (plain.mean(), nee.mean(), se)


def test_nee_reduces_variance_on_the_cornell_box(scene_path):
    sc = rtm.load_scene(scene_path("cornellBox"), "abc").with_dims(48, 26, 64, 2)
    ref = O.oracle_render(sc.with_dims(samples=2048), O.LEVEL_ABC, 7)[0]
    plain = O.oracle_render(sc, O.LEVEL_ABC, 8)[0]
    nee, rays, samples, shadow = O.oracle_render_nee(sc, 9)
    mse_plain, mse_nee = ((plain - ref) ** 2).mean(), ((nee - ref) ** 2).mean()
    assert mse_nee < mse_plain / 3, (mse_plain, mse_nee)
    assert 0.5 < shadow / samples < rays / samples          # at most one shadow ray per continued vertex


def test_nee_without_listed_lights_is_the_reference_estimator():
    # only the walls emit (radius 1000 < r): nothing is listed, every draw sequence is the plain one except
    # for the three extra uniforms NEE would consume — there are none when the list is empty
    sc = rtm.random_scene(3, 4, 24, 14, 8, 1)
    emi = sc.emission.copy()
    emi[:] = 0.0
    emi[0] = [3.0, 3.0, 3.0]
    rad = sc.radius.copy()
    rad[0] = 1001.0
    from dataclasses import replace
    sc = replace(sc, emission=emi, radius=rad).normalized()
    a = O.oracle_render(sc, O.LEVEL_ABC, 5)[0]
    b = O.oracle_render_nee(sc, 5)[0]
    assert np.array_equal(a, b)
