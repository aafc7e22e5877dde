"""The CPU oracle (oracle/rtm_oracle.c) against the golden vectors generated from
the reference's OWN object code (tests/golden/make_golden.py) and, when
oracle/_ref is built, against that object code directly.  This is what pins the
oracle: hit maps, double images (same per-row mt19937 seeds) and
ray counts must agree bit for bit."""
import numpy as np
import pytest

import oracle_lib as O
import raytracingmin_b200 as rtm
from conftest import SCENE_NAMES

LEVELS = ["head", "abc"]


def _scene(scene_path, name, level):
    return rtm.load_scene(scene_path(name), compat=level)


@pytest.mark.parametrize("level", LEVELS)
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_loader_matches_reference_parser(golden, scene_path, name, level):
    """scene.py against what png::LoadData::from_json produced (incl. F2 at HEAD)."""
    sc = _scene(scene_path, name, level)
    k = f"{name}.{level}."
    assert np.array_equal(sc.position, golden[k + "parsed_pos"])
    assert np.array_equal(sc.radius, golden[k + "parsed_rad"])
    assert np.array_equal(sc.color, golden[k + "parsed_col"])
    assert np.array_equal(sc.emission, golden[k + "parsed_emi"])
    assert np.float32(sc.fov) == golden[k + "parsed_fov"]
    cam9 = np.concatenate([sc.cam_origin, sc.cam_target, sc.cam_up])
    assert np.array_equal(cam9, golden[k + "parsed_cam9"])
    assert [sc.width, sc.height, sc.samples, sc.superSamples] == list(golden[k + "parsed_dims"])


@pytest.mark.parametrize("level", LEVELS)
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_oracle_hits_small_bit_exact(golden, scene_path, name, level):
    sc = _scene(scene_path, name, level).with_dims(64, 34, 3, 2)
    h = O.oracle_primary_hits(sc)
    assert np.array_equal(h, golden[f"{name}.{level}.hits_small"].astype(np.int32))


@pytest.mark.parametrize("level", LEVELS)
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_oracle_hist_480(golden, scene_path, name, level):
    sc = _scene(scene_path, name, level).with_dims(480, 252, 1, 2)
    h = O.oracle_primary_hits(sc)
    hist = np.bincount(h.ravel() + 1, minlength=sc.n + 1)
    assert np.array_equal(hist, golden[f"{name}.{level}.hist_480"])


# SURVEY.md §8c known answers (L-ABC, 480x252, ss=2; first entry = misses)
SURVEY_HISTS = {
    "cornellBox": [0, 27554, 164126, 163999, 22149, 49126, 56886, 0],
    "settingData": [440298, 27554, 7520, 8468],
    "simple1": [252422, 0, 0, 0, 12, 231406],
    "simple2": [0, 0, 0, 0, 0, 483840],
}


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_survey_histograms(golden, name):
    assert list(golden[f"{name}.abc.hist_480"]) == SURVEY_HISTS[name]


def test_head_cornell_tie_rule(golden):
    """Under F2 objects 1-4 collapse to identical spheres; only the lowest index is reported."""
    assert list(golden["cornellBox.head.hist_480"]) == [0, 60264, 6188, 0, 0, 0, 208694, 208694]


@pytest.mark.parametrize("level", LEVELS)
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_oracle_image_and_rays_bit_exact(golden, scene_path, name, level):
    """Same per-row mt19937 seeds => the restatement reproduces the reference image exactly."""
    sc = _scene(scene_path, name, level).with_dims(64, 34, 3, 2)
    img, rays, samples = O.oracle_render(sc, O.LEVEL_HEAD if level == "head" else O.LEVEL_ABC, 1234)
    k = f"{name}.{level}."
    assert samples == 64 * 34 * 3 * 4
    assert rays == int(golden[k + "rays_small"])
    assert np.array_equal(img, golden[k + "image_small"])


def test_oracle_intersect_known_answers(golden, scene_path):
    sc = rtm.load_scene(scene_path("simple2"), "abc")
    cases = golden["intersect_cases"]
    assert len(cases) >= 11
    lib = O.oracle()
    import ctypes as C
    n_hit = 0
    for row in cases:
        obj = int(row[0])
        org, d = np.ascontiguousarray(row[1:4]), np.ascontiguousarray(row[4:7])
        hit_ref, t_ref, n_ref = bool(row[7]), row[8], row[9:12]
        t = C.c_double()
        nrm = np.zeros(3)
        c = np.ascontiguousarray(sc.position[obj])
        hit = lib.rtmo_intersect_sphere(c.ctypes.data, C.c_float(float(sc.radius[obj])), org.ctypes.data, d.ctypes.data,
                                        C.byref(t), nrm.ctypes.data)
        assert bool(hit) == hit_ref
        if hit_ref:
            n_hit += 1
            assert t.value == t_ref
            assert np.array_equal(nrm, n_ref)
    assert 3 <= n_hit < len(cases)


def test_head_exact_image_equals_head_render(scene_path):
    """rtmo_render_head_exact (no RNG) == the L-HEAD path-traced image whenever RR survives or not:
    every sample is the emission of the primary hit on both branches."""
    for name in SCENE_NAMES:
        sc = rtm.load_scene(scene_path(name), "head").with_dims(48, 26, 2, 2)
        a = O.oracle_render_head_exact(sc)
        b, _, _ = O.oracle_render(sc, O.LEVEL_HEAD, 77)
        assert np.array_equal(a, b)


def test_to_u8_truncates_without_gamma():
    v = np.array([0.0, 0.5, 0.999, 1.0, 7.0, 1.0 / 255, 254.9999 / 255])
    assert list(O.oracle_to_u8(v)) == [0, 127, 254, 255, 255, 1, 254]


# ---- against the reference's own objects, when they are built (this container) ----
needs_ref = pytest.mark.skipif(not (O.ref_available("head") and O.ref_available("abc")),
                               reason="oracle/_ref not built (no /root/reference on this machine)")


@needs_ref
@pytest.mark.parametrize("level", LEVELS)
@pytest.mark.parametrize("seed,n,ss,closed", [(11, 9, 3, True), (3, 1, 1, False), (29, 40, 2, True), (57, 120, 1, False),
                                              (71, 6, 4, True)])
def test_oracle_vs_reference_objects_random_scene(level, seed, n, ss, closed):
    """The C restatement against the reference's own compiled objects on scenes neither was written for: hit maps,
    ray counts and the double image bit for bit (same mt19937 stream per row)."""
    sc = rtm.random_scene(seed=seed, n=n, width=40, height=24, samples=2, superSamples=ss, closed=closed)
    rs = O.RefScene(level, sc)
    assert np.array_equal(rs.primary_hits(0), O.oracle_primary_hits(sc))
    img_ref, rays_ref, _ = rs.render(4321)
    img, rays, _ = O.oracle_render(sc, O.LEVEL_HEAD if level == "head" else O.LEVEL_ABC, 4321)
    assert rays == rays_ref
    assert np.array_equal(img, img_ref)
    rs.close()


@needs_ref
def test_reference_parser_on_shipped_scene(scene_path, golden):
    rs = O.RefScene("abc", json_path=scene_path("cornellBox"))
    assert (rs.width, rs.height, rs.samples, rs.superSamples, rs.n) == (960, 504, 25, 4, 7)
    rs.close()
