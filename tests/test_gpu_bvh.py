"""Parity of the hierarchy (RTM_ACCEL_BVH) closest-hit path — config C5 of SURVEY.md §8d, the
10 000-sphere scene — against the reference's O(N) loop (src/Renderer.cpp:58-73).

  * primary-hit maps: BIT-EXACT against the oracle's brute force and the golden vectors
    generated from the reference's own objects (tests/golden/make_golden_synth10k.py);
  * closest hit on arbitrary rays (FP64): hit index and t bit for bit;
  * radiance: statistical against the reference objects' render of the same scene.
"""
import hashlib
import os

import numpy as np
import pytest

import oracle_lib as O
import raytracingmin_b200 as rtm

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def synth():
    return rtm.load_scene(os.path.join(ROOT, "scenes", "synth10k.json"), "abc")


@pytest.fixture(scope="module")
def golden10k():
    return np.load(os.path.join(ROOT, "tests", "golden", "synth10k_golden.npz"))


@pytest.mark.parametrize("seed,n,ss,closed", [(5, 1, 3, True), (7, 12, 2, True), (9, 60, 1, False), (11, 200, 2, True),
                                              (13, 700, 1, False), (15, 3, 1, False)])
def test_bvh_hits_equal_brute_force_random_scenes(seed, n, ss, closed):
    sc = rtm.random_scene(seed, n, width=61, height=37, samples=1, superSamples=ss, closed=closed)
    want = O.oracle_primary_hits(sc)
    for precision in ("f32", "f64"):
        with rtm.Renderer(sc, accel="bvh", precision=precision) as r:
            got = r.render(want_image=False, want_hits=True).hits
        assert np.array_equal(got, want), (precision, int((got != want).sum()))


@pytest.mark.parametrize("name", ["cornellBox", "simple1", "simple2", "settingData"])
@pytest.mark.parametrize("level", ["head", "abc"])
def test_bvh_forced_on_shipped_scenes(golden, scene_path, name, level):
    """n <= 7, all spheres similar or huge: degenerate trees must still give the reference's answer
    (at HEAD every sphere sits at (z,0,0), so exact ties decide: lowest index wins)."""
    sc = rtm.load_scene(scene_path(name), level).with_dims(64, 34, 3, 2)
    with rtm.Renderer(sc, compat=level, accel="bvh") as r:
        res = r.render(want_image=False, want_hits=True)
    assert np.array_equal(res.hits, golden[f"{name}.{level}.hits_small"].astype(np.int32))


@pytest.mark.parametrize("name", ["cornellBox", "simple1", "settingData"])
def test_bvh_forced_radiance_matches_brute_force(scene_path, name):
    """with accel=bvh on the Cornell box six of seven radii are 1e4, so "big = 8 x the median" classed none as big
    and the r = 1e4 walls went through the FP32 leaf test, whose direct form cancels at 1e8.  Big is now also an absolute
    radius (>= 256); the FP32 bounce radiance of the forced hierarchy must agree with the brute-force megakernel: same
    primary hits, rays/sample within 0.5 %, image means within 1 % (same RNG keys, different arithmetic of the leaf test)."""
    sc = rtm.load_scene(scene_path(name), "abc").with_dims(96, 54, 256, 1)
    with rtm.Renderer(sc, accel="brute", seed=5) as r:
        a = r.render(want_hits=True)
    with rtm.Renderer(sc, accel="bvh", seed=5) as r:
        b = r.render(want_hits=True)
    assert np.array_equal(a.hits, b.hits)
    ra, rb = a.stats.rays / a.stats.samples, b.stats.rays / b.stats.samples
    assert abs(ra - rb) < 0.005 * ra, (ra, rb)
    assert abs(float(a.image.mean()) - float(b.image.mean())) < 0.01 * float(a.image.mean()) + 1e-5


def test_more_big_spheres_than_the_hierarchy_path_holds_is_refused():
    sc = rtm.random_scene(9, 80, width=8, height=8, samples=1, superSamples=1, closed=False)
    sc.radius[:12] = 1000.0
    with pytest.raises(rtm.RtmError) as e:
        rtm.Renderer(sc.normalized(), accel="bvh")
    assert e.value.code == -4
    with rtm.Renderer(sc.normalized(), accel="bvh", precision="f64") as r:   # the exact walk takes them in its leaves
        assert r.render(want_image=False, want_hits=True).hits.shape == (8, 8, 1, 1)


def test_synth10k_hits_480_vs_reference_objects(synth, golden10k):
    sc = synth.with_dims(480, 270, 1, 1)
    with rtm.Renderer(sc) as r:                      # accel=auto -> hierarchy for n > 64
        res = r.render(want_image=False, want_hits=True)
    assert np.array_equal(res.hits[:, :, 0, 0], golden10k["hits_480"].astype(np.int32))
    assert res.stats.bvh_nodes > 0
    assert res.stats.sphere_tests < 300 * sc.n_sub   # the reference would run 10 000 per ray


def test_synth10k_hits_full_1080p(synth, golden10k):
    """BASELINE config 5 at its full size: 2.07 M primary rays, every index equal (checksum + histogram)."""
    sc = synth.with_dims(1920, 1080, 1, 1)
    with rtm.Renderer(sc) as r:
        h = r.render(want_image=False, want_hits=True).hits
    assert np.array_equal(np.bincount(h.ravel() + 1, minlength=sc.n + 1), golden10k["hist_1080"])
    sha = np.frombuffer(hashlib.sha256(np.ascontiguousarray(h, "<i4").tobytes()).digest(), np.uint8)
    assert np.array_equal(sha, golden10k["sha_1080"])


def test_synth10k_supersampled_crop_vs_oracle(synth):
    sc = synth.with_dims(96, 54, 1, 3)
    want = O.oracle_primary_hits(sc)
    with rtm.Renderer(sc) as r:
        got = r.render(want_image=False, want_hits=True).hits
    assert np.array_equal(got, want)


def test_closest_hit_f64_bvh_random_rays_bit_exact(synth):
    rng = np.random.default_rng(8)
    n = 1500
    org = rng.uniform(-9, 9, (n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.sqrt((d * d).sum(axis=1).astype(np.float32)).astype(np.float64)[:, None]    # Normalize(), Ray.h:67-72
    d[0] = np.nan              # degenerate directions take the brute-force loop inside the kernel
    d[1] = 0.0
    d[2] *= 3.0
    d[3] = [1.0, 0.0, 0.0]     # axis-parallel: 1/d = inf in two slabs
    d[4] = [0.0, -1.0, 0.0]
    hit_o, t_o, _ = O.oracle_closest_hit(synth, org, d)
    with rtm.Renderer(synth.with_dims(16, 9, 1, 1)) as r:
        hit, t = r.closest_hit(org, d, "f64")
    assert np.array_equal(hit, hit_o)
    m = hit_o >= 0
    assert np.array_equal(t[m], t_o[m])
    assert m.mean() > 0.9


def test_closest_hit_f32_bvh_agrees_with_f64_on_surface_rays(synth):
    rng = np.random.default_rng(2)
    n = 20000
    org0 = np.tile(synth.cam_origin, (n, 1))
    d0 = rng.normal(size=(n, 3))
    d0[:, 2] = np.abs(d0[:, 2]) + 0.2
    d0 /= np.linalg.norm(d0, axis=1, keepdims=True)
    with rtm.Renderer(synth.with_dims(16, 9, 1, 1)) as r:
        hit0, t0 = r.closest_hit(org0, d0, "f64")
        assert (hit0 >= 0).all()
        p = org0 + d0 * t0[:, None]
        nrm = p - synth.position[hit0]
        nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        nrm[np.einsum("ij,ij->i", nrm, d0) > 0] *= -1
        d1 = rng.normal(size=(n, 3))
        d1 /= np.linalg.norm(d1, axis=1, keepdims=True)
        d1[np.einsum("ij,ij->i", nrm, d1) < 0] *= -1
        d1 = d1 + 0.05 * nrm
        d1 /= np.linalg.norm(d1, axis=1, keepdims=True)
        h64, t64 = r.closest_hit(p, d1, "f64")
        h32, t32 = r.closest_hit(p, d1, "f32", origin_id=hit0)
    agree = (h64 == h32)
    assert agree.mean() > 0.995, agree.mean()          # disagreement only at silhouettes / seams
    both = agree & (h64 >= 0)
    assert np.median(np.abs(t64[both] - t32[both])) < 1e-4
    assert np.percentile(np.abs(t64[both] - t32[both]), 99) < 2e-2
    assert (t32[h32 >= 0] > 1e-3).all()


def _rmse(a, b):
    return np.sqrt(((a - b) ** 2).reshape(-1, 3).mean(axis=0))


def test_synth10k_radiance_vs_reference_objects(synth, golden10k):
    """48x27 at 512 spp from the reference's own objects vs the CUDA paths at 2048 spp: the difference
    must be explained by the reference render's own noise (estimated from the two CUDA precisions)."""
    ref = golden10k["image_48"]
    rps_ref = float(golden10k["rays_48"]) / float(golden10k["samples_48"])
    sc = synth.with_dims(48, 27, 2048, 1)
    imgs = {}
    for precision in ("f32", "f64"):
        with rtm.Renderer(sc, precision=precision, seed=11) as r:
            res = r.render()
        imgs[precision] = res.image.astype(np.float64)
        rps = res.stats.rays / res.stats.samples
        assert abs(rps - rps_ref) < 0.01 * rps_ref, (precision, rps, rps_ref)      # same estimator
        rel = abs(imgs[precision].mean() - ref.mean()) / ref.mean()
        assert rel < 0.03, (precision, rel)
    # the two device paths are 2048-spp renders of the same estimator: their mutual RMSE bounds the
    # noise of one of them; the 512-spp reference image is 2x noisier per pixel
    noise = _rmse(imgs["f32"], imgs["f64"]) / np.sqrt(2.0)
    for precision in ("f32", "f64"):
        rm = _rmse(imgs[precision], ref)
        assert (rm <= 4.0 * noise * np.sqrt(1.0 + 4.0)).all(), (precision, rm, noise)


def test_bvh_image_matches_brute_force_image_statistically():
    """Same scene through both closest-hit searches (n = 200 fits the brute-force kernel)."""
    sc = rtm.random_scene(21, 200, width=64, height=36, samples=256, superSamples=1)
    with rtm.Renderer(sc, accel="brute", seed=3) as r:
        a = r.render()
    with rtm.Renderer(sc, accel="bvh", seed=4) as r:
        b = r.render()
    ia, ib = a.image.astype(np.float64), b.image.astype(np.float64)
    assert abs(ia.mean() - ib.mean()) < 0.02 * ia.mean()
    ra, rb = a.stats.rays / a.stats.samples, b.stats.rays / b.stats.samples
    assert abs(ra - rb) < 0.01 * ra
    assert b.stats.sphere_tests < a.stats.sphere_tests / 3


def test_bvh_deterministic_and_slices_add_up(synth):
    sc = synth.with_dims(64, 36, 8, 2)
    with rtm.Renderer(sc, seed=1) as r:
        full = r.render()
        again = r.render()
    assert np.array_equal(full.image, again.image)
    parts, rays = np.zeros_like(full.image, dtype=np.float64), 0
    for g in range(4):
        plan = rtm.plan_partition(sc.samples, sc.superSamples, g, 4)
        with rtm.Renderer(sc, seed=1, **plan) as r:
            res = r.render()
        parts += res.image
        rays += res.stats.rays
    assert rays == full.stats.rays
    assert np.allclose(parts, full.image, rtol=2e-6, atol=1e-7)


def test_shared_and_local_stack_instantiations_render_the_same_image(synth):
    """radiance_bvh_kernel<..., SST>: the traversal stack in shared memory (host-selected for trees of depth + 2 <= RTM_BVH_SSTACK = 26)
    against the local-memory stack (rtm_options.debug_bvh_walk bit 0).  Same walk, same candidates, same RNG counters:
    bit-identical image and counters."""
    sc = synth.with_dims(64, 36, 8, 2)
    with rtm.Renderer(sc, seed=1) as r:
        here = r.render()
    with rtm.Renderer(sc, seed=1, debug_bvh_walk=1) as r:
        there = r.render()
    assert np.array_equal(there.image, here.image)
    assert there.stats.rays == here.stats.rays
    assert there.stats.sphere_tests == here.stats.sphere_tests
    assert there.stats.bvh_nodes == here.stats.bvh_nodes


def test_compact_half_precision_nodes_render_the_same_image(synth):
    """radiance_bvh_kernel<..., WK = 1>: 32-byte nodes with half-precision (centre, half-extent) child boxes (one L1 request
    per visit) against the 64-byte float nodes (bit 1).  The compact boxes contain the
    float ones, so the walk meets a superset of the candidates and keeps the same winner: same paths, bit-identical image
    and ray count; only the visit counters may differ (by the few percent the boxes are wider)."""
    sc = synth.with_dims(96, 54, 8, 2)
    with rtm.Renderer(sc, seed=4) as r:
        a = r.render()
    with rtm.Renderer(sc, seed=4, debug_bvh_walk=2) as r:
        b = r.render()
    assert np.array_equal(a.image, b.image)
    assert a.stats.rays == b.stats.rays and a.stats.rays_traced == b.stats.rays_traced
    assert b.stats.bvh_nodes <= a.stats.bvh_nodes <= 1.15 * b.stats.bvh_nodes
    assert b.stats.sphere_tests <= a.stats.sphere_tests <= 1.15 * b.stats.sphere_tests


@pytest.mark.parametrize("stack_bit", [0, 1])
def test_four_wide_compact_nodes_render_the_same_image_in_half_the_visits(synth, stack_bit):
    """radiance_bvh_kernel<..., WK = 2> (rtm_options.debug_bvh_walk bit 3): 64-byte nodes of up to four children, the
    collapse of the compact two-child tree (same boxes, same leaves).  Same winners, so the same paths and a bit-identical
    image; about half the node visits; a few percent more sphere tests (children beyond the nearest are taken in slot
    order, not by distance)."""
    sc = synth.with_dims(96, 54, 8, 2)
    with rtm.Renderer(sc, seed=4, debug_bvh_walk=8 | stack_bit) as r:
        a = r.render()
    with rtm.Renderer(sc, seed=4, debug_bvh_walk=stack_bit) as r:
        b = r.render()
    assert np.array_equal(a.image, b.image)
    assert a.stats.rays == b.stats.rays and a.stats.rays_traced == b.stats.rays_traced
    assert 0.4 * b.stats.bvh_nodes <= a.stats.bvh_nodes <= 0.62 * b.stats.bvh_nodes, (a.stats.bvh_nodes, b.stats.bvh_nodes)
    assert 0.98 * b.stats.sphere_tests <= a.stats.sphere_tests <= 1.12 * b.stats.sphere_tests, (a.stats.sphere_tests, b.stats.sphere_tests)


def test_breadth_first_and_treelet_numberings_render_the_same_image(synth):
    """rtm_options.debug_bvh_walk bit 2: the inner nodes numbered breadth-first at the top + 4-node treelets below instead of
    breadth-first throughout.  Only addresses change: identical image and identical counters."""
    sc = synth.with_dims(64, 36, 8, 2)
    for walk in (0, 8, 16):
        with rtm.Renderer(sc, seed=9, debug_bvh_walk=walk) as r:
            a = r.render()
        with rtm.Renderer(sc, seed=9, debug_bvh_walk=walk | 4) as r:
            b = r.render()
        assert np.array_equal(a.image, b.image)
        assert a.stats.bvh_nodes == b.stats.bvh_nodes and a.stats.sphere_tests == b.stats.sphere_tests


@pytest.mark.parametrize("which", ["synth10k", "synth10k_sliced", "cornell_forced", "random300", "tiny_frame"])
def test_ray_pool_kernel_renders_what_the_lane_owned_kernel_renders(synth, scene_path, which):
    """radiance_bvh_pool_kernel (rtm_options.debug_bvh_walk bit 4: 64 rays per warp parked in shared memory, phases that
    pick 32 rays in the same phase — a measured design alternative) against radiance_bvh_kernel (one ray per
    lane, the default).  The same device functions run per ray with the same RNG counters, so the image and all four
    counters are identical."""
    kw = {}
    if which == "synth10k":
        sc = synth.with_dims(128, 72, 8, 2)
    elif which == "synth10k_sliced":
        sc, kw = synth.with_dims(64, 36, 8, 2), dict(sub_begin=1, sub_end=3, sample_begin=2, sample_end=7)
    elif which == "cornell_forced":     # big spheres (the packed test in front of the walk), a hierarchy of three
        sc, kw = rtm.load_scene(scene_path("cornellBox"), "abc").with_dims(96, 54, 8, 2), dict(accel="bvh")
    elif which == "random300":
        sc = rtm.random_scene(11, 300, width=96, height=54, samples=8, superSamples=1, closed=True)
    else:                               # fewer sub-pixels than one block has slots
        sc = synth.with_dims(8, 5, 4, 1)
    with rtm.Renderer(sc, seed=21, debug_bvh_walk=16, **kw) as r:
        a = r.render()
    with rtm.Renderer(sc, seed=21, **kw) as r:
        b = r.render()
    assert np.array_equal(a.image, b.image), float(np.abs(a.image - b.image).max())
    for f in ("rays", "rays_traced", "sphere_tests", "bvh_nodes", "samples"):
        assert getattr(a.stats, f) == getattr(b.stats, f), f


# ---------------- wavefront variant (same estimator, state in global memory) ----------------
@pytest.mark.parametrize("which", ["cornell", "random200_brute", "random200_bvh", "synth10k"])
def test_wavefront_matches_megakernel(scene_path, synth, which):
    if which == "cornell":
        sc, kw = rtm.load_scene(scene_path("cornellBox"), "abc").with_dims(96, 50, 8, 2), {}
    elif which == "synth10k":
        sc, kw = synth.with_dims(96, 54, 16, 1), {}
    else:
        sc = rtm.random_scene(21, 200, width=64, height=36, samples=32, superSamples=1)
        kw = dict(accel="brute" if which.endswith("brute") else "bvh")
    with rtm.Renderer(sc, seed=9, **kw) as r:
        a = r.render(want_hits=True)
    with rtm.Renderer(sc, seed=9, variant="wavefront", **kw) as r:
        b = r.render(want_hits=True)
        b2 = r.render()
    assert np.array_equal(a.hits, b.hits)
    assert np.array_equal(b.image, b2.image)                       # deterministic
    assert b.stats.samples == a.stats.samples
    # same RNG keys and the same formulas: the paths coincide except where a rounding difference between
    # the two compilations flips a roulette/closest-hit decision
    assert abs(b.stats.rays - a.stats.rays) <= 2e-3 * a.stats.rays, (a.stats.rays, b.stats.rays)
    ia, ib = a.image.astype(np.float64), b.image.astype(np.float64)
    assert abs(ia.mean() - ib.mean()) <= 5e-3 * ia.mean()
    same = np.isclose(ia, ib, rtol=1e-4, atol=1e-6).all(axis=-1).mean()
    assert same > 0.9, same


def test_hierarchy_100k_spheres_hit_map_bit_exact_and_radiance_runs(gpu_lib):
    """size edge: 100 000 spheres (hierarchy depth 19, 59 k nodes) — exact hit map against the oracle's O(N) loop
    on a small image, FP32 and FP64 bounce loops agree on the mean"""
    rng = np.random.default_rng(1)
    n = 100000
    pos = rng.uniform(-50, 50, (n, 3))
    rad = (np.round(np.exp(rng.uniform(np.log(0.05), np.log(0.8), n)) * 1024) / 1024).astype(np.float32)
    emi = np.zeros((n, 3))
    emi[::50] = 6.0
    sc = rtm.Scene(width=64, height=36, samples=64, superSamples=1, cam_origin=np.array([0.0, 0.0, -60.0]), cam_target=np.zeros(3),
                   cam_up=np.array([0.0, 1.0, 0.0]), fov=float(np.float32(0.8)), position=pos, radius=rad,
                   color=rng.uniform(0.1, 0.9, (n, 3)), emission=emi).normalized()
    want = O.oracle_primary_hits(sc)
    out = {}
    for precision in ("f32", "f64"):
        with rtm.Renderer(sc, precision=precision) as r:
            res = r.render(want_hits=True)
        assert np.array_equal(res.hits, want), precision
        assert res.stats.bvh_nodes > 0 and np.isfinite(res.image).all()
        out[precision] = (float(res.image.astype(np.float64).mean()), res.stats.rays / res.stats.samples)
    assert (want >= 0).mean() > 0.5
    assert abs(out["f32"][0] - out["f64"][0]) < 0.05 * out["f64"][0] + 1e-4, out
    assert abs(out["f32"][1] - out["f64"][1]) < 0.02 * out["f64"][1], out
