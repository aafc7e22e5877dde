"""renderer_render_noalloc (include/rtm_b200.h): the FP32 pass with its scratch — sub-pixel records, the ACNE c' values, the
counters and the work cursor — in the library's fixed device array instead of the handle's cudaMalloc'd buffers.  No
counterpart in the reference (added for callers that may not allocate: graph capture, serving loops); what is checked is
that it renders EXACTLY what the existing entry point renders for the same (scene, seed) — the image is a pure function of
the two (test_gpu_parity.py compares same-seed renders with array_equal, so does this file) — on calls in a row that find
the array full of the previous call's data, and that it refuses what does not fit.

Collected last on purpose (a failure here must not hide another test).  Runs alone as well.
"""
import os

import numpy as np
import pytest

import raytracingmin_b200 as rtm

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _scene(name, W, H, S, ss):
    return rtm.load_scene(os.path.join(ROOT, "scenes", name), "abc").with_dims(W, H, S, ss)


def _noalloc(r, sc, stream=None, poison=True):
    import torch
    out = torch.empty((sc.height, sc.width, 4), dtype=torch.float32, device="cuda:0")
    if poison:
        out.fill_(float("nan"))   # the pass overwrites every pixel
    st = stream if stream is not None else torch.cuda.current_stream()
    with torch.cuda.stream(st):
        r.render_noalloc(out.data_ptr(), st.cuda_stream)
    st.synchronize()
    a = out.cpu().numpy()
    assert (a[..., 3] == 0).all()
    return a[..., :3]


CASES = {
    # brute-force megakernel, 4 sub-pixels per pixel: 230,400 records
    "cornell": ("cornellBox.head.json", 320, 180, 16, 2, {}),
    # the scene with very small spheres: the c' block of the scratch is in use as well
    "settingData": ("settingData.head.json", 256, 144, 8, 1, {}),
    # hierarchy kernel (10^4 spheres): its counters and guided work cursor live in the scratch too
    "synth10k": ("synth10k.json", 96, 54, 8, 2, {}),
    # brute-force scene pushed through the hierarchy kernel
    "simple2_bvh": ("simple2.head.json", 192, 108, 8, 1, {"accel": "bvh"}),
    # the extensions ride the same pass: PlaneObject squares, next-event estimation, and both behind the hierarchy / together
    "planes": ("cornellBoxPlanes.json", 160, 90, 8, 2, {}),
    "planes_bvh": ("cornellBoxPlanes.json", 160, 90, 8, 1, {"accel": "bvh"}),
    "nee": ("cornellBox.head.json", 160, 90, 8, 2, {"estimator": "nee"}),
    "planes_nee": ("cornellBoxPlanes.json", 160, 90, 8, 1, {"estimator": "nee"}),
}


@pytest.mark.parametrize("case", list(CASES))
def test_noalloc_pass_equals_the_handle_pass_twice_in_a_row(case):
    fname, W, H, S, ss, kw = CASES[case]
    sc = _scene(fname, W, H, S, ss)
    with rtm.Renderer(sc, seed=11, **kw) as r:
        want = r.render().image.copy()
        first = _noalloc(r, sc)
        second = _noalloc(r, sc)          # finds the array full of the first call's records and a spent work cursor
        assert np.array_equal(first, want), (case, float(np.abs(first - want).max()))
        assert np.array_equal(second, want), (case, float(np.abs(second - want).max()))
        # and the handle's own pass is untouched by it
        assert np.array_equal(r.render().image, want)


def test_noalloc_calls_of_different_handles_share_the_array_in_stream_order():
    """Two handles (different scenes, different record counts) alternate on ONE stream: each call must see only its own
    records although the array still holds the other scene's, and a side stream works like the current one."""
    import torch
    sa, sb = _scene("cornellBox.head.json", 240, 135, 8, 2), _scene("synth10k.json", 64, 36, 8, 1)
    with rtm.Renderer(sa, seed=5) as ra, rtm.Renderer(sb, seed=6) as rb:
        wa, wb = ra.render().image.copy(), rb.render().image.copy()
        side = torch.cuda.Stream()
        for st in (None, side):
            for _ in range(2):
                assert np.array_equal(_noalloc(ra, sa, st), wa)
                assert np.array_equal(_noalloc(rb, sb, st), wb)


def test_noalloc_call_leaves_the_free_device_memory_alone():
    import torch
    sc = _scene("cornellBox.head.json", 320, 180, 4, 2)
    with rtm.Renderer(sc, seed=1) as r:
        out = torch.empty((sc.height, sc.width, 4), dtype=torch.float32, device="cuda:0")
        r.render_noalloc(out.data_ptr(), torch.cuda.current_stream().cuda_stream)   # first use: module load, symbol lookup
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(0)
        for _ in range(4):
            r.render_noalloc(out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info(0)
        assert free1 >= free0, (free0, free1)


def test_noalloc_refuses_frames_beyond_the_fixed_array_and_other_paths():
    import torch
    out = torch.empty((1080, 1920, 4), dtype=torch.float32, device="cuda:0")
    ptr, st = out.data_ptr(), torch.cuda.current_stream().cuda_stream
    # 2,073,600 sub-pixels x 16 B > 16 MiB
    with rtm.Renderer(_scene("cornellBox.head.json", 1920, 1080, 1, 1)) as r:
        with pytest.raises(rtm.RtmError) as e:
            r.render_noalloc(ptr, st)
        assert e.value.code == -4 and "scratch" in str(e.value) and "2073600" in str(e.value) and "16777216" in str(e.value)
    # 1024 x 1024 = 1,048,576 > 1,048,560 by sixteen records; 1024 x 1023 fits
    with rtm.Renderer(_scene("cornellBox.head.json", 1024, 1024, 1, 1)) as r:
        with pytest.raises(rtm.RtmError):
            r.render_noalloc(ptr, st)
    sc = _scene("cornellBox.head.json", 1024, 1023, 1, 1)
    with rtm.Renderer(sc, seed=2) as r:
        want = r.render().image.copy()
        assert np.array_equal(_noalloc(r, sc), want)
    small = _scene("cornellBox.head.json", 64, 36, 2, 1)
    with rtm.Renderer(small, precision="f64") as r:
        with pytest.raises(rtm.RtmError) as e:
            r.render_noalloc(ptr, st)
        assert e.value.code == -4
    with rtm.Renderer(small, variant="wavefront") as r:
        with pytest.raises(rtm.RtmError):
            r.render_noalloc(ptr, st)
    with rtm.Renderer(small) as r:
        with pytest.raises(rtm.RtmError) as e:
            r.render_noalloc(0, st)
        assert e.value.code == -1
