"""The C++ host front end (raytracingmin_b200/host/): scene loader against what the
reference's own parser produced (golden), JSON round trip, image writers decoded with
PIL, and the CLI's reference-compatible flag handling.  No GPU needed: rendering itself
is covered by the gpu-marked test at the bottom."""
import ctypes as C
import json
import os
import sys
import subprocess

import numpy as np
import pytest

from conftest import SCENE_NAMES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "raytracingmin_b200", "host")
CLI = os.path.join(HOST, "rtm_b200")


@pytest.fixture(scope="module")
def hostlib():
    if not os.path.exists(os.path.join(HOST, "librtm_host.so")) or not os.path.exists(CLI):
        subprocess.check_call(["make", "-C", HOST], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HOST, "librtm_host.so"))
    lib.rtmh_last_error.restype = C.c_char_p
    lib.rtmh_load_scene.argtypes = [C.c_char_p, C.c_int] + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 4
    lib.rtmh_roundtrip_json.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    lib.rtmh_write_bmp.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    lib.rtmh_write_jpg.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    return lib


def _load(lib, path, head):
    dims = (C.c_int * 4)()
    cam = np.empty(9)
    fov = C.c_float()
    n = lib.rtmh_load_scene(path.encode(), int(head), dims, cam.ctypes.data, C.byref(fov), 0, None, None, None, None)
    if n < 0:
        raise ValueError(lib.rtmh_last_error().decode())
    pos, col, emi, rad = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 3)), np.empty(n, np.float32)
    lib.rtmh_load_scene(path.encode(), int(head), dims, cam.ctypes.data, C.byref(fov), n, pos.ctypes.data, rad.ctypes.data,
                        col.ctypes.data, emi.ctypes.data)
    return list(dims), cam, fov.value, pos, rad, col, emi


@pytest.mark.parametrize("level", ["head", "abc"])
@pytest.mark.parametrize("name", SCENE_NAMES)
def test_cpp_loader_matches_reference_parser(hostlib, golden, scene_path, name, level):
    dims, cam, fov, pos, rad, col, emi = _load(hostlib, scene_path(name), level == "head")
    k = f"{name}.{level}."
    assert dims == list(golden[k + "parsed_dims"])
    assert np.array_equal(cam, golden[k + "parsed_cam9"]) and np.float32(fov) == golden[k + "parsed_fov"]
    assert np.array_equal(pos, golden[k + "parsed_pos"]) and np.array_equal(rad, golden[k + "parsed_rad"])
    assert np.array_equal(col, golden[k + "parsed_col"]) and np.array_equal(emi, golden[k + "parsed_emi"])


def test_cpp_loader_superset_policy(hostlib, tmp_path):
    """legacy keys, missing objectType, `{}` objects, unknown keys"""
    d = {"00 width": 8, "00 height": 4, "00 sample": 7, "comment": "x",
         "01 camera": {"origin": [0, 0, -10], "target": [0, 0, 0], "upVec": [0, 1, 0], "fov": 90},
         "02 scene": {"00 object": [{}, {"00 position": [1, 2, 3], "01 size": 0.1, "comment": "c",
                                         "02 material": {"color": [0.5, 0.25, 0.125], "emission": [0, 0, 0]}}]}}
    p = tmp_path / "legacy.json"
    p.write_text(json.dumps(d))
    dims, cam, fov, pos, rad, col, emi = _load(hostlib, str(p), False)
    assert dims == [8, 4, 7, 1] and fov == 90.0 and len(rad) == 1
    assert rad[0] == np.float32(0.1) and pos.tolist() == [[1, 2, 3]]
    assert _load(hostlib, str(p), True)[3].tolist() == [[3, 0, 0]]            # F2 in --compat=head
    d["02 scene"]["00 object"][1]["00 objectType"] = 3
    p.write_text(json.dumps(d))
    with pytest.raises(ValueError, match="neither a sphere"):
        _load(hostlib, str(p), False)
    p.write_text('{"00 width": 8,')
    with pytest.raises(ValueError, match="json parse error"):
        _load(hostlib, str(p), False)
    del d["01 camera"]["fov"]
    d["02 scene"]["00 object"][1]["00 objectType"] = 1
    p.write_text(json.dumps(d))
    with pytest.raises(ValueError, match="fov"):
        _load(hostlib, str(p), False)


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_json_round_trip(hostlib, scene_path, tmp_path, name):
    buf = C.create_string_buffer(1 << 20)
    n = hostlib.rtmh_roundtrip_json(scene_path(name).encode(), buf, len(buf))
    assert n > 0
    p = tmp_path / "rt.json"
    p.write_bytes(buf.value)
    a, b = _load(hostlib, scene_path(name), False), _load(hostlib, str(p), False)
    assert a[0] == b[0] and a[2] == b[2]
    for x, y in zip(a[3:], b[3:]):
        assert np.array_equal(x, y)
    # and the Python twin reads the C++ writer's output identically
    import raytracingmin_b200 as rtm
    s1, s2 = rtm.load_scene(scene_path(name)), rtm.load_scene(str(p))
    assert np.array_equal(s1.position, s2.position) and np.array_equal(s1.radius, s2.radius)


def test_image_writers_decode(hostlib, tmp_path):
    from PIL import Image
    rng = np.random.default_rng(0)
    h, w = 37, 51                                        # not multiples of 8, odd row padding
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([xx * 255 // (w - 1), yy * 255 // (h - 1), (xx + yy) * 255 // (w + h - 2)], axis=2).astype(np.uint8)
    img[5:12, 7:20] = rng.integers(0, 256, (7, 13, 3), dtype=np.uint8)
    bmp, jpg = str(tmp_path / "o.bmp"), str(tmp_path / "o.jpg")
    assert hostlib.rtmh_write_bmp(bmp.encode(), img.ctypes.data, w, h) == 0
    assert hostlib.rtmh_write_jpg(jpg.encode(), img.ctypes.data, w, h, 60) == 0
    got = np.asarray(Image.open(bmp).convert("RGB"))
    assert np.array_equal(got, img)                      # BMP is lossless
    j = Image.open(jpg)
    assert j.format == "JPEG" and j.size == (w, h)
    dec = np.asarray(j.convert("RGB")).astype(int)
    smooth = np.ones((h, w), bool)
    smooth[3:14, 5:22] = False
    assert np.abs(dec - img.astype(int))[smooth].mean() < 4.0     # q60, 2x2 chroma subsampling, on a smooth gradient
    # same encoder settings as PIL at q60 give a similar file size (within 2x)
    ref = str(tmp_path / "ref.jpg")
    Image.fromarray(img).save(ref, quality=60, subsampling=2)
    assert 0.5 < os.path.getsize(jpg) / os.path.getsize(ref) < 2.0


@pytest.mark.parametrize("w,h,quality", [(51, 37, 60), (64, 48, 60), (17, 5, 60), (960, 504, 60), (33, 20, 95), (16, 16, 1)])
def test_image_writers_are_byte_identical_to_the_reference_stb(hostlib, tmp_path, w, h, quality):
    """north_star: "the same stb image output".  The reference writes result.jpg (quality 60) and result.bmp through
    stb_image_write v1.15 (src/Renderer.cpp:256-257; stb/stb_image_write.h:491-498, 1440-1560); the front end's own
    encoders must produce the same bytes as the stb object code the reference itself compiled (oracle/_ref)."""
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_lib as O
    if not O.ref_available("head") or not hasattr(O.ref_lib("head"), "refh_stb_write"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(w * 1000 + h)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([xx * 255 // max(1, w - 1), yy * 255 // max(1, h - 1), (xx + yy) * 255 // max(1, w + h - 2)], axis=2).astype(np.uint8)
    img[h // 4:h // 2, w // 4:w // 2] = rng.integers(0, 256, (h // 2 - h // 4, w // 2 - w // 4, 3), dtype=np.uint8)   # noise: long codes, 0xFF stuffing
    img[0, :, :] = 255
    img = np.ascontiguousarray(img)
    assert O.ref_lib("head").refh_stb_write(str(tmp_path / "ref").encode(), img.ctypes.data, w, h, quality) == 0
    assert hostlib.rtmh_write_bmp(str(tmp_path / "own.bmp").encode(), img.ctypes.data, w, h) == 0
    assert hostlib.rtmh_write_jpg(str(tmp_path / "own.jpg").encode(), img.ctypes.data, w, h, quality) == 0
    assert (tmp_path / "own.bmp").read_bytes() == (tmp_path / "ref.bmp").read_bytes()
    a, b = (tmp_path / "own.jpg").read_bytes(), (tmp_path / "ref.jpg").read_bytes()
    assert len(a) == len(b) and a == b, (len(a), len(b), next((i for i in range(min(len(a), len(b))) if a[i] != b[i]), None))


def test_cli_reference_flags(hostlib, tmp_path):
    r = subprocess.run([CLI, "-?"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0 and "-sampleJson" in r.stdout and "-json" in r.stdout
    r = subprocess.run([CLI, "-sampleJson"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0
    d = json.loads((tmp_path / "settingData.json").read_text())       # src/SettingData.cpp:14-24
    assert (d["00 width"], d["00 height"], d["00 samples"], d["00 superSamples"]) == (960, 540, 10, 4)
    assert d["01 camera"]["fov"] == 60 and d["01 camera"]["target"] == [0, 0, 1] and "02 scene" not in d


def test_cli_fails_loudly_without_gpu(hostlib, tmp_path, scene_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([CLI, "-json", scene_path("settingData")], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr
    assert not (tmp_path / "result.bmp").exists()


@pytest.mark.gpu
def test_cli_renders_result_files(hostlib, tmp_path, scene_path):
    """drop-in run: -json <file> -> result.jpg + result.bmp; pixels = the C ABI's u8 image"""
    from PIL import Image
    import raytracingmin_b200 as rtm
    hits = tmp_path / "hits.bin"
    r = subprocess.run([CLI, "-json", scene_path("cornellBox"), "--width=96", "--height=50", "--samples=4", "--superSamples=2",
                        "--stats", f"--dump-hits={hits}", f"--dump-pfm={tmp_path / 'linear.pfm'}"],
                       capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    bmp = np.asarray(Image.open(tmp_path / "result.bmp").convert("RGB"))
    assert Image.open(tmp_path / "result.jpg").size == (96, 50)
    sc = rtm.load_scene(scene_path("cornellBox")).with_dims(96, 50, 4, 2)
    with rtm.Renderer(sc) as rr:
        res = rr.render(want_hits=True)
        u8 = rr.read_u8()
    assert np.array_equal(bmp, u8)                                   # same seed, same image, bit for bit
    assert np.array_equal(np.fromfile(hits, np.int32).reshape(res.hits.shape), res.hits)
    with open(tmp_path / "linear.pfm", "rb") as f:                   # the linear image, before the 8-bit conversion
        assert f.readline() == b"PF\n" and f.readline() == b"96 50\n" and f.readline() == b"-1.0\n"
        pfm = np.frombuffer(f.read(), "<f4").reshape(50, 96, 3)[::-1]
    assert np.array_equal(pfm, res.image)
    st = json.loads(r.stdout.strip().splitlines()[-1])
    assert st["samples"] == 96 * 50 * 16 and st["rays"] == res.stats.rays


@pytest.mark.gpu
def test_cli_renders_the_plane_scene_with_both_estimators(hostlib, tmp_path):
    """extensions through the drop-in CLI: objectType 2 in the scene file, --estimator=nee"""
    from PIL import Image
    import raytracingmin_b200 as rtm
    scene = os.path.join(ROOT, "scenes", "cornellBoxPlanes.json")
    sc = rtm.load_scene(scene).with_dims(96, 50, 4, 2)
    for est in ("reference", "nee"):
        r = subprocess.run([CLI, "-json", scene, "--width=96", "--height=50", "--samples=4", "--superSamples=2", "--stats",
                            f"--estimator={est}"], capture_output=True, text=True, cwd=tmp_path)
        assert r.returncode == 0, r.stderr
        bmp = np.asarray(Image.open(tmp_path / "result.bmp").convert("RGB"))
        with rtm.Renderer(sc, estimator=est) as rr:
            rr.render()
            u8 = rr.read_u8()
        assert np.array_equal(bmp, u8), est
