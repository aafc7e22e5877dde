"""The C-ABI library: loads, exports every symbol include/rtm_b200.h declares, and
fails loudly (no CPU fallback) when there is no CUDA device.  No compute here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import raytracingmin_b200 as rtm
from raytracingmin_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "rtm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b((?:renderer|rtm_debug)_\w+)\s*\(", src)
    return sorted(set(names))


def test_header_symbols_are_exported():
    lib = rtm.load_library()
    names = _declared_functions()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rtm_b200.h but not exported"
    assert sorted(api.EXPORTS) == names, "api.EXPORTS out of sync with the header"


def test_abi_version_and_defaults():
    lib = rtm.load_library()
    assert lib.renderer_abi_version() == 4
    o = api._Options()
    lib.renderer_default_options(C.byref(o))
    assert (o.device, o.n_gpus, o.compat, o.precision, o.variant, o.accel) == (0, 1, api.COMPAT_ABC, api.PRECISION_F32,
                                                                                 api.VARIANT_MEGA, api.ACCEL_AUTO)
    assert o.seed == 0x5EED


def test_struct_layout_matches_header():
    """sizeof as the C compiler sees them (LP64): guards the ctypes mirrors."""
    assert C.sizeof(api._SceneDesc) == 16 + 72 + 8 + 32 + 8 + 48
    assert C.sizeof(api._Options) == 80
    assert C.sizeof(api._Stats) == 64


def test_invalid_arguments_are_rejected_before_touching_the_gpu():
    sc = rtm.random_scene(1, 3)
    for bad in (dict(width=0), dict(samples=0), dict(superSamples=-1)):
        with pytest.raises(rtm.RtmError) as e:
            rtm.Renderer(sc.with_dims(**bad))
        assert e.value.code == -1
    with pytest.raises(rtm.RtmError) as e:
        rtm.Renderer(sc, sample_begin=3, sample_end=2)
    assert e.value.code == -1


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(rtm.RtmError) as e:
        rtm.Renderer(rtm.random_scene(1, 3))
    assert e.value.code == -2
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under raytracingmin_b200/ may reference it."""
    pkg = os.path.join(ROOT, "raytracingmin_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or fn == "Makefile":
                txt = open(os.path.join(dp, fn), errors="replace").read()
                assert "rtm_oracle" not in txt and "oracle_lib" not in txt and "oracle/" not in txt, os.path.join(dp, fn)
