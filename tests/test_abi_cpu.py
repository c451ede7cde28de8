"""CPU tests (-m "not gpu"): the C-ABI library loads, exports every symbol include/rtrt.h declares,
and its host-only entry points (config loader, argument checks) behave.  No compute calls."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(pkg):
    import __graft_entry__ as g
    g.build()
    return pkg._lib.load()


def test_header_symbols_exported(lib, pkg):
    hdr = open(os.path.join(ROOT, "include", "rtrt.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(rtrt_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/rtrt.h but not exported by librtrt.so"
    assert declared == set(pkg._lib.SYMBOLS), "ctypes binding table out of sync with include/rtrt.h"
    assert lib.rtrt_abi_version() == 1


def test_config_defaults(pkg, lib):
    cfg = pkg.config_defaults()     # toml::find_or fallbacks, src/configLoader.cpp:13-27
    assert (cfg.width, cfg.height) == (1920, 1080)
    assert cfg.useDynamicResolution == 1 and cfg.targetFps == 60.0
    assert (cfg.maxWidth, cfg.maxHeight, cfg.minWidth, cfg.minHeight) == (3840, 2160, 640, 480)
    assert cfg.loadCameraAtInit == 0 and cfg.inputMeshFileName == b""


def test_config_toml(pkg, lib):
    cfg = pkg.load_config(os.path.join(ROOT, "tests", "golden", "config.toml"))
    assert (cfg.width, cfg.height) == (1280, 720)
    assert cfg.inputMeshFileName == b"meshes/terrain # not a comment.bin"
    assert cfg.numInputTextureFileNames == 2
    assert bytes(cfg.inputTextureFileNames[0]).split(b"\0")[0] == b"a.png"
    assert bytes(cfg.inputTextureFileNames[1]).split(b"\0")[0] == b"b,c.png"
    assert cfg.loadCameraAtInit == 1
    assert cfg.inputCameraFileName == b"cam_in.bin" and cfg.cameraSaveFileName == b"cam_out.bin"
    assert cfg.useDynamicResolution == 0 and abs(cfg.targetFps - 48.5) < 1e-6
    assert (cfg.maxWidth, cfg.maxHeight, cfg.minWidth) == (2560, 1440, 320)
    assert cfg.minHeight == 480      # missing key -> reference default


def test_config_missing_file(pkg, lib):
    with pytest.raises(pkg.RtrtError) as e:
        pkg.load_config("/nonexistent/config.toml")
    assert e.value.code == -5


def test_null_context_is_an_error_not_a_crash(lib):
    assert lib.rtrt_build_bvh(None) == -1
    assert lib.rtrt_destroy(None) == -1
    assert lib.rtrt_get_stream(None) is None


def test_create_without_gpu_fails_loudly(pkg, lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.RtrtError) as e:
        pkg.RayTracer(64, 64).init()
    assert e.value.code == -2       # RTRT_ERR_CUDA: no device, no CPU fallback
