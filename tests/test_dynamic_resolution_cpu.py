"""SURVEY 8f N3 — the dynamic-resolution controller (RayTracer::UpdateFrame, src/kernel.cu:78-101): the C-ABI entry point is a pure
host function, so it is checked here (no GPU) against the oracle's float32 restatement and against the properties the reference's
"safe resolution" block guarantees."""
import numpy as np
import pytest


def test_controller_matches_oracle(pkg, orc):
    cfg = pkg.config_defaults()
    assert cfg.useDynamicResolution and cfg.targetFps == 60.0 and (cfg.minWidth, cfg.maxWidth) == (640, 3840)
    rng = np.random.default_rng(7)
    widths = [640, 656, 1280, 1920, 1936, 2560, 3824, 3840] + [int(16 * rng.integers(40, 241)) for _ in range(40)]
    dts = [4.0, 8.0, 15.0, 16.0, 16.1, 16.6667, 17.0, 17.3, 20.0, 33.3, 50.0, 100.0, 1000.0] + list(rng.uniform(2.0, 80.0, 60).astype(np.float32))
    for w in widths:
        for dt in dts:
            got = pkg.dynamic_resolution_step(cfg, w, (w // 16) * 9, float(dt))
            want = orc.dynamic_resolution_step(w, dt, cfg.targetFps, cfg.minWidth, cfg.maxWidth)
            assert got == want, (w, dt, got, want)
            assert got[0] % 16 == 0 and cfg.minWidth <= got[0] <= cfg.maxWidth and got[1] == got[0] // 16 * 9


def test_controller_dead_band_and_direction(pkg):
    cfg = pkg.config_defaults()
    # inside [1000/62, 1000/58] ms the size is held; slower frames shrink, faster frames grow (sqrt of the time ratio)
    assert pkg.dynamic_resolution_step(cfg, 1920, 1080, 16.7) == (1920, 1080)
    w_slow, _ = pkg.dynamic_resolution_step(cfg, 1920, 1080, 33.3)
    w_fast, _ = pkg.dynamic_resolution_step(cfg, 1920, 1080, 8.0)
    assert w_slow < 1920 < w_fast
    assert abs(w_slow - 1920 * np.sqrt(16.6667 / 33.3)) <= 16 and abs(w_fast - 1920 * np.sqrt(16.6667 / 8.0)) <= 16
    assert pkg.dynamic_resolution_step(cfg, 640, 360, 500.0)[0] == cfg.minWidth
    assert pkg.dynamic_resolution_step(cfg, 3840, 2160, 1.0)[0] == cfg.maxWidth


def test_controller_disabled_and_bad_arguments(pkg):
    cfg = pkg.config_defaults()
    cfg.useDynamicResolution = 0
    assert pkg.dynamic_resolution_step(cfg, 1000, 777, 100.0) == (1000, 777)
    cfg.useDynamicResolution = 1
    with pytest.raises(pkg.RtrtError):
        pkg.dynamic_resolution_step(cfg, 1920, 1080, 0.0)
    with pytest.raises(pkg.RtrtError):
        pkg.dynamic_resolution_step(cfg, 0, 0, 16.0)


def test_controller_converges_on_a_synthetic_load(pkg, orc):
    """Closed loop with frame time proportional to the pixel count (the path is per-pixel work): the controller settles inside its
    dead band at the width whose frame time is the target, like the reference's loop does."""
    cfg = pkg.config_defaults()
    ms_per_mpx = 12.0                                      # 1920x1080 would take 24.9 ms
    w, h = 3840, 2160
    for _ in range(40):
        w, h = pkg.dynamic_resolution_step(cfg, w, h, ms_per_mpx * w * h / 1e6)
    dt = ms_per_mpx * w * h / 1e6
    assert 1000.0 / 62 - 0.6 <= dt <= 1000.0 / 58 + 0.6, (w, h, dt)
