"""CPU tests of the multi-GPU host logic of csrc/multi.cu (rtrt_multi_plan_rows: a pure host function): the exchange plans of every
pair of ranks are consistent — each row a rank needs arrives from exactly one peer, nothing is sent that is not needed."""
import numpy as np
import pytest


@pytest.mark.parametrize("H,world,strip", [(1080, 2, 8), (1080, 8, 8), (2160, 4, 8), (256, 3, 8), (270, 5, 16), (64, 4, 8)])
def test_halo_and_frame_blocks_are_single_row_ranges(pkg, H, world, strip):
    """What the direct (unpacked) exchange of csrc/multi.cu relies on: the rows one rank sends another for a halo exchange or the frame gather are
    ONE range of consecutive rows (bands are contiguous), so they can travel image to image; the strip-interleaved G-buffer blocks are not."""
    par = pkg.parallel
    interleaved = False
    for r in range(world):
        for d in range(world):
            for kind, halos in ((1, (1, 3, 6, 12, 24, 32)), (2, (0,))):
                for halo in halos:
                    s = par.plan_rows(H, world, strip, r, d, kind, halo)
                    assert s.size == 0 or s[-1] - s[0] + 1 == s.size, (r, d, kind, halo)
            s = par.plan_rows(H, world, strip, r, d, 0)
            interleaved |= s.size > 0 and s[-1] - s[0] + 1 != s.size
    assert interleaved or H // strip <= world


@pytest.mark.parametrize("H,world,strip", [(1080, 2, 8), (1080, 8, 8), (2160, 4, 8), (256, 3, 8), (270, 5, 16), (64, 4, 8)])
def test_plans_cover_exactly_what_each_rank_needs(pkg, H, world, strip):
    par = pkg.parallel
    bands = par.band_bounds(H, world)
    assert bands[0][0] == 0 and bands[-1][1] == H and all(b % 16 == 0 for b, _ in bands) and all(e > b for b, e in bands)
    rows = np.arange(H)
    owner = (rows // strip) % world
    for d in range(world):
        b, e = bands[d]
        in_band = (rows >= b) & (rows < e)
        # kind 0: G-buffers on band +- 32, from the ranks that traced those strips
        need = (rows >= b - 32) & (rows < e + 32) & (owner != d)
        got = np.zeros(H, np.int32)
        for r in range(world):
            s = par.plan_rows(H, world, strip, r, d, 0)
            assert (owner[s] == r).all() and (np.diff(s) > 0).all()
            got[s] += 1
        assert np.array_equal(got, need.astype(np.int32)), (d, "gbuf")
        # kind 1: halo rows around the band from the bands that own them
        for halo in (1, 3, 6, 12, 24, 32):
            need = ((rows >= b - halo) & (rows < b)) | ((rows >= e) & (rows < e + halo))
            got = np.zeros(H, np.int32)
            for r in range(world):
                s = par.plan_rows(H, world, strip, r, d, 1, halo)
                rb, re = bands[r]
                assert ((s >= rb) & (s < re)).all()
                got[s] += 1
            assert np.array_equal(got, need.astype(np.int32)), (d, "halo", halo)
        # kind 2: the frame gather = every row outside the band, from its owner band
        got = np.zeros(H, np.int32)
        for r in range(world):
            got[par.plan_rows(H, world, strip, r, d, 2)] += 1
        assert np.array_equal(got, (~in_band).astype(np.int32)), (d, "frame")
    assert par.plan_rows(H, world, strip, 0, 0, 2).size == 0


def test_plan_rows_rejects_bad_arguments(pkg):
    with pytest.raises(ValueError):
        pkg.parallel.plan_rows(32, 4, 8, 0, 1, 0)          # fewer 16-row tiles than ranks
    with pytest.raises(ValueError):
        pkg.parallel.plan_rows(256, 2, 8, 2, 0, 0)         # rank out of range
