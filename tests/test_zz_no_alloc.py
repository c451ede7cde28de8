"""GPU test of rtrt_traverse_host_noalloc (csrc/traverse_noalloc.cu): the allocation-free variant of rtrt_traverse_host whose device
staging and visit counters live in a file-scope `static __device__` array (address via cudaGetSymbolAddress, once per device).

In a file of its own that pytest collects LAST: the new entry point is compared, through the C ABI,
with the existing one on two calls in a row — bit for bit, which is the tolerance of every traversal test in this suite — and the second
call's counters must equal the first's (the shared array keeps the last call's contents; what must start at zero is cleared per call).
Passes alone in a fresh process and after the other tests (it builds its own context and scene)."""
import numpy as np
import pytest

from helpers import assert_hits_equal

pytestmark = pytest.mark.gpu


def _rays(pkg, n, seed):
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, pkg.RAY_DTYPE)
    rays["orig"] = rng.uniform((-60, -5, -60), (60, 25, 60), size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    rays["dir"] = d / np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    return rays


def test_noalloc_traverse_matches_traverse_host_twice_in_a_row(pkg, orc):
    L = pkg._lib
    pos, nrm, idx = pkg.scenes.terrain(64, 80, seed=1)                    # 10 240 triangles, 10 batches
    cam = pkg.scenes.make_camera((0, 30, -60), 0.0, -0.4, 160, 160)
    batch_a = np.concatenate([pkg.scenes.primary_rays(cam, 160, 160), _rays(pkg, 20000, 3)])       # 45 600 rays
    batch_b = _rays(pkg, L.NOALLOC_MAX_RAYS, 4)                                                    # exactly the maximum
    with pkg.RayTracer(64, 64) as rt:
        rt.set_mesh(pos, idx, nrm); rt.build_bvh()
        want_a, want_b = rt.traverse(batch_a), rt.traverse(batch_b)        # the existing entry point (rtrt_traverse_host)
        want_any = rt.traverse(batch_a, any_hit=True)
        # two calls in a row on different batches, then the first batch again: stale contents of the shared array must not leak
        got_a, nv_a, lt_a = rt.traverse_noalloc(batch_a)
        got_b, nv_b, lt_b = rt.traverse_noalloc(batch_b)
        got_a2, nv_a2, lt_a2 = rt.traverse_noalloc(batch_a)
        got_any, _, _ = rt.traverse_noalloc(batch_a, any_hit=True)
        # a short batch right after a long one: entries behind n belong to the previous call and must not be returned or counted
        got_s, nv_s, lt_s = rt.traverse_noalloc(batch_a[:1000])
        # sizes beyond the static maximum are rejected with a clear error, nothing is launched
        with pytest.raises(pkg.RtrtError) as e:
            rt.traverse_noalloc(_rays(pkg, L.NOALLOC_MAX_RAYS + 1, 5))
        assert e.value.code == -3 and "RTRT_NOALLOC_MAX_RAYS" in str(e.value)
        # the wide layout is never used (it may have to be built = allocate): same hits with the option on
        rt.set_option("wide_bvh", 1)
        got_w, _, _ = rt.traverse_noalloc(batch_a)
        rt.set_option("wide_bvh", 0)
        empty, nv0, lt0 = rt.traverse_noalloc(batch_a[:0])
    assert_hits_equal(got_a, want_a)
    assert_hits_equal(got_b, want_b)
    assert_hits_equal(got_a2, want_a)
    assert_hits_equal(got_any, want_any)
    assert_hits_equal(got_s, want_a[:1000])
    assert_hits_equal(got_w, want_a)
    assert empty.shape[0] == 0 and (nv0, lt0) == (0, 0)
    assert (nv_a, lt_a) == (nv_a2, lt_a2), "visit counters of a repeated call differ: the static scratch was not cleared"
    assert nv_a > batch_a.shape[0] and lt_a > 0 and nv_b > 0 and (nv_b, lt_b) != (nv_a, lt_a)
    assert 0 < nv_s < nv_a and lt_s <= lt_a
    # and against the checker: the CPU oracle's traversal of the same rays
    ob = orc.build_bvh(pos, nrm, idx)
    assert_hits_equal(got_a, orc.traverse(ob, batch_a))
