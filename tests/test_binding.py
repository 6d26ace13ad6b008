"""tests/test_binding.py -- the reference-side binding, RUN on the reference's own primitives.

oracle/_ref/libnnbvh_ref_binding.so holds the unmodified reference translation units, the driver, and
integration/pbrt/gpubvh.h (pbrt::GPUBVHAggregate, the class a maintainer adds as the "gpubvh" accelerator) linked
against libnnbvh.so.  The binding is constructed over the very std::vector<Primitive> -- SimplePrimitive over
Triangle and over Sphere -- that the reference's BVHAggregate of the same scene got, and every
ShapeIntersection it returns (tHit and the SurfaceInteraction: pi with its error interval, n, uv, dpdu/dpdv,
dndu/dndv, the shading frame, faceIndex, wo) is compared with BVHAggregate::Intersect's, bit for bit; likewise
IntersectP and Bounds.  This is the drop-in claim itself: same inputs, the reference's types on both sides.
"""
import numpy as np
import pytest

from conftest import rand_rays


@pytest.fixture(scope="module")
def binding():
    from oracle import pyoracle
    if not pyoracle.binding_available():
        pytest.skip("oracle/_ref/libnnbvh_ref_binding.so not built (no /root/reference on this machine)")
    return pyoracle.Binding(threads=1)


def _scene(rng, T, S):
    from tools.make_golden import random_spheres
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.03) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    sph = random_spheres(rng, S, scale=0.1, center_box=0.0)
    for k in range(S):
        m = np.eye(4)
        m[:3] = sph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        if k % 4 == 0:
            m[:3, 0] *= -1
        sph[k, :12] = m[:3].astype(np.float32).ravel()
        sph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    sph[1::3, 27] = rng.uniform(40, 350, len(sph[1::3])).astype(np.float32)       # phimax-clipped ones
    return tris, sph


@pytest.mark.gpu
@pytest.mark.parametrize("split", ["sah", "hlbvh"])
def test_binding_returns_the_reference_shape_intersections(binding, split):
    rng = np.random.default_rng(71)
    tris, sph = _scene(rng, 6000, 50)
    sc = binding.scene_mixed(tris, sph, 4, split)
    b = binding.create(sc, 4, split)
    try:
        N = 40_000
        rays = rand_rays(rng, N)
        aim = rng.integers(0, len(sph), N // 2)
        rays[:N // 2, 4:7] = sph[aim][:, [3, 7, 11]] + rng.normal(size=(N // 2, 3)).astype(np.float32) * 0.05 \
            - rays[:N // 2, :3]
        rays[:, 7] = rng.random(N)
        rays[N // 2::2, 3] = rng.random(len(rays[N // 2::2])) * 1.5          # finite tMax on a quarter
        whit, wt, w42 = binding.intersect_surface(sc, rays)                   # BVHAggregate::Intersect
        assert 0.2 * N < whit.sum() < 0.9 * N
        ghit, gt, g42 = binding.surface(b, rays, batch=True)                  # IntersectBatch + Resolve
        assert np.array_equal(ghit, whit)
        assert np.array_equal(gt.view(np.uint32)[whit], wt.view(np.uint32)[whit])
        bad = np.argwhere(g42.view(np.uint32)[whit] != w42.view(np.uint32)[whit])
        assert bad.size == 0, f"{len(bad)} words differ, first {bad[0]}"
        n1 = 1500                                                             # Primitive's single-ray surface
        shit, st, s42 = binding.surface(b, rays[:n1], batch=False)
        assert np.array_equal(shit, whit[:n1]) and np.array_equal(s42.view(np.uint32)[shit], w42.view(np.uint32)[:n1][shit])
        assert np.array_equal(binding.intersect_p_binding(b, rays[-n1:]),
                              binding.intersect_p(sc, rays[-n1:], parallel=False))
        nodes, _ = binding.export(sc, len(tris) + len(sph))
        assert np.array_equal(binding.bounds(b).view(np.uint32), nodes[0, :6])   # BVHAggregate::Bounds = root box
    finally:
        binding.destroy(b)


@pytest.mark.gpu
def test_wavefront_subclass_runs_on_the_reference_queues(binding):
    """pbrt::GPUBVHWavefrontAggregate (integration/pbrt/gpubvh_wavefront.h) called through the reference's
    WavefrontAggregate interface with the reference's own RayQueue / EscapedRayQueue / ShadowRayQueue /
    SOA<PixelSampleState>, beside CPUAggregate's loops (wavefront/aggregate.cpp:42-67) over the reference
    BVHAggregate on the same queues.  What the reference's consumers leave behind must be identical: which
    rays escaped, the ray EnqueueWorkAfterIntersection spawned for every hit (intr.SpawnRay: pi's error
    bounds and n), and the radiance RecordShadowRayResult added per pixel."""
    rng = np.random.default_rng(5)
    T, N = 8000, 50_000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.03) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    sc = binding.scene(tris, 4, "sah")
    rays = rand_rays(rng, N)
    rays[:, 7] = rng.random(N)
    want_kind, want_next = binding.wavefront_closest(sc, rays, use_binding=False)
    assert set(np.unique(want_kind)) == {0, 1} and 0.1 * N < (want_kind == 1).sum() < 0.6 * N
    kind, nxt = binding.wavefront_closest(sc, rays, use_binding=True)
    assert np.array_equal(kind, want_kind)
    assert np.array_equal(nxt.view(np.uint32)[kind == 1], want_next.view(np.uint32)[kind == 1])
    rays[N // 2:, 3] = rng.random(N - N // 2)
    want_L = binding.wavefront_shadow(sc, rays, use_binding=False)
    assert 0.05 * N < (want_L == 0).sum() < 0.6 * N
    assert np.array_equal(binding.wavefront_shadow(sc, rays, use_binding=True), want_L)


def test_reference_loops_over_reference_queues_are_consistent(binding):
    """The checker's half of the test above, without a GPU: CPUAggregate's loops over real queues agree with the
    reference's per-ray answers (escaped <=> no hit, the continued ray keeps d / time and starts on the surface,
    an unoccluded shadow ray adds Ld / 2)."""
    rng = np.random.default_rng(15)
    T, N = 3000, 20_000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.04) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    sc = binding.scene(tris, 4, "sah")
    rays = rand_rays(rng, N)
    rays[:, 7] = rng.random(N)
    kind, nxt = binding.wavefront_closest(sc, rays, use_binding=False)
    hit, _, s42 = binding.intersect_surface(sc, rays)
    assert np.array_equal(kind == 1, hit) and np.array_equal(kind == 0, ~hit) and hit.sum() > 2000
    assert np.array_equal(nxt[hit][:, 4:8], rays[hit][:, 4:8])
    p = (s42[:, 0:3] + s42[:, 3:6]) / 2
    assert np.abs(nxt[hit, :3] - p[hit]).max() < 1e-5
    rays[N // 2:, 3] = rng.random(N - N // 2)
    L = binding.wavefront_shadow(sc, rays, use_binding=False)
    occ = binding.intersect_p(sc, rays, parallel=False)
    assert np.array_equal(L, np.where(occ != 0, 0, (np.arange(N) % 7 + 1) / 2).astype(np.float32))
