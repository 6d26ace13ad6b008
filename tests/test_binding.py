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


def _walk_scene(rng, T=6000, N=20_000):
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.05) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    rays = rand_rays(rng, N)
    # shadow rays as the integrator makes them: d = pLight - o (unnormalised), tMax = 1 - ShadowEpsilon
    target = rng.random((N, 3)).astype(np.float32) * 1.5 - 0.25
    rays[:, 4:7] = target - rays[:, :3]
    rays[:, 3] = np.float32(0.9999)
    rays[:, 7] = rng.random(N)
    rays[::97, 4:7] = 0.0          # degenerate: the walk does not start (`while (ray.d != 0)`)
    seg6 = np.concatenate([rays[:, :3], target], axis=1).astype(np.float32)
    return tris, rays, seg6


def test_reference_transmittance_and_subsurface_loops_walk_several_surfaces():
    """CPU half of the test below: CPUAggregate::IntersectShadowTr / IntersectOneRandom (wavefront/aggregate.cpp:70-124)
    over the reference BVHAggregate.  With null materials every surface is an interface, so the walks cross several
    surfaces -- the multi-round case the batched binding has to reproduce -- and run on several host threads."""
    from oracle import pyoracle
    if not pyoracle.binding_available():
        pytest.skip("oracle/_ref/libnnbvh_ref_binding.so not built (no /root/reference on this machine)")
    b = pyoracle.Binding(threads=4)
    rng = np.random.default_rng(23)
    tris, rays, seg6 = _walk_scene(rng, 3000, 5000)
    sc = b.scene(tris, 4, "sah")
    L, seg, _ = b.wavefront_shadow_tr(sc, rays, use_binding=False)
    assert seg.max() >= 4 and (seg[::97] == 0).all() and (seg[1::97] >= 1).all()
    # nothing is opaque and there is no medium: every walk reaches the light with T_ray = 1 -> L = Ld / 2
    assert np.array_equal(L, ((np.arange(len(rays)) % 7 + 1) / 2).astype(np.float32))
    pdf, pn, _ = b.wavefront_one_random(sc, seg6, use_binding=False)
    assert 0.2 * len(pdf) < (pdf > 0).sum() and (pdf[pdf > 0] <= 1).all() and (pdf[pdf > 0] < 1).sum() > 100
    b.set_threads(1)


@pytest.mark.gpu
def test_batched_shadow_tr_and_one_random_match_the_reference_loops():
    """pbrt::GPUBVHWavefrontAggregate::IntersectShadowTr / IntersectOneRandom -- one GPU batch per ROUND of segments
    (integration/pbrt/gpubvh_wavefront.h) -- through the WavefrontAggregate interface on the reference's queues,
    with pbrt's ParallelFor on 4 host threads, against CPUAggregate's loops on the same queues: the same radiance in
    SOA<PixelSampleState>, the same reservoir probabilities and chosen SubsurfaceInteractions, and no more GPU
    batches than the longest walk has segments (round-1: one single-ray GPU batch per segment per ray, from
    unsynchronised worker threads)."""
    from oracle import pyoracle
    if not pyoracle.binding_available():
        pytest.skip("oracle/_ref/libnnbvh_ref_binding.so not built (no /root/reference on this machine)")
    b = pyoracle.Binding(threads=4)
    rng = np.random.default_rng(29)
    tris, rays, seg6 = _walk_scene(rng)
    sc = b.scene(tris, 4, "sah")
    want_L, seg, _ = b.wavefront_shadow_tr(sc, rays, use_binding=False)
    got_L, _, batches = b.wavefront_shadow_tr(sc, rays, use_binding=True)
    assert np.array_equal(got_L, want_L)
    assert seg.max() >= 4 and batches == seg.max()
    want_pdf, want_pn, _ = b.wavefront_one_random(sc, seg6, use_binding=False)
    got_pdf, got_pn, batches = b.wavefront_one_random(sc, seg6, use_binding=True)
    assert np.array_equal(got_pdf, want_pdf) and np.array_equal(got_pn.view(np.uint32), want_pn.view(np.uint32))
    assert (want_pdf > 0).sum() > 2000 and 2 <= batches <= 64
    # the single-ray surface used concurrently: Primitive::Intersect from 4 ParallelFor workers at once (the
    # engine is not re-entrant; the binding's single-ray calls take a lock)
    bb = b.create(sc, 4, "sah")
    try:
        sub = rays[:2000][(rays[:2000, 4:7] != 0).any(axis=1)]   # d = 0 "hits" with tHit = NaN on both sides; the NaN's
        whit, wt, w42 = b.intersect_surface(sc, sub)             # sign / payload is the host's vs the device's
        ghit, gt, g42 = b.surface(bb, sub, batch=False, parallel=True)
        assert np.array_equal(ghit, whit) and np.array_equal(gt.view(np.uint32)[whit], wt.view(np.uint32)[whit])
        bad = np.argwhere(g42.view(np.uint32)[whit] != w42.view(np.uint32)[whit])
        assert bad.size == 0, f"{len(bad)} words differ: rows {np.unique(bad[:, 0])[:8]} columns {np.unique(bad[:, 1])}"
    finally:
        b.destroy(bb)
        b.set_threads(1)


@pytest.mark.gpu
@pytest.mark.parametrize("split", ["sah", "middle"])
def test_binding_over_one_list_with_every_primitive_kind(binding, split):
    """pbrt::GPUBVHAggregate over the list the reference's general scene holds: SimplePrimitive over Triangle and
    Sphere, GeometricPrimitive with a constant alpha, TransformedPrimitive over per-object BVHAggregates (whose
    flattened trees go to the device as the reference built them).  Every ShapeIntersection -- for an instance the
    object-space interaction carried back by renderFromPrimitive, cpu/primitive.cpp:112-130 -- IntersectP and the
    wavefront subclass's IntersectClosest over the reference's queues must equal the reference's."""
    from test_general_scene import _rays, _scene as general_scene
    rng = np.random.default_rng(83)
    tris, alpha, sph, soups, inst24, inst_obj = general_scene(rng)
    binding.tag_materials(False)        # null materials: the queue harness takes EnqueueWorkAfterIntersection's !material branch
    try:
        sc = binding.scene_general(tris, alpha, sph, soups, inst24, inst_obj, 4, split)["handle"]
    finally:
        binding.tag_materials(True)
    b = binding.create(sc, 4, split)
    try:
        N = 60_000
        rays = _rays(rng, N, sph, inst24)
        rays[:, 7] = rng.random(N)
        whit, wt, w42 = binding.intersect_surface(sc, rays)                   # BVHAggregate::Intersect
        assert 0.2 * N < whit.sum() < 0.9 * N
        ghit, gt, g42 = binding.surface(b, rays, batch=True)                  # one batch + Resolve per hit
        assert np.array_equal(ghit, whit)
        assert np.array_equal(gt.view(np.uint32)[whit], wt.view(np.uint32)[whit])
        bad = np.argwhere(g42.view(np.uint32)[whit] != w42.view(np.uint32)[whit])
        assert bad.size == 0, f"{len(bad)} words differ, first {bad[0]}"
        n1 = 1500                                                             # Primitive's single-ray surface
        shit, st, s42 = binding.surface(b, rays[N // 4:N // 4 + n1], batch=False)      # rays aimed at instances
        assert np.array_equal(shit, whit[N // 4:N // 4 + n1])
        assert np.array_equal(s42.view(np.uint32)[shit], w42.view(np.uint32)[N // 4:N // 4 + n1][shit])
        assert np.array_equal(binding.intersect_p_binding(b, rays[-n1:]),
                              binding.intersect_p(sc, rays[-n1:], parallel=False))
        nodes, _ = binding.export(sc, len(tris) + len(sph) + len(inst24))
        assert np.array_equal(binding.bounds(b).view(np.uint32), nodes[0, :6])
        sub = rays[:20_000].copy()
        sub[:, 3] = np.inf                                                    # the queues carry no tMax
        want_kind, want_next = binding.wavefront_closest(sc, sub, use_binding=False, split=split)
        kind, nxt = binding.wavefront_closest(sc, sub, use_binding=True, split=split)
        assert np.array_equal(kind, want_kind) and (kind == 1).sum() > 4000
        assert np.array_equal(nxt.view(np.uint32)[kind == 1], want_next.view(np.uint32)[kind == 1])
    finally:
        binding.destroy(b)
