"""tests/test_wavefront.py -- the wavefront batch backend (include/nnbvh_wavefront.hpp).

CPU part: the host-side restatement of Triangle::InteractionFromIntersection
(/root/reference/src/pbrt/shapes.h:884-1010) against golden answers produced by the reference's own
function (tools/make_golden.py -> tests/golden/interaction.npz) and, where oracle/_ref is built,
against the reference live on fresh inputs.  Bit-exact: every one of the 39 words.

GPU part: GPUBVHWavefrontAggregate::IntersectClosest / IntersectShadow over SoA ray queues
(CPUAggregate::IntersectClosest / IntersectShadow, wavefront/aggregate.cpp:34-68): every ray reaches
the sink exactly once, as a miss or with the oracle's primitive / tHit and the reference's surface
geometry, for queues in host memory and in device memory.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, rand_rays

CASES = ("plain", "uv", "uv_n", "n_only", "uv_n_s", "reversed", "degenerate_uv", "zero_ns")


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


@pytest.fixture(scope="module")
def shim():
    subprocess.check_call(["sh", os.path.join(ROOT, "tests", "cpp", "build.sh")], stdout=subprocess.DEVNULL)
    L = ctypes.CDLL(os.path.join(ROOT, "tests", "cpp", "libwavefront_shim.so"))
    L.wf_last_error.restype = ctypes.c_char_p
    L.wf_create.restype = ctypes.c_void_p
    L.wf_create.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    L.wf_destroy.argtypes = [ctypes.c_void_p]
    L.wf_interaction.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + \
        [ctypes.c_void_p] * 3 + [ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_int64, ctypes.c_void_p]
    L.wf_sphere_interaction.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_int64, ctypes.c_void_p]
    L.wf_interaction_wo.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
    L.wf_intersect_closest.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + \
        [ctypes.c_void_p] * 5
    L.wf_intersect_shadow.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_void_p, ctypes.c_int]
    return L


def _interaction(L, c):
    f = lambda k: np.ascontiguousarray(c[k]) if c.get(k) is not None else None
    verts, idx, tri, b, wt = f("verts"), f("idx"), f("tri"), f("b"), f("wo_time")
    n, s, uv = f("n"), f("s"), f("uv")
    out = np.zeros((tri.shape[0], 39), dtype=np.uint32)
    rc = L.wf_interaction(_p(verts), verts.shape[0], _p(idx), idx.shape[0], _p(n), _p(s), _p(uv),
                          int(bool(c["reverse"])), _p(tri), _p(b), _p(wt), tri.shape[0], _p(out))
    assert rc == 0
    return out


NAMES39 = (["pi.low"] * 3 + ["pi.high"] * 3 + ["n"] * 3 + ["uv"] * 2 + ["dpdu"] * 3 + ["dpdv"] * 3 +
           ["dndu"] * 3 + ["dndv"] * 3 + ["shading.n"] * 3 + ["shading.dpdu"] * 3 + ["shading.dpdv"] * 3 +
           ["shading.dndu"] * 3 + ["shading.dndv"] * 3 + ["faceIndex"])


def _same(got, want, what):
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"{what}: {len(bad)} words differ, first at query {bad[0][0]} field {NAMES39[bad[0][1]]}"


@pytest.mark.parametrize("name", CASES)
def test_interaction_matches_reference_golden(shim, name):
    g = np.load(os.path.join(GOLDEN, "interaction.npz"))
    c = {k: (g[f"{name}_{k}"] if f"{name}_{k}" in g else None)
         for k in ("verts", "idx", "uv", "n", "s", "tri", "b", "wo_time", "reverse")}
    _same(_interaction(shim, c), g[name + "_want"], name)


def test_interaction_matches_reference_live(shim, reference):
    from tools.make_golden import interaction_cases
    for seed in (1, 2, 3):
        for name, c in interaction_cases(np.random.default_rng(seed)).items():
            want = reference.triangle_interaction(c["verts"], c["idx"], c["tri"], c["b"], c["wo_time"],
                                                  normals=c["n"], tangents=c["s"], uvs=c["uv"],
                                                  reverse_orientation=c["reverse"])
            _same(_interaction(shim, c), want, f"{name} seed {seed}")


NAMES42 = NAMES39 + ["wo"] * 3


def _sphere_interaction(L, sphere28, reverse, p_obj, rays):
    """The host restatement on the engine's sphere hit records: pObj + (wo = -d, time)."""
    n = rays.shape[0]
    wt = np.ascontiguousarray(np.concatenate([-rays[:, 4:7], rays[:, 7:8]], axis=1), dtype=np.float32)
    p_obj = np.ascontiguousarray(p_obj, dtype=np.float32)
    s = np.ascontiguousarray(sphere28, dtype=np.float32)
    out = np.zeros((n, 42), dtype=np.uint32)
    assert L.wf_sphere_interaction(_p(s), int(bool(reverse)), _p(p_obj), _p(wt), n, _p(out)) == 0
    return out


def _same42(got, want, what):
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"{what}: {len(bad)} words differ, first at ray {bad[0][0]} field {NAMES42[bad[0][1]]}"


def test_sphere_interaction_matches_reference_golden(shim):
    """Sphere::InteractionFromIntersection (shapes.h:238-280) + Transform::operator()(SurfaceInteraction)
    (util/transform.cpp:229-261): all 42 words, bit for bit, for mirrored / clipped / reversed spheres."""
    g = np.load(os.path.join(GOLDEN, "sphere_interaction.npz"))
    total = 0
    for i, s in enumerate(g["spheres"]):
        hit = g["hit"][i]
        got = _sphere_interaction(shim, s, g["reverse"][i], g["p_obj"][i], g["rays"][i])
        _same42(got[hit], g["surf"][i][hit], f"sphere {i}")
        total += int(hit.sum())
    assert total > 2000


def test_sphere_interaction_matches_reference_live(shim, reference, oracle):
    from tools.make_golden import random_spheres, sphere_rays
    rng = np.random.default_rng(77)
    spheres = random_spheres(rng, 12)
    spheres[::3, 0:12:4] *= -1          # first column of renderFromObject: a mirror
    spheres[1::2, 27] = rng.uniform(60, 350, len(spheres[1::2])).astype(np.float32)     # u = phi / phiMax
    for i, s in enumerate(spheres):
        m = np.eye(4)
        m[:3] = s[:12].reshape(3, 4).astype(np.float64)
        s[12:24] = np.linalg.inv(m)[:3].ravel()
        rays = sphere_rays(rng, np.repeat(s[None], 300, 0))
        hit, t_hit, want = reference.sphere_interaction(s, i % 2 == 1, rays)
        p_obj = np.zeros((300, 3), dtype=np.float32)
        for k in np.nonzero(hit)[0]:
            h, o4 = oracle.sphere(rays[k], s)       # the record the engine returns (tests/test_spheres.py)
            assert h and o4[:1].view(np.uint32)[0] == t_hit[k:k + 1].view(np.uint32)[0]
            p_obj[k] = o4[1:]
        got = _sphere_interaction(shim, s, i % 2 == 1, p_obj, rays)
        _same42(got[hit], want.view(np.uint32)[hit], f"sphere {i}")


def test_phimax_degrees_round_trip(shim):
    """A binding that only sees Sphere::phiMax (radians) gets back a `phimax` that the constructor maps to the
    same float (nnbvh::SpherePhiMaxDegrees)."""
    rng = np.random.default_rng(4)
    deg = np.concatenate([rng.uniform(0, 360, 200_000), [0, 90, 180, 270, 359.99, 360, 1e-3, 45.5]]).astype(np.float32)
    k = np.float32(np.float32(3.14159265358979323846) / np.float32(180))
    rad = (k * deg).astype(np.float32)
    out = np.zeros_like(rad)
    shim.wf_phimax_degrees.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
    shim.wf_phimax_degrees.restype = None
    shim.wf_phimax_degrees(_p(rad), rad.size, _p(out))
    assert np.array_equal((k * np.clip(out, 0, 360)).astype(np.float32).view(np.uint32), rad.view(np.uint32))


def _normalize_f32(v):
    """Normalize(Vector3f) (vecmath.h:948-961) in float32 steps."""
    v = np.ascontiguousarray(v, dtype=np.float32)
    ls = v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2]
    return v / np.sqrt(ls)[:, None]


def test_interaction_stores_normalized_wo(shim):
    """Interaction(pi, n, uv, wo, time) keeps Normalize(wo) (interaction.h:30-32)."""
    g = np.load(os.path.join(GOLDEN, "sphere_interaction.npz"))
    wo = np.ascontiguousarray(g["wo"])
    v9 = np.array([0, 0, 0, 1, 0, 0, 0, 1, 0], dtype=np.float32)
    out = np.zeros_like(wo)
    shim.wf_interaction_wo(_p(v9), _p(wo), wo.shape[0], _p(out))
    assert np.array_equal(out.view(np.uint32), g["wo_stored"].view(np.uint32))
    assert np.array_equal(_normalize_f32(wo).view(np.uint32), g["wo_stored"].view(np.uint32))


def test_backend_fails_loudly_without_gpu(shim):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    verts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32)
    idx = np.array([[0, 1, 2]], dtype=np.int32)
    h = shim.wf_create(_p(verts), 3, _p(idx), 1, None, None, 4, 0, 1)
    assert not h and b"no CPU fallback" in shim.wf_last_error()


def _mesh(rng, V=4000, T=12000):
    """An indexed mesh of small triangles (shared vertices) with uv and normals."""
    verts = rng.random((V, 3)).astype(np.float32)
    a = rng.integers(0, V, T)
    # neighbours in a spatial sort keep triangles small
    order = np.lexsort((verts[:, 2], verts[:, 1], (verts[:, 0] * 16).astype(int)))
    inv = np.empty(V, dtype=np.int64)
    inv[order] = np.arange(V)
    b = order[np.clip(inv[a] + rng.integers(1, 6, T), 0, V - 1)]
    c = order[np.clip(inv[a] - rng.integers(1, 6, T), 0, V - 1)]
    idx = np.stack([a, b, c], axis=1).astype(np.int32)
    idx = idx[(idx[:, 0] != idx[:, 1]) & (idx[:, 1] != idx[:, 2]) & (idx[:, 0] != idx[:, 2])]
    uv = rng.random((V, 2)).astype(np.float32)
    n = rng.normal(size=(V, 3)).astype(np.float32)
    n = (n / np.linalg.norm(n, axis=1, keepdims=True)).astype(np.float32) + np.float32(0)
    return verts, np.ascontiguousarray(idx), uv, n


def _plane(addr, n):
    return np.ctypeslib.as_array((ctypes.c_float * n).from_address(addr))


def test_queue_views_are_the_reference_queue_planes(reference):
    """RayQueueView / ShadowRayQueueView name the arrays of the reference's OWN queue types: a real
    pbrt::RayQueue (WorkQueue<RayWorkItem>, workitems.h:330-343) and ShadowRayQueue (:165-172, :386), with
    SOA<> generated by the reference's soac, filled through PushCameraRay / Push.  Plane k of the view holds
    component k of ray i at index i -- nothing to gather or repack."""
    rng = np.random.default_rng(12)
    N = 5000
    rays = rand_rays(rng, N)
    rays[:, 7] = rng.random(N)
    rays[:, 3] = rng.random(N)
    q, planes, size = reference.ray_queue(rays, max_queue_size=N + 100)
    try:
        assert size == N and len(set(planes)) == 7
        for k, col in enumerate((0, 1, 2, 4, 5, 6, 7)):            # ox oy oz dx dy dz time
            assert np.array_equal(_plane(planes[k], N).view(np.uint32), rays[:, col].view(np.uint32)), k
        for i in (0, 1, N // 2, N - 1):                            # what (*rayQ)[i].ray gives the reference's loop
            got = reference.ray_queue_get(q, i)
            assert np.array_equal(got[[0, 1, 2, 4, 5, 6, 7]], rays[i, [0, 1, 2, 4, 5, 6, 7]])
    finally:
        reference.ray_queue_destroy(q)
    q, planes, size = reference.shadow_queue(rays)
    try:
        assert size == N and len(set(planes)) == 8
        for k, col in enumerate((0, 1, 2, 4, 5, 6, 7, 3)):         # ... time tMax
            assert np.array_equal(_plane(planes[k], N).view(np.uint32), rays[:, col].view(np.uint32)), k
    finally:
        reference.shadow_queue_destroy(q)


@pytest.mark.gpu
@pytest.mark.parametrize("where", ["host", "device", "reference_queue"])
def test_wavefront_queues_closest_and_shadow(shim, oracle, reference, where):
    import torch
    rng = np.random.default_rng(34)
    verts, idx, uv, nrm = _mesh(rng)
    T = idx.shape[0]
    h = shim.wf_create(_p(verts), verts.shape[0], _p(idx), T, _p(nrm), _p(uv), 4, 0, 4)
    assert h, shim.wf_last_error()
    try:
        N = 200_000
        rays = rand_rays(rng, N)
        rays[:, 7] = rng.random(N)                     # ray.time travels to the interaction
        rays[N // 2:, 3] = rng.random(N - N // 2)      # shadow tMax
        planes_np = [np.ascontiguousarray(rays[:, k]) for k in (0, 1, 2, 4, 5, 6, 7, 3)]
        ref_q = None
        if where == "device":
            keep = [torch.from_numpy(p).cuda() for p in planes_np]
            ptrs = [t.data_ptr() for t in keep]
        elif where == "reference_queue":   # the memory of the reference's own RayQueue / ShadowRayQueue
            ref_q = (reference.ray_queue(rays), reference.shadow_queue(rays))
            ptrs = ref_q[0][1] + [None]
        else:
            ptrs = [p.ctypes.data for p in planes_np]
        arr = (ctypes.c_void_p * 8)(*ptrs)
        arr_s = (ctypes.c_void_p * 8)(*ref_q[1][1]) if ref_q else arr

        kind = np.full(N, -1, dtype=np.int8)
        prim = np.zeros(N, dtype=np.uint32)
        t_hit = np.zeros(N, dtype=np.float32)
        surf = np.zeros((N, 39), dtype=np.uint32)
        wo_t = np.zeros((N, 4), dtype=np.float32)
        # a queue larger than maxRays is the caller's error (queues are allocated for maxQueueSize)
        assert shim.wf_intersect_closest(h, N - 1, arr, N, _p(kind), _p(prim), _p(t_hit), _p(surf), _p(wo_t)) < 0
        calls = shim.wf_intersect_closest(h, N, arr, N, _p(kind), _p(prim), _p(t_hit), _p(surf), _p(wo_t))
        assert calls == N, shim.wf_last_error()

        # the same rays through the oracle on the tree the reference builds (tMax = inf for closest)
        tris9 = verts[idx].reshape(T, 9)
        sc = reference.scene(tris9, 4, "sah")
        nodes, order = reference.export(sc, T)
        r_inf = rays.copy()
        r_inf[:, 3] = np.inf
        want = oracle.intersect(nodes, tris9[order], r_inf)
        miss = want[:, 0] == 0xFFFFFFFF
        assert np.array_equal(kind == 0, miss) and np.array_equal(kind == 1, ~miss)
        hit = np.nonzero(~miss)[0]
        assert len(hit) > N // 20
        assert np.array_equal(prim[hit], order[want[hit, 0]].astype(np.uint32))      # caller's primitive index
        assert np.array_equal(t_hit[hit].view(np.uint32), want[hit, 1])
        assert np.array_equal(wo_t[hit, :3], _normalize_f32(-rays[hit, 4:7])) and np.array_equal(wo_t[hit, 3], rays[hit, 7])
        b = want[hit, 2:5].view(np.float32)
        ref_surf = reference.triangle_interaction(verts, idx, prim[hit].astype(np.int32), b, wo_t[hit],
                                                  normals=nrm, uvs=uv)
        _same(surf[hit], ref_surf, "EnqueueWorkAfterIntersection geometry")

        found = np.full(N, 7, dtype=np.uint8)
        for tr in (0, 1):
            assert shim.wf_intersect_shadow(h, N, arr_s, N, _p(found), tr) == N, shim.wf_last_error()
            assert np.array_equal(found, oracle.intersect_p(nodes, tris9[order], rays))
        # an empty queue is a no-op
        assert shim.wf_intersect_shadow(h, N, arr_s, 0, _p(found), 0) == 0
    finally:
        shim.wf_destroy(h)
        if ref_q:
            reference.ray_queue_destroy(ref_q[0][0])
            reference.shadow_queue_destroy(ref_q[1][0])


@pytest.mark.gpu
def test_wavefront_closest_hits_on_spheres(shim, oracle, reference):
    """A primitive list of triangles + `Shape "sphere"` entries behind the wavefront backend: sphere
    leaves reach EnqueueWorkAfterIntersection with Sphere::Intersect's SurfaceInteraction
    (shapes.h:130-139), triangles with Triangle::Intersect's, every word equal to the reference's."""
    from tools.make_golden import random_spheres
    rng = np.random.default_rng(91)
    verts, idx, uv, nrm = _mesh(rng, V=2000, T=5000)
    T, S = idx.shape[0], 60
    sph = random_spheres(rng, S, scale=0.1, center_box=0.0)
    for k in range(S):
        m = np.eye(4)
        m[:3] = sph[k, :12].reshape(3, 4)
        m[:3, 3] = rng.random(3)
        if k % 4 == 0:
            m[:3, 0] *= -1                                   # swaps handedness
        sph[k, :12] = m[:3].astype(np.float32).ravel()
        sph[k, 12:24] = np.linalg.inv(m)[:3].astype(np.float32).ravel()
    rev = (np.arange(S) % 3 == 0).astype(np.uint8)
    shim.wf_create_mixed.restype = ctypes.c_void_p
    shim.wf_create_mixed.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + \
        [ctypes.c_void_p] * 3 + [ctypes.c_int, ctypes.c_void_p] + [ctypes.c_int] * 3
    h = shim.wf_create_mixed(_p(verts), verts.shape[0], _p(idx), T, _p(nrm), _p(uv), _p(sph), S, _p(rev), 4, 0, 4)
    assert h, shim.wf_last_error()
    try:
        N = 100_000
        rays = rand_rays(rng, N)
        aim = rng.integers(0, S, N // 2)                     # half the rays aimed at a sphere
        tgt = sph[aim][:, [3, 7, 11]] + rng.normal(size=(N // 2, 3)).astype(np.float32) * 0.05
        rays[:N // 2, 4:7] = tgt - rays[:N // 2, :3]
        rays[:, 7] = rng.random(N)
        rays[:, 3] = np.inf
        planes_np = [np.ascontiguousarray(rays[:, k]) for k in (0, 1, 2, 4, 5, 6, 7)]
        arr = (ctypes.c_void_p * 7)(*[p.ctypes.data for p in planes_np])
        kind = np.full(N, -1, dtype=np.int8)
        prim = np.zeros(N, dtype=np.uint32)
        t_hit = np.zeros(N, dtype=np.float32)
        surf = np.zeros((N, 39), dtype=np.uint32)
        wo_t = np.zeros((N, 4), dtype=np.float32)
        assert shim.wf_intersect_closest(h, N, arr, N, _p(kind), _p(prim), _p(t_hit), _p(surf), _p(wo_t)) == N, \
            shim.wf_last_error()

        tris9 = verts[idx].reshape(T, 9)
        sc = reference.scene_mixed(tris9, sph, 4, "sah")
        _, order = reference.export(sc, T + S)
        want = reference.intersect(sc, rays, parallel=False)
        miss = want[:, 0] == 0xFFFFFFFF
        assert np.array_equal(kind == 0, miss) and np.array_equal(kind == 1, ~miss)
        hit = np.nonzero(~miss)[0]
        assert np.array_equal(prim[hit], order[want[hit, 0]].astype(np.uint32))
        assert np.array_equal(t_hit[hit].view(np.uint32), want[hit, 1])
        on_tri = hit[prim[hit] < T]
        on_sph = hit[prim[hit] >= T]
        assert len(on_tri) > 1000 and len(on_sph) > 1000
        ref_surf = reference.triangle_interaction(verts, idx, prim[on_tri].astype(np.int32),
                                                  want[on_tri, 2:5].view(np.float32), wo_t[on_tri], normals=nrm, uvs=uv)
        _same(surf[on_tri], ref_surf, "triangle hits")
        got42 = np.concatenate([surf, wo_t[:, :3].view(np.uint32)], axis=1)
        for k in np.unique(prim[on_sph] - T):
            sel = on_sph[prim[on_sph] - T == k]
            rhit, rt, r42 = reference.sphere_interaction(sph[k], rev[k], rays[sel])
            assert rhit.all() and np.array_equal(rt.view(np.uint32), t_hit[sel].view(np.uint32))
            _same42(got42[sel], r42.view(np.uint32), f"sphere {k}")
        assert np.array_equal(wo_t[hit, 3], rays[hit, 7])
    finally:
        shim.wf_destroy(h)


# ---- instances (TransformedPrimitive::Intersect, cpu/primitive.cpp:112-125) --------------------------------
def _instance_interactions(L, inst24, inst_object, objs, obj_orders, order, n_top, rays, hits, inner):
    """The host restatement on the engine's records of instance hits -> (ray indices, 42 words each)."""
    L.wf_instance_interaction.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3 + \
        [ctypes.c_int64, ctypes.c_void_p]
    on_inst = np.nonzero(inner != 0xFFFFFFFF)[0]
    k_of = order[hits[on_inst, 0]] - n_top
    got = np.zeros((len(on_inst), 42), dtype=np.uint32)
    for k in np.unique(k_of):
        m = k_of == k
        sel, j = on_inst[m], int(inst_object[k])
        tri = np.ascontiguousarray(obj_orders[j][inner[sel]], dtype=np.int32)
        b = np.ascontiguousarray(hits[sel, 2:5]).view(np.float32)
        d_time = np.ascontiguousarray(np.concatenate([rays[sel, 4:7], rays[sel, 7:8]], axis=1), dtype=np.float32)
        out = np.zeros((len(sel), 42), dtype=np.uint32)
        soup = np.ascontiguousarray(objs[j], dtype=np.float32)
        assert L.wf_instance_interaction(_p(np.ascontiguousarray(inst24[k])), _p(soup), len(soup), _p(tri), _p(b),
                                         _p(d_time), len(sel), _p(out)) == 0
        got[m] = out
    return on_inst, got


def test_instance_interaction_matches_reference_golden(shim):
    g = np.load(os.path.join(GOLDEN, "instanced_surface.npz"))
    J = int(g["n_objects"])
    objs, orders = [g[f"obj{j}"] for j in range(J)], [g[f"obj{j}_order"] for j in range(J)]
    idx, got = _instance_interactions(shim, g["inst24"], g["inst_object"], objs, orders, g["order"], len(g["top"]),
                                      g["rays"], g["hits"], g["inner"])
    assert len(idx) > 800
    _same42(got, g["surf"][idx], "instance hits")


def test_instance_interaction_matches_reference_live(shim, reference):
    from tools.make_golden import instanced_scene
    rng = np.random.default_rng(58)
    top, objs, inst, io = instanced_scene(rng, T=600, K=20, obj_sizes=(200, 1, 700))
    inst[::3, 1:12:4] *= -1                 # second column of renderFromPrimitive: mirrored instances
    for k in range(len(inst)):
        m = np.eye(4)
        m[:3] = inst[k, :12].reshape(3, 4).astype(np.float64)
        inst[k, 12:] = np.linalg.inv(m)[:3].ravel()
    sc = reference.scene_instanced(top, objs, inst, io, 4, "sah")
    rays = rand_rays(rng, 8000)
    rays[:, 3] = np.inf
    rays[:, 4:7] *= np.exp(rng.uniform(-1, 1, (8000, 1))).astype(np.float32)
    rays[:, 7] = rng.random(8000)
    hits, inner = reference.intersect_instanced(sc["handle"], rays)
    _, _, surf = reference.intersect_surface(sc["handle"], rays)
    idx, got = _instance_interactions(shim, inst, io, objs, sc["obj_orders"], sc["order"], len(top), rays, hits, inner)
    assert len(idx) > 500
    _same42(got, surf.view(np.uint32)[idx], "instance hits")


@pytest.mark.gpu
def test_wavefront_closest_hits_on_instances(shim, reference):
    """An instanced scene behind the wavefront backend: top-level triangles and TransformedPrimitive
    instances; every ray's kind / tHit / SurfaceInteraction (42 words) equals BVHAggregate::Intersect's."""
    from tools.make_golden import instanced_scene
    rng = np.random.default_rng(63)
    top, objs, inst, io = instanced_scene(rng, T=4000, K=60, obj_sizes=(1500, 1, 3000))
    inst[::4, 0:12:4] *= -1
    for k in range(len(inst)):
        m = np.eye(4)
        m[:3] = inst[k, :12].reshape(3, 4).astype(np.float64)
        m[:3, :3] *= 0.6
        inst[k, :12] = m[:3].ravel()
        inst[k, 12:] = np.linalg.inv(m)[:3].ravel()
    obj_all = np.ascontiguousarray(np.concatenate(objs), dtype=np.float32)
    start = np.concatenate([[0], np.cumsum([len(o) for o in objs])]).astype(np.int32)
    shim.wf_create_instanced.restype = ctypes.c_void_p
    shim.wf_create_instanced.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int] * 4
    top = np.ascontiguousarray(top, dtype=np.float32)
    h = shim.wf_create_instanced(_p(top), len(top), _p(obj_all), _p(start), len(objs), _p(inst), _p(io), len(inst), 4, 0, 4)
    assert h, shim.wf_last_error()
    try:
        N = 60_000
        rays = rand_rays(rng, N)
        rays[:, 3] = np.inf
        rays[:, 7] = rng.random(N)
        planes_np = [np.ascontiguousarray(rays[:, k]) for k in (0, 1, 2, 4, 5, 6, 7)]
        arr = (ctypes.c_void_p * 7)(*[p.ctypes.data for p in planes_np])
        kind = np.full(N, -1, dtype=np.int8)
        prim = np.zeros(N, dtype=np.uint32)
        t_hit = np.zeros(N, dtype=np.float32)
        surf = np.zeros((N, 39), dtype=np.uint32)
        wo_t = np.zeros((N, 4), dtype=np.float32)
        assert shim.wf_intersect_closest(h, N, arr, N, _p(kind), _p(prim), _p(t_hit), _p(surf), _p(wo_t)) == N, \
            shim.wf_last_error()
        sc = reference.scene_instanced(top, objs, inst, io, 4, "sah")
        whits, winner = reference.intersect_instanced(sc["handle"], rays)
        rhit, rt, r42 = reference.intersect_surface(sc["handle"], rays)
        assert np.array_equal(kind == 1, rhit) and np.array_equal(kind == 0, ~rhit)
        hit = np.nonzero(rhit)[0]
        assert (winner[hit] != 0xFFFFFFFF).sum() > 3000 and (winner[hit] == 0xFFFFFFFF).sum() > 3000
        assert np.array_equal(prim[hit], sc["order"][whits[hit, 0]].astype(np.uint32))
        assert np.array_equal(t_hit[hit].view(np.uint32), rt[hit].view(np.uint32))
        got42 = np.concatenate([surf, wo_t[:, :3].view(np.uint32)], axis=1)
        _same42(got42[hit], r42.view(np.uint32)[hit], "instanced scene")
    finally:
        shim.wf_destroy(h)
