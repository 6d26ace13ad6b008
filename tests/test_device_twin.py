"""tests/test_device_twin.py -- the device math SOURCE, run on the host.

tests/cpp/device_twin.cpp includes nn_bvh_b200/csrc/{exact_math,sphere,atan2_host,alpha_math}.cuh unchanged and
compiles them with g++ against tests/cpp/cuda_stub/cuda_runtime.h, where every CUDA intrinsic they use has an
IEEE-754 equivalent (_rn operations = plain operations under -ffp-contract=off, __fmaf_rn = std::fma, bit casts =
memcpy).  So the statements the GPU executes -- the watertight triangle test, the slab test, Sphere::BasicIntersect
in interval arithmetic, glibc's atan2f, Transform::ApplyInverse(Ray), the alpha test's hash / spawned origin /
re-trace -- run here on the reference's golden vectors and beside the checker, bit for bit, without a GPU.
What this cannot show is that the GPU's instructions are IEEE-exact; the -m gpu tests do.
"""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, rand_rays


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


@pytest.fixture(scope="module")
def twin():
    subprocess.check_call(["sh", os.path.join(ROOT, "tests", "cpp", "build.sh")], stdout=subprocess.DEVNULL)
    L = ctypes.CDLL(os.path.join(ROOT, "tests", "cpp", "libdevice_twin.so"))
    for f in (L.dt_triangle, L.dt_sphere, L.dt_spawn_origin):
        f.argtypes = [ctypes.c_void_p] * 3
    L.dt_box.argtypes = [ctypes.c_void_p] * 2
    L.dt_atan2.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
    L.dt_atan2.restype = None
    L.dt_hash_float.argtypes = [ctypes.c_void_p] * 2
    L.dt_hash_float.restype = ctypes.c_float
    L.dt_alpha_triangle.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_float, ctypes.c_void_p]
    L.dt_instance_ray.argtypes = [ctypes.c_void_p] * 3
    L.dt_instance_ray.restype = None
    return L


def _pairs(L, fn, rays, other, gold_hit, gold_out):
    rays, other = np.ascontiguousarray(rays), np.ascontiguousarray(other)
    n_hit = 0
    for i in range(len(rays)):
        out = np.zeros(4, dtype=np.float32)
        h = fn(_p(rays[i]), _p(other[i]), _p(out))
        assert bool(h) == bool(gold_hit[i]), i
        if h:
            n_hit += 1
            assert np.array_equal(out.view(np.uint32), gold_out[i].view(np.uint32)), i
    return n_hit


def test_device_triangle_and_box_source_on_reference_golden(twin, golden):
    tv = golden["tri_vectors"]
    assert _pairs(twin, twin.dt_triangle, tv["rays"], tv["tris"], tv["hit"], tv["out"]) > 3000
    bv = golden["box_vectors"]
    rays, boxes = np.ascontiguousarray(bv["rays"]), np.ascontiguousarray(bv["boxes"])
    got = np.array([twin.dt_box(_p(rays[i]), _p(boxes[i])) for i in range(len(rays))], dtype=np.uint8)
    assert np.array_equal(got, bv["hit"].astype(np.uint8)) and 1000 < got.sum() < 5500


@pytest.mark.parametrize("name", ["sphere_vectors", "sphere_phi"])
def test_device_sphere_source_on_reference_golden(twin, name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    assert _pairs(twin, twin.dt_sphere, g["rays"], g["spheres"], g["hit"], g["out"]) > 1500


def test_device_atan2_source_matches_host_libm(twin, oracle):
    from test_spheres import _atan2_inputs
    y, x = _atan2_inputs(np.random.default_rng(21), 2_000_000)
    got = np.zeros_like(y)
    twin.dt_atan2(_p(y), _p(x), y.size, _p(got))
    want = oracle.atan2f(y, x)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got.view(np.uint32)[~nan], want.view(np.uint32)[~nan])


def test_device_alpha_source_on_reference_golden(twin, oracle, golden):
    g = np.load(os.path.join(GOLDEN, "alpha.npz"))
    od = np.ascontiguousarray(g["od"])
    got = np.array([twin.dt_hash_float(_p(v[:3].copy()), _p(v[3:].copy())) for v in od], dtype=np.float32)
    assert np.array_equal(got.view(np.uint32), g["hashes"].view(np.uint32))
    tv = golden["tri_vectors"]
    assert _pairs(twin, twin.dt_spawn_origin, tv["rays"], tv["tris"], g["spawn_hit"], g["spawn"]) > 3000
    # GeometricPrimitive::Intersect over one triangle: the device loop against the checker's recursion
    rng = np.random.default_rng(9)
    rays, tris = np.ascontiguousarray(tv["rays"]), np.ascontiguousarray(tv["tris"])
    alpha = rng.random(len(rays)).astype(np.float32)
    alpha[::5] = 0.0
    alpha[1::5] = 1.0
    ignored = 0
    for i in range(len(rays)):
        a, b = np.zeros(4, dtype=np.float32), np.zeros(4, dtype=np.float32)
        h = twin.dt_alpha_triangle(_p(rays[i]), _p(tris[i]), float(alpha[i]), _p(a))
        w = oracle.lib.nnbvh_oracle_alpha_triangle(_p(rays[i]), _p(tris[i]), ctypes.c_float(float(alpha[i])), _p(b))
        assert bool(h) == bool(w), i
        if h:
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), i
        ignored += bool(tv["hit"][i]) and not h
    assert ignored > 800            # the alpha test did remove hits


def test_device_instance_ray_source_matches_checker(twin, oracle):
    from tools.make_golden import instanced_scene
    rng = np.random.default_rng(31)
    _, _, inst, _ = instanced_scene(rng, T=10, K=200, obj_sizes=(5,))
    rays = rand_rays(rng, 4000)
    rays[2000:, 3] = rng.random(2000)
    oracle.lib.nnbvh_oracle_instance_ray.argtypes = [ctypes.c_void_p] * 3
    oracle.lib.nnbvh_oracle_instance_ray.restype = None
    for i in range(len(rays)):
        m = np.ascontiguousarray(inst[i % len(inst), 12:])
        a, b = np.zeros(8, dtype=np.float32), np.zeros(8, dtype=np.float32)
        twin.dt_instance_ray(_p(m), _p(rays[i]), _p(a))
        oracle.lib.nnbvh_oracle_instance_ray(_p(m), _p(rays[i]), _p(b))
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), i


def _traverse(L, nodes, ordered_tris9, rays, sphere_of=None, spheres=None, alpha=None, any_hit=False):
    L.dt_traverse.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                              ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                              ctypes.c_void_p, ctypes.c_void_p]
    L.dt_traverse.restype = None
    nodes = np.ascontiguousarray(nodes)
    t = np.ascontiguousarray(ordered_tris9, dtype=np.float32)
    rays = np.ascontiguousarray(rays, dtype=np.float32)
    n = len(rays)
    hits = np.zeros((n, 5), dtype=np.uint32)
    occ = np.zeros(n, dtype=np.uint8)
    so = None if sphere_of is None else np.ascontiguousarray(sphere_of, dtype=np.int32)
    sp = None if spheres is None else np.ascontiguousarray(spheres, dtype=np.float32)
    al = None if alpha is None else np.ascontiguousarray(alpha, dtype=np.float32)
    L.dt_traverse(_p(nodes), len(nodes), _p(t), len(t), None if so is None else _p(so), None if sp is None else _p(sp),
                  0 if sp is None else len(sp), None if al is None else _p(al), _p(rays), n, _p(hits),
                  _p(occ) if any_hit else None)
    return occ if any_hit else hits


def test_device_kernels_whole_traversal_on_reference_golden(twin, golden):
    """BasicKernel and AlphaKernel themselves -- the one-thread-per-ray kernels, called as ordinary functions with
    blockIdx / threadIdx set per ray -- over the reference's trees: closest hits and IntersectP equal the reference's."""
    g = golden["scene_small"]
    for split in ("sah", "middle"):
        order = g[f"{split}_order"]
        ot = g["tris"][order]
        assert np.array_equal(_traverse(twin, g[f"{split}_nodes"], ot, g["rays"]), g[f"{split}_hits"])
        assert np.array_equal(_traverse(twin, g[f"{split}_nodes"], ot, g["rays"], any_hit=True), g[f"{split}_occ"])
    a = np.load(os.path.join(GOLDEN, "alpha.npz"))
    order = a["order"]
    assert np.array_equal(_traverse(twin, a["nodes"], a["tris"][order], a["rays"], alpha=a["alpha"][order]), a["hits"])
    assert np.array_equal(_traverse(twin, a["nodes"], a["tris"][order], a["rays"], alpha=a["alpha"][order], any_hit=True),
                          a["occ"])
    s = np.load(os.path.join(GOLDEN, "sphere_vectors.npz"))
    T = len(s["mixed_tris"])
    order = s["mixed_order"]
    so = np.where(order >= T, order - T, -1).astype(np.int32)
    ot = np.zeros((len(order), 9), dtype=np.float32)
    ot[order < T] = s["mixed_tris"][order[order < T]]
    assert np.array_equal(_traverse(twin, s["mixed_nodes"], ot, s["mixed_rays"], sphere_of=so, spheres=s["mixed_spheres"]),
                          s["mixed_hits"])
    assert np.array_equal(_traverse(twin, s["mixed_nodes"], ot, s["mixed_rays"], sphere_of=so, spheres=s["mixed_spheres"],
                                    any_hit=True), s["mixed_occ"])


def test_device_instanced_kernel_on_reference_golden(twin):
    """BasicKernel with the two-level traversal (InstanceRay + InnerTraverse) over the reference's instanced golden
    scene: hit records, the primitive inside the object and IntersectP."""
    g = np.load(os.path.join(GOLDEN, "instanced.npz"))
    f = twin.dt_traverse_instanced
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_uint32, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    c = lambda k, dt: np.ascontiguousarray(g[k], dtype=dt)
    nodes, tris, inst_of = c("sc_nodes", np.uint32), c("sc_ordered_tris9", np.float32), c("sc_instance_of", np.int32)
    inst24, inst_obj = c("sc_inst24", np.float32), c("sc_inst_object", np.int32)
    obj_nodes, obj_tris = c("sc_obj_nodes", np.uint32), c("sc_obj_tris", np.float32)
    J = int(g["n_objects"])
    nbase = np.ascontiguousarray(np.concatenate([g["sc_obj_node_base"], [len(obj_nodes)]]), dtype=np.int32)
    pbase = np.ascontiguousarray(np.concatenate([g["sc_obj_prim_base"], [len(obj_tris)]]), dtype=np.int32)
    rays = c("rays", np.float32)
    n = len(rays)
    hits, inner, occ = np.zeros((n, 5), np.uint32), np.zeros(n, np.uint32), np.zeros(n, np.uint8)
    args = (_p(nodes), len(nodes), _p(tris), len(tris), _p(inst_of), _p(inst24), _p(inst_obj), len(inst24), _p(obj_nodes),
            _p(nbase), _p(obj_tris), _p(pbase), J, _p(rays), n)
    f(*args, _p(hits), _p(inner), None)
    assert np.array_equal(hits, g["hits"]) and np.array_equal(inner, g["inner"])
    f(*args, _p(hits), _p(inner), _p(occ))
    assert np.array_equal(occ, g["occ"])


def test_device_alpha_kernel_matches_checker_at_size(twin, oracle):
    """The seeded scene of tests/test_alpha_gpu.py (smaller ray count), AlphaKernel's source on the host against
    the checker: product-built tree, unnormalised directions, finite tMax, alpha in {0, (0,1), 1}."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(44)
    T, N = 60_000, 80_000
    c = rng.random((T, 1, 3), dtype=np.float32)
    tris = ((c + np.float32(0.02) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)
            + np.float32(0)).astype(np.float32)
    a = rng.random(T).astype(np.float32)
    a[::3] = 1.0
    a[1::9] = 0.0
    nodes, order = nb.build_bvh(tris, 4, "sah")
    nodes_u = nodes.view(np.uint32).reshape(-1, 8)
    ot, oa = tris[order], a[order]
    rays = rand_rays(rng, N)
    rays[N // 2:, 3] = rng.random(N - N // 2)
    rays[::5, 4:7] *= np.float32(3.7)
    want = oracle.intersect_alpha(nodes_u, ot, oa, rays)
    assert np.array_equal(_traverse(twin, nodes_u, ot, rays, alpha=oa), want)
    assert (oracle.intersect(nodes_u, ot, rays) != want).any(axis=1).sum() > 2000
    assert np.array_equal(_traverse(twin, nodes_u, ot, rays, alpha=oa, any_hit=True),
                          oracle.intersect_alpha(nodes_u, ot, oa, rays, any_hit=True))


def test_device_kernel_counters_on_reference_golden(twin, golden):
    """BasicKernel's COUNT instantiation: the totals of bvhNodesVisited / nTriTests the reference itself reported."""
    g = golden["scene_small"]
    f = twin.dt_traverse_counters
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_int64,
                  ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    for split in ("sah", "equal"):
        nodes = np.ascontiguousarray(g[f"{split}_nodes"])
        ot = np.ascontiguousarray(g["tris"][g[f"{split}_order"]], dtype=np.float32)
        rays = np.ascontiguousarray(g["rays"], dtype=np.float32)
        n = len(rays)
        tot = []
        for any_hit in (0, 1):
            nv, tt = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
            f(_p(nodes), len(nodes), _p(ot), len(ot), _p(rays), n, any_hit, _p(nv), _p(tt))
            tot += [int(nv.sum()), int(tt.sum())]
        assert tuple(tot) == tuple(int(v) for v in g[f"{split}_stats"])


def test_device_triangle_delta_t_shortcut_never_changes_the_outcome(twin, oracle):
    """exact_math.cuh decides `t <= deltaT` (shapes.cpp:250-269) with a cheap upper bound first and falls back to
    the reference's expression.  Rays that start within a few error bounds of the triangle -- t of the order of
    deltaT, across 40 orders of magnitude of scene scale -- must give the verbatim answer (hit flag and t, b bits)."""
    rng = np.random.default_rng(977)
    n = 30000
    rays = np.zeros((n, 8), dtype=np.float32)
    tris = np.zeros((n, 9), dtype=np.float32)
    accepted = rejected = 0
    for i in range(n):
        scale = 10.0 ** rng.uniform(-21, 18)      # down to where the products go denormal, up to overflow
        tri = (rng.normal(size=(3, 3)) * scale).astype(np.float32)
        if i % 5 == 0:
            tri += np.float32(scale * 10.0 ** rng.uniform(0, 6))        # far from the origin: large |coordinates|, small edges
        b = rng.dirichlet([1, 1, 1])
        p = (b[:, None] * tri.astype(np.float64)).sum(0)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        d *= 10.0 ** rng.uniform(-3, 3)                                 # unnormalised directions too
        eps = np.abs(p).max() * 10.0 ** rng.uniform(-9, -3) if i % 3 else np.abs(p).max() * 10.0 ** rng.uniform(-7.5, -5.5)
        o = p - d / np.linalg.norm(d) * eps
        rays[i, :3], rays[i, 4:7], rays[i, 3] = o, d, np.inf
        tris[i] = tri.ravel()
    for i in range(n):
        got = np.zeros(4, dtype=np.float32)
        h = twin.dt_triangle(_p(rays[i]), _p(tris[i]), _p(got))
        wh, want = oracle.triangle(rays[i], tris[i])
        assert bool(h) == bool(wh), (i, rays[i], tris[i])
        if wh:
            accepted += 1
            assert np.array_equal(got.view(np.uint32), np.asarray(want, dtype=np.float32).view(np.uint32)), i
        else:
            rejected += 1
    assert accepted > 5000 and rejected > 5000
