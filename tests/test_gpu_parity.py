"""GPU: the CUDA engine (through libnnbvh.so's C ABI) against the oracle and the
committed golden vectors.  Bit-exact: hit/miss, primitive index, t, b0..b2 (the
north star allows 2 ulp on t/b; the engine is held to 0), counters included."""
import numpy as np
import pytest

from conftest import rand_rays, soup

pytestmark = pytest.mark.gpu

KERNELS = {"basic": 1, "wide": 2, "step": 3}


def _agg(tris=None, kernel="wide", **kw):
    import nn_bvh_b200 as nb
    return nb.GPUBVHAggregate(tris, kernel=KERNELS[kernel], **kw)


def _as_words(h):
    return np.ascontiguousarray(h).view(np.uint32).reshape(-1, 5)


def test_device_triangle_golden(golden):
    g = golden["tri_vectors"]
    agg = _agg(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), max_prims_in_node=4)
    hit, out = agg.test_triangles(g["rays"], g["tris"])
    assert np.array_equal(hit, g["hit"])
    assert np.array_equal(out.view(np.uint32), g["out"].view(np.uint32))
    assert hit[0] == 0   # Triangle.BadCases, shapes_test.cpp:435-449


def test_device_box_golden(golden):
    g = golden["box_vectors"]
    agg = _agg(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), max_prims_in_node=4)
    assert np.array_equal(agg.test_boxes(g["rays"], g["boxes"]), g["hit"])


@pytest.mark.parametrize("kernel", ["basic", "wide", "step"])
@pytest.mark.parametrize("split", ["sah", "middle", "equal"])
def test_scene_golden(golden, kernel, split):
    g = golden["scene_small"]
    agg = _agg(g["tris"], kernel, max_prims_in_node=4, split_method=split)
    assert np.array_equal(agg.nodes.view(np.uint32).reshape(-1, 8), g[f"{split}_nodes"])
    hits = agg.IntersectBatch(g["rays"])
    assert np.array_equal(_as_words(hits), g[f"{split}_hits"])
    assert np.array_equal(agg.IntersectPBatch(g["rays"]), g[f"{split}_occ"])
    st = g[f"{split}_stats"]
    h2, nv, tt = agg.IntersectCounters(g["rays"])
    assert np.array_equal(_as_words(h2), g[f"{split}_hits"])
    assert (int(nv.sum()), int(tt.sum())) == (st[0], st[1])
    o2, nv, tt = agg.IntersectCounters(g["rays"], any_hit=True)
    assert np.array_equal(o2, g[f"{split}_occ"])
    assert (int(nv.sum()), int(tt.sum())) == (st[2], st[3])


@pytest.mark.parametrize("kernel", ["basic", "wide", "step"])
@pytest.mark.parametrize("T,s,mp", [(1, 0.3, 4), (2, 0.3, 4), (37, 0.2, 4), (20000, 0.02, 4),
                                     (20000, 0.05, 1), (60000, 0.01, 255)])
def test_random_soups_vs_oracle(oracle, kernel, T, s, mp):
    rng = np.random.default_rng(T * 7 + mp)
    tris = soup(rng, T, s)
    agg = _agg(tris, kernel, max_prims_in_node=mp)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    rays = rand_rays(rng, 100_000)
    rays[50_000:, 3] = rng.random(50_000)
    want, wnv, wtt = oracle.intersect(nodes, agg.ordered_tris9, rays, counters=True)
    assert np.array_equal(_as_words(agg.IntersectBatch(rays)), want)
    assert np.array_equal(agg.IntersectPBatch(rays), oracle.intersect_p(nodes, agg.ordered_tris9, rays))
    _, nv, tt = agg.IntersectCounters(rays)
    assert np.array_equal(nv, wnv) and np.array_equal(tt, wtt)


@pytest.mark.parametrize("kernel", ["basic", "wide", "step"])
def test_edge_inputs(oracle, kernel):
    rng = np.random.default_rng(3)
    tris = soup(rng, 5000, 0.05)
    agg = _agg(tris, kernel, max_prims_in_node=4)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    # empty batch
    assert len(agg.IntersectBatch(np.zeros((0, 8), np.float32))) == 0
    # ragged sizes around the warp / block / chunk boundaries
    for n in (1, 31, 32, 33, 255, 257, 4097):
        rays = rand_rays(rng, n)
        assert np.array_equal(_as_words(agg.IntersectBatch(rays)),
                              oracle.intersect(nodes, agg.ordered_tris9, rays)), n
    # axis-aligned directions incl. -0.0, origins on box planes, tMax = 0, inf, tiny
    rays = rand_rays(rng, 6000)
    rays[:2000, 4:7] = 0
    rays[np.arange(2000), 4 + np.arange(2000) % 3] = np.where(np.arange(2000) % 2, 1.0, -1.0)
    rays[1000:2000, 4:7] = np.where(rays[1000:2000, 4:7] == 0, np.float32(-0.0), rays[1000:2000, 4:7])
    rays[2000:3000, 0] = nodes[0, 0:1].view(np.float32)   # origin.x on the root's pMin.x plane
    rays[3000:3500, 3] = 0.0
    rays[3500:4000, 3] = 1e-30
    want = oracle.intersect(nodes, agg.ordered_tris9, rays)
    assert np.array_equal(_as_words(agg.IntersectBatch(rays)), want)
    assert np.array_equal(agg.IntersectPBatch(rays), oracle.intersect_p(nodes, agg.ordered_tris9, rays))


@pytest.mark.parametrize("kernel", ["basic", "wide", "step"])
def test_empty_scene_misses(kernel):
    import nn_bvh_b200 as nb
    agg = _agg(None, kernel)
    rays = rand_rays(np.random.default_rng(0), 100)
    h = agg.IntersectBatch(rays)
    assert np.all(h["prim"] == nb.MISS) and not agg.IntersectPBatch(rays).any()
    with pytest.raises(RuntimeError):
        agg.Bounds()


def test_deep_tree_rejected():
    import nn_bvh_b200 as nb
    k = 70
    nodes = np.zeros(2 * k + 1, dtype=nb.NODE_DTYPE)
    for i in range(k):
        nodes[i]["offset"] = 2 * k - i
        nodes[i]["bounds"] = [0, 0, 0, 1, 1, 1]
    for j in range(k, 2 * k + 1):
        nodes[j]["nPrimitives"] = 1
        nodes[j]["offset"] = j - k
        nodes[j]["bounds"] = [0, 0, 0, 1, 1, 1]
    tris = np.tile(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), (k + 1, 1))
    with pytest.raises(nb.NnbvhError) as ei:
        nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris)
    assert ei.value.code == -4


def test_watertight_sphere(oracle):
    """Triangle.Watertight (shapes_test.cpp:31-131, disabled upstream): rays from inside a
    jittered triangulated sphere, many aimed exactly at vertices, must always hit."""
    rng = np.random.default_rng(12111)
    nt, npk = 16, 16
    th = np.linspace(0, np.pi, nt + 1)
    ph = np.linspace(0, 2 * np.pi, npk + 1)[:-1]
    P = np.array([[np.sin(t) * np.cos(p), np.sin(t) * np.sin(p), np.cos(t)] for t in th for p in ph])
    P = (P * (1 + 0.05 * rng.random((len(P), 1)))).astype(np.float32)
    P[:npk] = P[0]
    P[-npk:] = P[-1]
    idx = []
    for i in range(nt):
        for j in range(npk):
            a, b = i * npk + j, i * npk + (j + 1) % npk
            c, d = a + npk, b + npk
            idx += [[a, c, b], [b, c, d]]
    tris = P[np.array(idx)].reshape(-1, 9) + np.float32(0)
    agg = _agg(tris, "step", max_prims_in_node=4)
    n = 100_000
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = rng.random((n, 3)) * 0.2 - 0.1
    d = rng.normal(size=(n, 3))
    aim = rng.integers(0, len(P), size=n)
    d[: n // 2] = P[aim[: n // 2]] - rays[: n // 2, :3]
    rays[:, 4:7] = d
    rays[:, 3] = np.inf
    h = agg.IntersectBatch(rays)
    import nn_bvh_b200 as nb
    assert np.all(h["prim"] != nb.MISS)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    assert np.array_equal(_as_words(h), oracle.intersect(nodes, agg.ordered_tris9, rays))


def test_single_ray_surface_and_tensors(oracle):
    import torch
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(8)
    tris = soup(rng, 3000, 0.05)
    agg = nb.GPUBVHAggregate.Create(tris, {"maxnodeprims": 4, "splitmethod": "sah"})
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    rays = rand_rays(rng, 50_000)
    want = oracle.intersect(nodes, agg.ordered_tris9, rays)
    # single-ray BVHAggregate::Intersect / IntersectP surface
    for i in range(20):
        r = agg.Intersect((rays[i, :3], rays[i, 4:7]), rays[i, 3])
        if want[i, 0] == nb.MISS:
            assert r is None and not agg.IntersectP((rays[i, :3], rays[i, 4:7]))
        else:
            assert r[0] == want[i, 0] and np.float32(r[1]).view(np.uint32) == want[i, 1]
            assert agg.IntersectP((rays[i, :3], rays[i, 4:7]))
    # device-resident tensors (no copies inside the call) and pinned host tensors
    dr = torch.from_numpy(rays).cuda()
    dh = torch.zeros((len(rays), 5), dtype=torch.int32, device="cuda")
    agg.IntersectBatch(dr, hits=dh)
    torch.cuda.synchronize()
    assert agg.last_kernel_ms() > 0
    assert np.array_equal(dh.cpu().numpy().view(np.uint32), want)
    pr = torch.from_numpy(rays).pin_memory()
    ph = torch.zeros((len(rays), 5), dtype=torch.int32).pin_memory()
    agg.IntersectBatch(pr, hits=ph)
    assert np.array_equal(ph.numpy().view(np.uint32), want)
    # SoA entry points (wavefront queue layout)
    cols = [np.ascontiguousarray(rays[:, k]) for k in (0, 1, 2, 4, 5, 6, 3)]
    hs = agg.IntersectBatchSoA(*cols)
    assert np.array_equal(_as_words(hs), want)
    occ = agg.IntersectBatchSoA(*cols, any_hit=True)
    assert np.array_equal(occ, oracle.intersect_p(nodes, agg.ordered_tris9, rays))
    assert agg.launch_count() > 0


@pytest.mark.parametrize("kernel", ["wide", "step"])
@pytest.mark.parametrize("mp", [1, 4, 32])
def test_axis_aligned_room_near_ties(oracle, kernel, mp):
    """Flat axis-aligned geometry: leaf boxes are flat, so box-entry distances tie with hit
    distances to the ulp, and rays aimed at grid vertices / edges hit several triangles at the
    same t (shared edges).  This is where 'tMax grows by an ulp' (shapes.cpp:239-247) and the
    in-leaf order matter; results must still be bit-identical."""
    from nn_bvh_b200 import scenes
    tris, _ = scenes.procedural_room(n_objects=6, sphere_level=3, seed=7)
    agg = _agg(tris, kernel, max_prims_in_node=mp)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    rng = np.random.default_rng(11 + mp)
    n = 300_000
    rays = np.zeros((n, 8), np.float32)
    o = np.stack([rng.uniform(-2.9, 2.9, n), rng.uniform(0.1, 2.9, n), rng.uniform(-2.9, 2.9, n)], 1)
    # targets: wall grid vertices (64x64 grid over [-3,3]x[0,3]x[-3,3]) on a random wall
    g = np.linspace(0.0, 1.0, 65)
    u, v = rng.integers(0, 65, n), rng.integers(0, 65, n)
    axis, side = rng.integers(0, 3, n), rng.integers(0, 2, n)
    lo, hi = np.array([-3.0, 0.0, -3.0]), np.array([3.0, 3.0, 3.0])
    tgt = np.zeros((n, 3))
    for a in range(3):
        oth = [k for k in range(3) if k != a]
        m = axis == a
        tgt[m, oth[0]] = lo[oth[0]] + g[u[m]] * (hi[oth[0]] - lo[oth[0]])
        tgt[m, oth[1]] = lo[oth[1]] + g[v[m]] * (hi[oth[1]] - lo[oth[1]])
        tgt[m, a] = np.where(side[m] == 1, hi[a], lo[a])
    tgt[n // 2:] += rng.normal(scale=1e-3, size=(n - n // 2, 3)) * (rng.random((n - n // 2, 1)) < 0.5)
    rays[:, :3] = o
    rays[:, 4:7] = (tgt - o)
    rays[: n // 4, 4:7] /= np.linalg.norm(rays[: n // 4, 4:7], axis=1, keepdims=True)
    rays[:, 3] = np.inf
    rays[n // 3: n // 2, 3] = 1.0    # shadow-style: d = target - origin, tMax = 1 (hit exactly at tMax)
    want = oracle.intersect(nodes, agg.ordered_tris9, rays)
    assert np.array_equal(_as_words(agg.IntersectBatch(rays)), want)
    assert np.array_equal(agg.IntersectPBatch(rays), oracle.intersect_p(nodes, agg.ordered_tris9, rays))


@pytest.mark.parametrize("any_hit", [False, True])
def test_binning_prepass_never_changes_results(oracle, any_hit, monkeypatch):
    """The ray-ordering pre-pass (binning.cuh) is invisible: forced on, off, or decided per batch on
    the device (threshold 0 = always sort, 33 = never), hits[i] still belongs to rays[i], bit for bit.
    Batch sizes straddle the radix tile (2048) and the 64 Ki-ray minimum."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(31)
    tris = soup(rng, 30_000, 0.02)
    nodes, order = nb.build_bvh(tris, 4, "sah")
    ot = tris[order]
    n = 65_536 + 2048 + 77
    incoherent = rand_rays(rng, n)
    incoherent[n // 2:, 3] = rng.random(n - n // 2)
    coherent = np.zeros((n, 8), np.float32)            # a pinhole: one origin, slowly varying direction
    coherent[:, :3] = [0.5, 0.5, -1.0]
    u = (np.arange(n) % 512) / 512 - 0.5
    v = (np.arange(n) // 512) / (n // 512 + 1) - 0.5
    coherent[:, 4:7] = np.stack([u, v, np.ones(n)], 1)
    coherent[:, 3] = np.inf
    for rays in (incoherent, coherent):
        want = (oracle.intersect_p if any_hit else oracle.intersect)(nodes.view(np.uint32).reshape(-1, 8), ot, rays)
        for env in ({"NNBVH_BIN": "0"}, {"NNBVH_BIN": "1"}, {"NNBVH_BIN": "2"},
                    {"NNBVH_BIN": "2", "NNBVH_BIN_THRESHOLD": "0"},
                    {"NNBVH_BIN": "2", "NNBVH_BIN_THRESHOLD": "33"},
                    {"NNBVH_BIN": "1", "NNBVH_BIN_CELL_BITS": "8", "NNBVH_BIN_DIR_BITS": "4"},
                    {"NNBVH_BIN": "1", "NNBVH_BIN_CELL_BITS": "2", "NNBVH_BIN_DIR_BITS": "0"}):
            for k in ("NNBVH_BIN", "NNBVH_BIN_THRESHOLD", "NNBVH_BIN_CELL_BITS", "NNBVH_BIN_DIR_BITS"):
                monkeypatch.delenv(k, raising=False)
            monkeypatch.setenv("NNBVH_BIN_MIN_RAYS", "65536")   # defaults skip batches / scenes this small
            monkeypatch.setenv("NNBVH_BIN_MIN_NODES", "1")
            for k, val in env.items():
                monkeypatch.setenv(k, val)
            agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=ot)
            if any_hit:
                got = agg.IntersectPBatch(rays)
            else:
                got = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
            assert np.array_equal(got, want), env
            agg.close()


def test_binning_decision_is_made_on_the_device(monkeypatch):
    """Mode 2 sorts incoherent batches and leaves coherent ones alone; after two coherent batches in a
    row the host stops enqueueing the sort kernels (launch counts: 2 decision kernels + 2 traversal
    launches instead of 2 + key + 3 x 4 + 2)."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(32)
    tris = soup(rng, 30_000, 0.02)
    nodes, order = nb.build_bvh(tris, 4, "sah")
    n = 1 << 17
    incoherent = rand_rays(rng, n)
    coherent = np.zeros((n, 8), np.float32)
    coherent[:, :3] = [0.5, 0.5, -1.0]
    coherent[:, 4:7] = np.stack([(np.arange(n) % 512) / 512 - 0.5, (np.arange(n) // 512) / 256 - 0.5, np.ones(n)], 1)
    coherent[:, 3] = np.inf
    monkeypatch.setenv("NNBVH_BIN", "2")
    monkeypatch.setenv("NNBVH_BIN_MIN_RAYS", "65536")
    monkeypatch.setenv("NNBVH_BIN_MIN_NODES", "1")
    import torch
    agg = nb.GPUBVHAggregate(nodes=nodes, ordered_tris9=tris[order])
    hits = torch.zeros((n, 5), dtype=torch.int32, device="cuda")

    def launches(rays):
        d = torch.from_numpy(rays).cuda()
        c0 = agg.launch_count()
        agg.IntersectBatch(d, hits=hits, n=n)
        torch.cuda.synchronize()
        return agg.launch_count() - c0

    full = 2 + 1 + 3 * 4 + 2
    assert launches(incoherent) == full
    assert launches(incoherent) == full
    assert launches(coherent) == full            # last decision seen by the host: "sort"
    assert launches(coherent) == full            # one coherent batch seen
    assert launches(coherent) == 4               # two in a row: sort kernels no longer enqueued
    assert launches(coherent) == 4
    assert launches(incoherent) == 4             # the device says "sort", the host learns it afterwards
    assert launches(incoherent) == full
    agg.close()


def test_read_bandwidth_probe_orders_l2_above_hbm():
    """The roofline denominators bench.py reports next to the HBM peak: an L2-resident buffer streams
    faster than one far larger than L2, and both are in the range a B200 can produce."""
    import nn_bvh_b200 as nb
    agg = nb.GPUBVHAggregate(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), 4)
    l2 = agg.read_bandwidth_probe(32 << 20, 20)
    hbm = agg.read_bandwidth_probe(2 << 30, 2)
    assert 1000 < hbm < 9000 and l2 > hbm
    with pytest.raises(nb.NnbvhError):
        agg.read_bandwidth_probe(0, 1)
    agg.close()


@pytest.mark.gpu
def test_clone_shares_the_scene_and_runs_concurrently(oracle):
    """nnbvh_clone: a second engine on the same uploaded scene with its own stream; batches issued on both at
    once give the oracle's records; the clone survives the original and a re-upload into the original."""
    import torch
    import nn_bvh_b200 as nb
    from conftest import rand_rays, soup
    rng = np.random.default_rng(77)
    tris = soup(rng, 30_000, 0.02)
    a = nb.GPUBVHAggregate(tris, 4, "sah")
    b = a.clone()
    N = 300_000
    rays = rand_rays(rng, N)
    rays[N // 2:, 3] = rng.random(N - N // 2)
    nodes = a.nodes.view(np.uint32).reshape(-1, 8)
    want = oracle.intersect(nodes, a.ordered_tris9, rays)
    wocc = oracle.intersect_p(nodes, a.ordered_tris9, rays)
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    a.set_stream(sa.cuda_stream)
    b.set_stream(sb.cuda_stream)
    d = torch.from_numpy(rays).cuda()
    hits = torch.zeros((N, 5), dtype=torch.int32, device="cuda")
    occ = torch.zeros(N, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(3):
        a.IntersectBatch(d, hits=hits, n=N)
        b.IntersectPBatch(d, occluded=occ, n=N)
    torch.cuda.synchronize()
    assert np.array_equal(hits.cpu().numpy().view(np.uint32), want)
    assert np.array_equal(occ.cpu().numpy(), wocc)
    # a new scene in the original leaves the clone on the old one; then the original goes away
    a2 = nb.GPUBVHAggregate(tris[:1000], 4, "sah")
    from nn_bvh_b200 import capi
    capi.check(capi.lib().nnbvh_upload_scene(a._h, capi.ptr(a2.nodes), len(a2.nodes), capi.ptr(a2.ordered_tris9),
                                             len(a2.ordered_tris9)))
    small = oracle.intersect(a2.nodes.view(np.uint32).reshape(-1, 8), a2.ordered_tris9, rays[:50_000])
    assert np.array_equal(np.ascontiguousarray(a.IntersectBatch(rays[:50_000])).view(np.uint32).reshape(-1, 5), small)
    a.close()
    assert np.array_equal(np.ascontiguousarray(b.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5), want)
    b.close()
