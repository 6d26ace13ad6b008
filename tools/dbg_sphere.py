import sys, numpy as np
sys.path.insert(0, '.')
import nn_bvh_b200 as nb
from oracle.pyoracle import Oracle
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
n = 100_000
rays = np.zeros((n, 8), np.float32)
rays[:, :3] = rng.random((n, 3)) * 0.2 - 0.1
d = rng.normal(size=(n, 3))
aim = rng.integers(0, len(P), size=n)
d[: n // 2] = P[aim[: n // 2]] - rays[: n // 2, :3]
rays[:, 4:7] = d
rays[:, 3] = np.inf
O = Oracle()
for kern in (1, 2):
    agg = nb.GPUBVHAggregate(tris, 4, kernel=kern)
    nodes = agg.nodes.view(np.uint32).reshape(-1, 8)
    h = np.ascontiguousarray(agg.IntersectBatch(rays)).view(np.uint32).reshape(-1, 5)
    w = O.intersect(nodes, agg.ordered_tris9, rays)
    bad = np.where((h != w).any(1))[0]
    print("kernel", kern, "mismatches", len(bad), bad[:10])
    for i in bad[:5]:
        print(i, "ray", rays[i], "\n  got ", h[i], h[i, 1:].view(np.float32), "\n  want", w[i], w[i, 1:].view(np.float32))
        print("  tri got", agg.ordered_tris9[h[i, 0]], "\n  tri want", agg.ordered_tris9[w[i, 0]])
        hh, oo = agg.test_triangles(np.repeat(rays[i:i+1], 2, 0), agg.ordered_tris9[[h[i,0], w[i,0]]])
        print("  device tri test:", hh, oo.view(np.uint32))
        print("  oracle tri test:", O.triangle(rays[i], agg.ordered_tris9[h[i,0]]), O.triangle(rays[i], agg.ordered_tris9[w[i,0]]))
