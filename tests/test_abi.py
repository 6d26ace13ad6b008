"""CPU: the C-ABI library loads, exports what include/nnbvh.h declares, the host
builder reproduces the reference tree, and compute fails loudly without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, soup


def test_library_exports_every_declared_symbol():
    from nn_bvh_b200 import capi
    L = capi.lib()
    hdr = open(os.path.join(ROOT, "include", "nnbvh.h")).read()
    declared = set(re.findall(r"\b(nnbvh_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in L.nnbvh_version()


def test_struct_sizes_match_header():
    from nn_bvh_b200 import capi
    assert capi.NODE_DTYPE.itemsize == 32   # sizeof(LinearBVHNode), aggregates.cpp:129
    assert capi.RAY_DTYPE.itemsize == 32
    assert capi.HIT_DTYPE.itemsize == 20
    assert ctypes.sizeof(capi.Options) == 32


@pytest.mark.parametrize("split", ["sah", "middle", "equal"])
def test_host_builder_matches_golden_reference_tree(golden, split):
    import nn_bvh_b200 as nb
    g = golden["scene_small"]
    nodes, order = nb.build_bvh(g["tris"], 4, split)
    assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), g[f"{split}_nodes"])
    assert np.array_equal(order, g[f"{split}_order"].astype(np.uint32))


def test_host_builder_matches_oracle_and_threads(oracle):
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(5)
    tris = soup(rng, 200_000, 0.01)   # > 64K spans: exercises the threaded path
    n1, o1 = nb.build_bvh(tris, 4, "sah", threads=1)
    n8, o8 = nb.build_bvh(tris, 4, "sah", threads=8)
    on, oo = oracle.build_bvh(tris, 4, "sah")
    assert np.array_equal(n1, n8) and np.array_equal(o1, o8)
    assert np.array_equal(n1.view(np.uint32).reshape(-1, 8), on)
    assert np.array_equal(o1, oo.astype(np.uint32))
    assert nb.tree_stack_depth(n1) <= 64


@pytest.mark.parametrize("split", ["sah", "middle", "equal"])
def test_host_builder_teams_and_forks_are_deterministic(oracle, split):
    """Large spans are worked on by several threads (chunked reductions, the two-pointer partition replayed from
    per-chunk position lists) and large children are forked: the tree must not depend on the thread count.
    The soup is lopsided (a dense cluster in a sparse room) and full of -0.0 / +0.0 coordinates, whose order
    decides which zero std::min keeps."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(17)
    tris = np.concatenate([soup(rng, 150_000, 0.002) * np.float32(0.05), soup(rng, 150_000, 0.02)]).astype(np.float32)
    tris[rng.random(tris.shape) < 0.02] = 0.0
    tris[rng.random(tris.shape) < 0.02] = -0.0
    tris = tris[rng.permutation(len(tris))]
    on, oo = oracle.build_bvh(tris, 4, split)
    for threads in (1, 3, 8):
        n, o = nb.build_bvh(tris, 4, split, threads=threads)
        assert np.array_equal(n.view(np.uint32).reshape(-1, 8), on), threads
        assert np.array_equal(o, oo.astype(np.uint32)), threads


def test_builder_edge_cases():
    import nn_bvh_b200 as nb
    one = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32)
    nodes, order = nb.build_bvh(one, 4, "sah")
    assert len(nodes) == 1 and nodes[0]["nPrimitives"] == 1 and order[0] == 0
    same = np.tile(one, (7, 1))   # identical centroids -> one leaf of 7 (aggregates.cpp:225)
    nodes, _ = nb.build_bvh(same, 4, "sah")
    assert len(nodes) == 1 and nodes[0]["nPrimitives"] == 7
    with pytest.raises(nb.NnbvhError):
        nb.build_bvh(np.zeros((0, 9), np.float32), 4, "sah")
    nodes, order = nb.build_bvh(one, 4, "hlbvh")   # one treelet, one leaf
    assert len(nodes) == 1 and nodes[0]["nPrimitives"] == 1 and order[0] == 0
    # identical primitives share one Morton code: a single leaf however large maxPrims is not
    # reached by bit splits (emitLBVH bitIndex == -1, aggregates.cpp:458)
    nodes, _ = nb.build_bvh(same, 4, "hlbvh")
    assert len(nodes) == 1 and nodes[0]["nPrimitives"] == 7


@pytest.mark.parametrize("name,mp", [("small4", 4), ("soup4", 4), ("soup1", 1), ("soup255", 255)])
def test_host_hlbvh_matches_reference_golden(golden, name, mp):
    """The host HLBVH builder against trees the unmodified reference produced
    (tools/make_golden.py: hlbvh), byte for byte, at any thread count."""
    import nn_bvh_b200 as nb
    g = golden["hlbvh"]
    tris = golden["scene_small"]["tris"] if name == "small4" else g["soup"]
    for threads in (1, 5):
        nodes, order = nb.build_bvh(tris, mp, "hlbvh", threads=threads)
        assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), g[f"{name}_nodes"])
        assert np.array_equal(order, g[f"{name}_order"].astype(np.uint32))
    assert nb.tree_stack_depth(nodes) <= 64


def test_host_hlbvh_matches_oracle_large(oracle):
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(6)
    tris = soup(rng, 250_000, 0.004)
    nodes, order = nb.build_bvh(tris, 4, "hlbvh")
    on, oo = oracle.build_bvh(tris, 4, "hlbvh")
    assert np.array_equal(nodes.view(np.uint32).reshape(-1, 8), on)
    assert np.array_equal(order, oo.astype(np.uint32))


def test_stack_depth_of_chain():
    import nn_bvh_b200 as nb
    # left-deep chain of k interior nodes
    k = 70
    nodes = np.zeros(2 * k + 1, dtype=nb.NODE_DTYPE)
    # pre-order: interior i at index i (i<k), its first child i+1, second child = leaf at 2k-i
    for i in range(k):
        nodes[i]["offset"] = 2 * k - i
    for j in range(k, 2 * k + 1):
        nodes[j]["nPrimitives"] = 1
        nodes[j]["offset"] = j - k
    assert nb.tree_stack_depth(nodes) == k


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import nn_bvh_b200 as nb
    with pytest.raises(nb.NnbvhError) as ei:
        nb.GPUBVHAggregate(np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0]], np.float32), 4)
    assert ei.value.code == -2 and "no CPU fallback" in str(ei.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "nn_bvh_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".sh")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "pyoracle" not in txt and "nnbvh_oracle" not in txt and "_ref/" not in txt, f


def _build_cpp_test():
    import subprocess
    subprocess.check_call(["sh", os.path.join(ROOT, "tests", "cpp", "build.sh")],
                          stdout=subprocess.DEVNULL)
    return os.path.join(ROOT, "tests", "cpp", "host_mirror_test")


def test_cpp_host_mirror_compiles_and_fails_loudly_without_gpu():
    """include/nnbvh.hpp (the C++ mirror of BVHAggregate) compiles against the C ABI;
    without a device the program reports NNBVH_ERR_CUDA (exit 3), never a CPU answer."""
    import subprocess
    import torch
    exe = _build_cpp_test()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by test_cpp_host_mirror_gpu")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_host_mirror_gpu():
    import subprocess
    r = subprocess.run([_build_cpp_test()], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("PASS"), r.stdout + r.stderr


def test_host_builders_match_oracle_on_many_small_inputs(oracle):
    """Every split method on 300 tiny soups (1..40 triangles, duplicates, flat and collinear sets, huge
    and tiny coordinates): nodes and primitive order byte for byte."""
    import nn_bvh_b200 as nb
    rng = np.random.default_rng(2026)
    for case in range(300):
        T = int(rng.integers(1, 41))
        scale = np.float32(10.0 ** rng.integers(-6, 7))
        tris = (rng.random((T, 9), dtype=np.float32) * scale).astype(np.float32)
        kind = case % 6
        if kind == 1:
            tris[T // 2:] = tris[: T - T // 2]            # duplicates
        elif kind == 2:
            tris[:, 2::3] = tris[0, 2]                    # planar
        elif kind == 3:
            tris[:, 1::3] = tris[0, 1]
            tris[:, 2::3] = tris[0, 2]                    # collinear centroids
        elif kind == 4:
            tris[:] = tris[0]                             # all identical
        mp = int(rng.choice([1, 2, 4, 8, 255]))
        for split in ("sah", "hlbvh", "middle", "equal"):
            try:
                n, o = nb.build_bvh(tris, mp, split)
            except nb.NnbvhError:
                # only HLBVH may refuse: treelet roots with one shared centroid (the reference CHECKs)
                assert split == "hlbvh", (case, split)
                with pytest.raises(Exception):
                    oracle.build_bvh(tris, mp, split)
                continue
            on, oo = oracle.build_bvh(tris, mp, split)
            assert np.array_equal(n.view(np.uint32).reshape(-1, 8), on), (case, split, T, mp)
            assert np.array_equal(o, oo.astype(np.uint32)), (case, split, T, mp)
