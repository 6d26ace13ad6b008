"""oracle/nn_tree_oracle.py -- TEST INFRASTRUCTURE (the checker), not product code.

Plain-Python restatement of the fork's NN-split tree semantics, statement for statement, for
small cases.  Paths relative to /root/reference/machine_learning:
  split           nn_BVH.py:32-105      (BVHNode.split, tight_bounds=True branch)
  aabb_of         nn_AABB.py:61-91      (get_AABB_from_vertices / _from_primitives)
  build_from_planes  nn_tree_bench.py:44-77  (build_tree_from_nn_prediction)
  to_list         nn_BVH.py:108-127
  sah             nn_loss.py:165-192
  epo             nn_loss.py:31-135
Parity status: PINNED against tests/golden/nn_tree.npz, which tools/make_golden_nn.py produced
by running the reference's own modules in the dev container.

Only tests/ may import this module.
"""
import math
import sys

import numpy as np

MAX_PRIMITIVES_PER_LEAF = 2   # nn_BVH.py:9
C_INN, C_TRI = 1.2, 1.0       # nn_loss.py:111-115


class Node:
    def __init__(self, aabb, prims):
        self.aabb, self.prims = aabb, prims        # aabb: [xmin,ymin,zmin,xmax,ymax,zmax]; prims: indices
        self.is_leaf = False
        self.parent = self.left = self.right = None


def aabb_of(tris, idx):
    """nn_AABB.py:61-91.  NB the reference seeds the maxima with sys.float_info.min (the smallest
    POSITIVE double), not the lowest one, and returns an all-zero box for an empty list."""
    if len(idx) == 0:
        return [0.0] * 6
    b = [sys.float_info.max] * 3 + [sys.float_info.min] * 3
    for i in idx:
        for v in range(3):
            for k in range(3):
                x = float(tris[i][3 * v + k])
                if x < b[k]:
                    b[k] = x
                if x > b[3 + k]:
                    b[3 + k] = x
    return b


def split(tris, node, axis, pos):
    """nn_BVH.py:32-105."""
    left, right = [], []
    for i in node.prims:
        c = [float(tris[i][3 * v + axis]) for v in range(3)]
        mx, mn = max(c), min(c)
        if mx <= pos:
            left.append(i)
        elif mn > pos:
            right.append(i)
        elif pos - mn >= mx - pos:
            left.append(i)
        else:
            right.append(i)
    node.left, node.right = Node(aabb_of(tris, left), left), Node(aabb_of(tris, right), right)
    node.left.parent = node.right.parent = node


def build_from_planes(tris, planes):
    """nn_tree_bench.py:44-77: level-order rows [onehot axis(3), offset].  Reproduces the shipped
    behaviour: only LEFT children with <= 2 primitives are flagged leaf early (`:74` is a no-op
    expression for the right child), every slot is split regardless, and the last level is
    flagged leaf."""
    axes = []
    for row in planes:
        oh = [float(x) for x in row[:3]]
        if oh not in ([1, 0, 0], [0, 1, 0], [0, 0, 1]):
            raise ValueError("Invalid split axis.")
        axes.append(oh.index(1.0))
    levels = int(math.log2(len(planes) + 1))
    root = Node(aabb_of(tris, list(range(len(tris)))), list(range(len(tris))))
    h = [None] * (2 ** (levels + 1) - 1)
    h[0] = root
    k = 1
    for i in range(len(planes)):
        split(tris, h[i], axes[i], float(planes[i][3]))
        h[k], h[k + 1] = h[i].left, h[i].right
        k += 2
        if len(h[i].left.prims) <= MAX_PRIMITIVES_PER_LEAF:
            h[i].left.is_leaf = True
    for i in range(2 ** levels - 1, len(h)):
        h[i].is_leaf = True
    return root, h


def to_list(root):
    """nn_BVH.py:108-127 (for a non-leaf root)."""
    inner, leaves, todo = [], [], [root]
    while todo:
        cur = todo.pop()
        while not cur.is_leaf:
            inner.append(cur)
            todo.append(cur.right)
            cur = cur.left
        leaves.append(cur)
    return inner, leaves


def _sa(b):
    x, y, z = b[3] - b[0], b[4] - b[1], b[5] - b[2]
    return 2 * (x * y + x * z + y * z)


def sah(root):
    """nn_loss.py:165-192."""
    inner, leaves = to_list(root)
    s = C_INN * sum(_sa(n.aabb) for n in inner) + C_TRI * sum(_sa(n.aabb) * len(n.prims) for n in leaves)
    return s / _sa(root.aabb)


def _area(tris, idx):
    """nn_loss.py:5-18."""
    a = 0.0
    for i in idx:
        p = np.asarray(tris[i], dtype=np.float64).reshape(3, 3)
        a += np.linalg.norm(np.cross(p[1] - p[0], p[2] - p[0]))
    return 0.5 * a


def _overlap(a, b):
    return all(a[k] <= b[3 + k] and a[3 + k] >= b[k] for k in range(3))


def _inside(tris, box, idx):
    """nn_loss.py:31-50: primitives with at least one vertex inside the box."""
    out = []
    for i in idx:
        for v in range(3):
            if all(box[3 + k] >= float(tris[i][3 * v + k]) >= box[k] for k in range(3)):
                out.append(i)
                break
    return out


def external_prims_inside(tris, node):
    """nn_loss.py:52-107: (S \\ Q(n)) intersected with n."""
    if node.parent is None:
        return []
    todo, cur, par = [], node, node.parent
    while par is not None:
        sib = par.right if par.left is cur else par.left
        if _overlap(node.aabb, sib.aabb):
            todo.append(sib)
        cur, par = par, par.parent
    res = []
    while todo:
        cur = todo.pop()
        dead = False
        while not cur.is_leaf:
            lo, ro = _overlap(node.aabb, cur.left.aabb), _overlap(node.aabb, cur.right.aabb)
            if lo:
                if ro:
                    todo.append(cur.right)
                cur = cur.left
            elif ro:
                cur = cur.right
            else:
                dead = True
                break
        if dead or not cur.is_leaf:
            continue
        res.extend(_inside(tris, node.aabb, cur.prims))
    return res


def epo(tris, root):
    """nn_loss.py:119-135."""
    inner, leaves = to_list(root)
    s = C_INN * sum(_area(tris, external_prims_inside(tris, n)) for n in inner)
    s += C_TRI * sum(_area(tris, external_prims_inside(tris, n)) for n in leaves)
    return s / _area(tris, root.prims)
