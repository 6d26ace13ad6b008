"""tools/pbrt_scene.py -- minimal pbrt-v4 scene reader: triangles + camera only.

Reads just enough of a .pbrt file (transform stack, Include, plymesh /
trianglemesh / loopsubdiv shapes, LookAt / Camera / Film) to hand the engine a
world-space triangle soup and a pinhole camera.  It is the host-side stand-in
for the reference's parser + BasicScene::CreateAggregate
(/root/reference/src/pbrt/scene.cpp:1352-1591), which are out of scope
(SURVEY.md section 2.1): materials, lights, media, textures and non-triangle shapes
(spheres ...) are skipped, so oracle and engine always see the same
triangle-only geometry.

Coordinates are WORLD space (pbrt's `--render-coord-sys world`); pbrt's
default CameraWorld space only translates them by the camera position.
"""
from __future__ import annotations

import math
import os
import re

import numpy as np

_TOKEN = re.compile(r'"[^"]*"|\[|\]|[^\s\[\]"]+')


def _tokenize(path):
    with open(path, "r", errors="replace") as f:
        for line in f:
            line = line.split("#", 1)[0]
            for m in _TOKEN.finditer(line):
                yield m.group(0)


def _translate(x, y, z):
    m = np.eye(4)
    m[:3, 3] = (x, y, z)
    return m


def _scale(x, y, z):
    return np.diag([x, y, z, 1.0])


def _rotate(deg, x, y, z):
    a = np.array([x, y, z], dtype=np.float64)
    a /= np.linalg.norm(a)
    s, c = math.sin(math.radians(deg)), math.cos(math.radians(deg))
    m = np.eye(4)
    m[0, 0] = a[0] * a[0] + (1 - a[0] * a[0]) * c
    m[0, 1] = a[0] * a[1] * (1 - c) - a[2] * s
    m[0, 2] = a[0] * a[2] * (1 - c) + a[1] * s
    m[1, 0] = a[0] * a[1] * (1 - c) + a[2] * s
    m[1, 1] = a[1] * a[1] + (1 - a[1] * a[1]) * c
    m[1, 2] = a[1] * a[2] * (1 - c) - a[0] * s
    m[2, 0] = a[0] * a[2] * (1 - c) - a[1] * s
    m[2, 1] = a[1] * a[2] * (1 - c) + a[0] * s
    m[2, 2] = a[2] * a[2] + (1 - a[2] * a[2]) * c
    return m


def _lookat(e, l, u):
    e, l, u = (np.array(v, dtype=np.float64) for v in (e, l, u))
    d = l - e
    d /= np.linalg.norm(d)
    right = np.cross(u / np.linalg.norm(u), d)
    right /= np.linalg.norm(right)
    newup = np.cross(d, right)
    world_from_cam = np.eye(4)
    world_from_cam[:3, 0] = right
    world_from_cam[:3, 1] = newup
    world_from_cam[:3, 2] = d
    world_from_cam[:3, 3] = e
    return np.linalg.inv(world_from_cam)


def read_ply(path):
    """Binary-little-endian / ascii PLY -> (verts (V,3) f32, tri indices (T,3) i32)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    header = data[:end].decode("ascii", "replace").splitlines()
    fmt = None
    elements = []
    for ln in header:
        t = ln.split()
        if not t:
            continue
        if t[0] == "format":
            fmt = t[1]
        elif t[0] == "element":
            elements.append({"name": t[1], "count": int(t[2]), "props": []})
        elif t[0] == "property":
            elements[-1]["props"].append(t[1:])
    tmap = {"float": "<f4", "float32": "<f4", "double": "<f8", "uchar": "u1", "uint8": "u1",
            "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2",
            "uint16": "<u2", "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}
    if fmt != "binary_little_endian":
        raise ValueError(f"{path}: unsupported PLY format {fmt}")
    off = end
    verts = idx = None
    for el in elements:
        if el["name"] == "vertex":
            dt = np.dtype([(p[-1], tmap[p[0]]) for p in el["props"]])
            a = np.frombuffer(data, dtype=dt, count=el["count"], offset=off)
            off += dt.itemsize * el["count"]
            verts = np.stack([a["x"], a["y"], a["z"]], axis=1).astype(np.float32)
        elif el["name"] == "face":
            lists = [p for p in el["props"] if p[0] == "list"]
            if len(el["props"]) != 1 or len(lists) != 1:
                # face_indices etc. -- handle the common (list, scalar...) layouts
                pass
            if el["count"] == 0:
                idx = np.zeros((0, 3), dtype=np.int32)
                continue
            fields = []
            for p in el["props"]:
                if p[0] == "list":
                    fields.append((p[-1] + "_n", tmap[p[1]]))
                    fields.append((p[-1], tmap[p[2]], (3,)))
                else:
                    fields.append((p[-1], tmap[p[0]]))
            dt = np.dtype(fields)
            a = np.frombuffer(data, dtype=dt, count=el["count"], offset=off)
            lname = lists[0][-1]
            if not np.all(a[lname + "_n"] == 3):
                raise ValueError(f"{path}: non-triangle faces (quads become bilinear patches in "
                                 "pbrt and are outside the triangle-only contract)")
            off += dt.itemsize * el["count"]
            idx = a[lname].astype(np.int32)
        else:
            raise ValueError(f"{path}: unexpected element {el['name']}")
    return verts, idx


class Scene:
    def __init__(self):
        self.tris = []       # list of (T,9) float32 world-space soups
        self.skipped = {}    # shape type -> count
        self.spheres = []    # Shape "sphere": dicts {render_from_object (3x4), radius, zmin, zmax, phimax}
        self.camera = {}     # eye/look/up or matrix, fov, res

    def soup(self):
        return np.concatenate(self.tris, axis=0) if self.tris else np.zeros((0, 9), np.float32)


def load(path, subdivide=None) -> Scene:
    """subdivide(levels, indices, verts) -> (indices, verts) implements loopsubdiv (the
    harness passes the reference's own LoopSubdivide through oracle/_ref)."""
    sc = Scene()
    ctm = np.eye(4)
    stack = []
    _walk(path, sc, [ctm], stack, subdivide)
    return sc


def _walk(path, sc, ctm_box, stack, subdivide):
    toks = list(_tokenize(path))
    base = os.path.dirname(path)
    i = 0
    n = len(toks)

    def is_directive(t):
        return t[0].isalpha() and t[0].isupper() and not t.startswith('"')

    def floats(k):
        nonlocal i
        out = [float(toks[i + j]) for j in range(k)]
        i += k
        return out

    def params():
        """Parse `"type name" value|[values]` pairs until the next directive."""
        nonlocal i
        out = {}
        while i < n and toks[i].startswith('"'):
            decl = toks[i].strip('"').split()
            i += 1
            vals = []
            if toks[i] == "[":
                i += 1
                while toks[i] != "]":
                    vals.append(toks[i])
                    i += 1
                i += 1
            else:
                vals.append(toks[i])
                i += 1
            out[decl[-1]] = (decl[0], vals)
        return out

    while i < n:
        t = toks[i]
        i += 1
        if t == "AttributeBegin" or t == "TransformBegin":
            stack.append(ctm_box[0].copy())
        elif t == "AttributeEnd" or t == "TransformEnd":
            ctm_box[0] = stack.pop()
        elif t == "Identity":
            ctm_box[0] = np.eye(4)
        elif t == "Translate":
            ctm_box[0] = ctm_box[0] @ _translate(*floats(3))
        elif t == "Scale":
            ctm_box[0] = ctm_box[0] @ _scale(*floats(3))
        elif t == "Rotate":
            ctm_box[0] = ctm_box[0] @ _rotate(*floats(4))
        elif t == "LookAt":
            v = floats(9)
            sc.camera["lookat"] = v
            ctm_box[0] = ctm_box[0] @ _lookat(v[0:3], v[3:6], v[6:9])
        elif t in ("Transform", "ConcatTransform"):
            assert toks[i] == "["
            i += 1
            m = np.array(floats(16)).reshape(4, 4).T
            assert toks[i] == "]"
            i += 1
            ctm_box[0] = m if t == "Transform" else ctm_box[0] @ m
        elif t == "Camera":
            kind = toks[i].strip('"')
            i += 1
            p = params()
            sc.camera["type"] = kind
            sc.camera["fov"] = float(p.get("fov", ("float", ["90"]))[1][0])
            sc.camera["camera_from_world"] = ctm_box[0].copy()
        elif t == "Film":
            i += 1
            p = params()
            sc.camera["xres"] = int(p.get("xresolution", ("integer", ["1280"]))[1][0])
            sc.camera["yres"] = int(p.get("yresolution", ("integer", ["720"]))[1][0])
        elif t == "WorldBegin":
            ctm_box[0] = np.eye(4)
        elif t == "Include":
            inc = toks[i].strip('"')
            i += 1
            _walk(os.path.join(base, inc), sc, ctm_box, stack, subdivide)
        elif t == "Shape":
            kind = toks[i].strip('"')
            i += 1
            p = params()
            verts = idx = None
            if kind == "plymesh":
                verts, idx = read_ply(os.path.join(base, p["filename"][1][0].strip('"')))
            elif kind in ("trianglemesh", "loopsubdiv"):
                verts = np.array([float(x) for x in p["P"][1]], dtype=np.float32).reshape(-1, 3)
                idx = np.array([int(x) for x in p["indices"][1]], dtype=np.int32).reshape(-1, 3)
                if kind == "loopsubdiv":
                    levels = int(p.get("levels", ("integer", ["3"]))[1][0])
                    if subdivide is None:
                        raise RuntimeError("loopsubdiv shape needs a subdivide callback")
                    fi, verts = subdivide(levels, idx, verts)
                    idx = np.asarray(fi, dtype=np.int32).reshape(-1, 3)
            elif kind == "sphere":
                # Sphere::Create (shapes.cpp): radius 1, zmin -radius, zmax radius, phimax 360 by default
                f1 = lambda k, d: float(p[k][1][0]) if k in p else d
                radius = f1("radius", 1.0)
                sc.spheres.append({"render_from_object": ctm_box[0][:3, :].tolist(), "radius": radius,
                                   "zmin": f1("zmin", -radius), "zmax": f1("zmax", radius),
                                   "phimax": f1("phimax", 360.0)})
                sc.skipped[kind] = sc.skipped.get(kind, 0) + 1   # not part of the triangle soup
            else:
                sc.skipped[kind] = sc.skipped.get(kind, 0) + 1
            if verts is not None and len(idx):
                m = ctm_box[0]
                v = verts.astype(np.float64) @ m[:3, :3].T + m[:3, 3]
                v = v.astype(np.float32)
                sc.tris.append(v[idx].reshape(-1, 9))
        elif is_directive(t):
            # Material, Texture, LightSource, MakeNamedMaterial, ... : skip type strings + params
            while i < n and toks[i].startswith('"') and " " not in toks[i].strip('"'):
                i += 1
            params()
        # stray tokens (values of skipped directives) are ignored
