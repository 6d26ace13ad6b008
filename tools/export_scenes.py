#!/usr/bin/env python3
"""tools/export_scenes.py -- turn the reference's shipped scenes into triangle soups.

Runs only where /root/reference exists (this container).  Writes
scene_cache/<name>.tris.npy ((T,9) float32, world space) and <name>.json
(camera, counts).  scene_cache/ is git-ignored but travels to the GPU box with
the gpurun snapshot, exactly like built .so files; __graft_entry__.build() calls
main() for every cache file that is missing.

killeroos needs loop subdivision: it is done by the reference's own
LoopSubdivide through oracle/_ref (tools may use the oracle; the product may not).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import pbrt_scene  # noqa: E402

REF_SCENES = "/root/reference/scenes"
SCENES = {
    "killeroos": "killeroos/killeroo-simple-v4.pbrt",
    "coffee_maker": "coffee_maker/scene-v4.pbrt",
    "bathroom": "bathroom/scene-v4.pbrt",
    "crown": "crown/crown.pbrt",
}


def main(names):
    from oracle.pyoracle import Reference
    ref = Reference()
    out = os.path.join(ROOT, "scene_cache")
    os.makedirs(out, exist_ok=True)
    for name in names:
        sc = pbrt_scene.load(os.path.join(REF_SCENES, SCENES[name]), subdivide=ref.loop_subdivide)
        tris = sc.soup() + np.float32(0.0)   # -0.0 -> +0.0, as pbrt's mesh transform does (util/mesh.cpp:36-39)
        np.save(os.path.join(out, name + ".tris.npy"), tris)
        cam = dict(sc.camera)
        if "camera_from_world" in cam:
            cam["camera_from_world"] = np.asarray(cam["camera_from_world"]).tolist()
        meta = {"source": SCENES[name], "triangles": int(tris.shape[0]),
                "skipped_shapes": sc.skipped, "camera": cam, "spheres": sc.spheres}
        with open(os.path.join(out, name + ".json"), "w") as f:
            json.dump(meta, f, indent=1)
        print(name, tris.shape, sc.skipped)


if __name__ == "__main__":
    main(sys.argv[1:] or list(SCENES))
