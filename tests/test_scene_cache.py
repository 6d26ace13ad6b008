"""tests/test_scene_cache.py -- the BASELINE.json scenes must reach the GPU box.

scene_cache/ is git-ignored and travels with the working tree like the built .so files;
__graft_entry__.build() exports it from the reference's shipped scenes.  Where the reference exists
(this container) the export must be complete, so that bench.py and the BASELINE-config GPU tests run on
the real scenes, never on a stand-in (round-1 VERDICT item 1).
"""
import json
import os

import numpy as np
import pytest

from conftest import ROOT

TRIANGLES = {"killeroos": 66532, "coffee_maker": 235271, "bathroom": 592188, "crown": 3540214}


@pytest.mark.skipif(not os.path.isdir("/root/reference/scenes"), reason="no /root/reference on this machine")
def test_build_exports_every_baseline_scene():
    import __graft_entry__
    __graft_entry__.export_scenes()
    for name, T in TRIANGLES.items():
        path = os.path.join(ROOT, "scene_cache", name + ".tris.npy")
        assert os.path.exists(path), path
        assert np.load(path, mmap_mode="r").shape == (T, 9)
        with open(os.path.join(ROOT, "scene_cache", name + ".json")) as f:
            assert json.load(f)["triangles"] == T


def test_missing_scene_is_an_error_not_a_stand_in(tmp_path, monkeypatch):
    from nn_bvh_b200 import scenes
    monkeypatch.setattr(scenes, "CACHE", str(tmp_path))
    with pytest.raises(FileNotFoundError):
        scenes.load("bathroom")
    tris, cam, name = scenes.load("bathroom", allow_procedural=True)
    assert name.startswith("procedural_room(stand-in")


def test_scene_cache_is_not_gpurun_ignored():
    with open(os.path.join(ROOT, ".gpurunignore")) as f:
        live = [l.strip() for l in f if l.strip() and not l.startswith("#")]
    assert not any(l.startswith("scene_cache") for l in live), live
