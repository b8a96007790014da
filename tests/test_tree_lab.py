"""profiles/tree_lab.py -- the CPU laboratory that picked the device tree builder -- kept honest: the trees its numpy builders
emit are valid trees in the reference's layout, the oracle's refit reproduces their boxes bit for bit, and the oracle's frame
from them equals its frame from the reference-built tree in everything a tie between two triangles cannot change.  (Also a
size-independent property of the oracle itself: the answer must not depend on which valid tree it walks.)"""
import importlib.util
import os

import numpy as np
import pytest

from conftest import ROOT
from oracle import port


@pytest.fixture(scope="module")
def lab():
    spec = importlib.util.spec_from_file_location("tree_lab", os.path.join(ROOT, "profiles", "tree_lab.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def valid(nodes, idx, n):
    assert nodes.shape[0] % 2 == 1 and sorted(idx.tolist()) == list(range(n))
    seen = np.zeros(nodes.shape[0], np.int32)
    slots = np.zeros(n, np.int32)
    stack = [0]
    while stack:
        i = stack.pop()
        seen[i] += 1
        nd = nodes[i]
        if nd["primitive_count"]:
            slots[nd["first_child_or_primitive"]: nd["first_child_or_primitive"] + nd["primitive_count"]] += 1
        else:
            c = int(nd["first_child_or_primitive"])
            stack += [c, c + 1]
    assert (seen == 1).all() and (slots == 1).all()


@pytest.mark.parametrize("variant", ["base", "bin100000+ad16/4", "top512"])
def test_lab_trees_are_valid_and_render_like_the_reference_tree(lab, bunny_pod, variant):
    import ceres_raytracer_b200.configs as configs
    n = bunny_pod["triangles"].shape[0]
    nodes, prim, sah = lab.build_variant(variant, bunny_pod["triangles"])
    valid(nodes, prim, n)
    pod = dict(bunny_pod)
    pod["nodes"], pod["prim_indices"] = nodes, prim
    assert np.array_equal(port.refit(pod).view(np.uint32), nodes.view(np.uint32))
    cam, sun = configs.BUNNY_CAMERA.as_array(), np.array([-50, -20, 0], np.float32)
    a = port.render(bunny_pod, cam, sun, 240, 135, 0)
    b = port.render(pod, cam, sun, 240, 135, 0)
    assert (a["stats"]["rays"], a["stats"]["hits"]) == (b["stats"]["rays"], b["stats"]["hits"])
    assert np.array_equal(a["state"], b["state"])
    same = a["ids"] == b["ids"]
    assert same.mean() > 0.999
    assert np.array_equal(a["pixels"][same].view(np.uint32), b["pixels"][same].view(np.uint32))
    assert sah > 0
