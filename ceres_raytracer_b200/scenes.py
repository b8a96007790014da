"""Synthetic scenes for the BVH memory / L2 stress configuration (BASELINE.json configs[4]).

instanced_grid(pod, nx, ny, nz) replicates a reference-built scene (the POD arrays of bvh::Bvh<float>, Triangle<float>,
tri_norms) on an nx x ny x nz grid, deterministic, no RNG:

  * triangles   every vertex (p0, p1() = p0 - e1, p2() = p0 + e2; triangle.hpp:36-37) is moved by the instance offset
                in binary32 and the Triangle is rebuilt as its constructor does (e1 = p0 - p1, e2 = p2 - p0,
                n = cross(e1, e2); triangle.hpp:30-34);
  * BVH         ONE flat tree in the reference's node layout (bvh.hpp:25-30, siblings adjacent, root at 0): a balanced
                top tree over the instances (median split of the grid cells along the longest axis) whose leaves are
                copies of the scene's own tree with shifted child / slot indices;
  * bounds      left as placeholders: the boxes of ALL nodes come from a refit (ceres_scene_refit on the GPU, or
                the checker's CPU refit in tests), exactly as HierarchyRefitter would produce them for the moved triangles.

Scene synthesis is host-side input preparation, outside the render path (the reference builds its scene on the host
too, anim.cpp:38-65)."""
import numpy as np

from .render import Camera, NODE_DT


def _top_tree(cells):
    """Balanced binary tree over the instance cells.  Returns (first_child[2K-1] (0 = leaf), leaf_instance[2K-1])."""
    K = len(cells)
    first_child = np.zeros(2 * K - 1, np.int64)
    leaf_inst = np.full(2 * K - 1, -1, np.int64)
    nxt = 1
    work = [(0, np.arange(K))]
    while work:
        node, ids = work.pop()
        if len(ids) == 1:
            leaf_inst[node] = ids[0]
            continue
        c = cells[ids]
        axis = int(np.argmax(c.max(axis=0) - c.min(axis=0)))
        order = ids[np.argsort(c[:, axis], kind="stable")]
        half = len(order) // 2
        first_child[node] = nxt
        work.append((nxt + 1, order[half:]))
        work.append((nxt, order[:half]))
        nxt += 2
    return first_child, leaf_inst


def instanced_grid(pod, nx, ny, nz, spacing=1.25):
    """Returns (pod of the grid scene with placeholder bounds, info dict)."""
    nodes = np.ascontiguousarray(pod["nodes"])
    if nodes.dtype != NODE_DT:
        nodes = nodes.view(NODE_DT).reshape(-1)
    tris = np.ascontiguousarray(pod["triangles"], np.float32).reshape(-1, 12)
    norms = np.ascontiguousarray(pod["norms"], np.float32).reshape(-1, 9)
    idx = np.ascontiguousarray(pod["prim_indices"], np.uint64)
    Nn, T = nodes.shape[0], tris.shape[0]
    if nodes[0]["primitive_count"] != 0:
        raise ValueError("the base scene's root must be an inner node")
    K = nx * ny * nz
    cells = np.stack(np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij"), -1).reshape(-1, 3)
    b = nodes[0]["bounds"].astype(np.float32)
    ext = np.array([b[1] - b[0], b[3] - b[2], b[5] - b[4]], np.float32)
    offsets = (cells.astype(np.float32) * (np.float32(spacing) * ext)).astype(np.float32)        # K x 3

    # ---- nodes: top tree (2K-1 nodes) + K copies of the base tree without its root ------------------------------
    top_fc, top_leaf = _top_tree(cells)
    n_top = 2 * K - 1
    out = np.zeros(n_top + K * (Nn - 1), NODE_DT)
    base_of = n_top + np.arange(K, dtype=np.int64) * (Nn - 1)            # old node i >= 1 of instance k -> base_of[k] + i - 1
    inner_top = top_leaf < 0
    out["first_child_or_primitive"][:n_top][inner_top] = top_fc[inner_top]
    k_leaf = top_leaf[~inner_top]
    out["first_child_or_primitive"][:n_top][~inner_top] = base_of[k_leaf] + int(nodes[0]["first_child_or_primitive"]) - 1
    body = nodes[1:]
    is_leaf = body["primitive_count"] != 0
    first = body["first_child_or_primitive"].astype(np.int64)
    for k in range(K):
        dst = out[base_of[k]: base_of[k] + Nn - 1]
        dst["primitive_count"] = body["primitive_count"]
        dst["first_child_or_primitive"] = np.where(is_leaf, first + k * T, first + base_of[k] - 1)
    # placeholder boxes (never used for rendering: refit first)
    out["bounds"][:] = np.array([0, 0, 0, 0, 0, 0], np.float32)

    # ---- triangles, normals, indices ---------------------------------------------------------------------------
    p0 = tris[:, 0:3]
    p1 = p0 - tris[:, 3:6]
    p2 = p0 + tris[:, 6:9]
    big = np.empty((K * T, 12), np.float32)
    for k in range(K):
        o = offsets[k]
        q0, q1, q2 = p0 + o, p1 + o, p2 + o
        e1, e2 = q0 - q1, q2 - q0
        dst = big[k * T:(k + 1) * T]
        dst[:, 0:3], dst[:, 3:6], dst[:, 6:9] = q0, e1, e2
        dst[:, 9] = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
        dst[:, 10] = e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2]
        dst[:, 11] = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
    big_idx = (idx[None, :] + (np.arange(K, dtype=np.uint64) * np.uint64(T))[:, None]).reshape(-1)
    big_norms = np.tile(norms, (K, 1))
    lo = np.array([b[0], b[2], b[4]], np.float32)
    info = {"instances": K, "grid": (nx, ny, nz), "extent": ext, "origin": lo,
            "span": (np.array([nx - 1, ny - 1, nz - 1], np.float32) * np.float32(spacing) + 1) * ext}
    return {"nodes": out, "prim_indices": big_idx, "triangles": big, "norms": big_norms}, info


def grid_camera(info):
    """camera on the -y side at 1.1 * (xspan / 2) / tan(30 deg) from the front face, centred on
    the grid in x and z, looking +y, up +z, fov 60; sun = (-50, -20, 0) * 10."""
    lo, span = info["origin"].astype(np.float64), info["span"].astype(np.float64)
    dist = 1.1 * (span[0] / 2) / np.tan(np.radians(30.0))
    eye = (float(lo[0] + span[0] / 2), float(lo[1] - dist), float(lo[2] + span[2] / 2))
    return Camera(eye=eye, dir=(0.0, 1.0, 0.0), up=(0.0, 0.0, 1.0), fov=60.0), (-500.0, -200.0, 0.0)
