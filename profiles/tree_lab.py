#!/usr/bin/env python
"""CPU laboratory for the topology of the GPU-built tree: a numpy restatement of the device
builder (Morton order -> locally-ordered clustering -> SAH leaf collapsing, lbvh_build.cuh) plus candidate improvements, judged
by what the ORACLE's traverser does on the resulting tree: node-pair steps S and triangle tests T of the dragon frame
(single_ray_traverser.hpp:158-163 statistics), next to the reference's SweepSah tree.  No GPU needed: the ideas that lower
S and T here are the ones worth a device implementation and GPU minutes.

  python profiles/tree_lab.py [--size 1280 720] [--variants base,top4096,...]
"""
import argparse, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import port

NODE_DT = np.dtype([("bounds", "<f4", (6,)), ("primitive_count", "<u4"), ("first_child_or_primitive", "<u4")])


def tri_boxes(tris12):
    p0 = tris12[:, 0:3]; p1 = p0 - tris12[:, 3:6]; p2 = p0 + tris12[:, 6:9]
    lo = np.minimum(np.minimum(p0, p1), p2); hi = np.maximum(np.maximum(p0, p1), p2)
    b = np.empty((len(tris12), 6), np.float32)
    b[:, 0::2] = lo; b[:, 1::2] = hi
    return b


def spread(v, bits):
    out = np.zeros_like(v, dtype=np.uint64)
    for k in range(bits):
        out |= ((v >> np.uint64(k)) & np.uint64(1)) << np.uint64(3 * k)
    return out


def morton_order(box, bits=10):
    c = 0.5 * (box[:, 0::2] + box[:, 1::2])
    lo, hi = c.min(0), c.max(0)
    n = float(1 << bits)
    q = np.clip((c - lo) / np.where(hi > lo, hi - lo, 1) * n, 0, n - 1).astype(np.uint64)
    code = (spread(q[:, 0], bits) << np.uint64(2)) | (spread(q[:, 1], bits) << np.uint64(1)) | spread(q[:, 2], bits)
    return np.argsort(code, kind="stable")


def half_area(b):
    d = b[..., 1::2] - b[..., 0::2]
    return d[..., 0] * d[..., 1] + d[..., 1] * d[..., 2] + d[..., 2] * d[..., 0]


def union(a, b):
    u = np.empty_like(a)
    u[..., 0::2] = np.minimum(a[..., 0::2], b[..., 0::2]); u[..., 1::2] = np.maximum(a[..., 1::2], b[..., 1::2])
    return u


class Tree:
    """Binary tree over n leaves: node ids 0..n-1 are the leaves (sorted position), n.. are inner nodes."""
    def __init__(self, leaf_box, leaf_tri):
        self.n = len(leaf_box)
        self.box = [None] * 0
        self.boxes = np.zeros((2 * self.n - 1, 6), np.float32)
        self.boxes[: self.n] = leaf_box
        self.left = np.full(2 * self.n - 1, -1, np.int64)
        self.right = np.full(2 * self.n - 1, -1, np.int64)
        self.next = self.n
        self.leaf_tri = leaf_tri
        self.root = -1

    def new_nodes(self, L, R):
        k = len(L)
        ids = np.arange(self.next, self.next + k)
        self.left[ids] = L; self.right[ids] = R
        self.boxes[ids] = union(self.boxes[L], self.boxes[R])
        self.next += k
        return ids


def ploc(tree, cid, radius=8, stop_at=1):
    """Rounds of locally-ordered clustering over the cluster array `cid` (node ids, in Morton order) until stop_at are left."""
    while len(cid) > stop_at:
        m = len(cid)
        cb = tree.boxes[cid]
        best = np.full(m, np.inf, np.float32); bj = np.arange(m)
        for d in range(1, radius + 1):
            if d >= m:
                break
            a = half_area(union(cb[:-d], cb[d:]))
            # candidate j = i + d for i, and j = i - d for i + d; ties go to the nearer position, then to i ^ 1
            for (idx, jdx) in ((np.arange(m - d), np.arange(d, m)), (np.arange(d, m), np.arange(m - d))):
                better = (a < best[idx]) | ((a == best[idx]) & (np.abs(bj[idx] - idx) == d) & (jdx == (idx ^ 1)))
                best[idx] = np.where(better, a, best[idx]); bj[idx] = np.where(better, jdx, bj[idx])
        i = np.arange(m)
        mutual = (bj != i) & (bj[bj] == i)
        lower = mutual & (i < bj)
        gone = mutual & (i > bj)
        if not lower.any():
            raise RuntimeError("no merge")
        new = tree.new_nodes(cid[lower], cid[bj[lower]])
        out = cid.copy()
        out[lower] = new
        cid = out[~gone]
    return cid


def sah_top(tree, cid, weight, max_leaf=1):
    """Top-down full-sweep SAH over the clusters (their boxes as primitives, weighted by their triangle counts), as
    sweep_sah_builder.hpp does over triangles: the tree above the clusters."""
    cb = tree.boxes[cid]
    cen = 0.5 * (cb[:, 0::2] + cb[:, 1::2])

    def build(ix):
        if len(ix) == 1:
            return int(cid[ix[0]])
        best = (np.inf, None, None)
        for ax in range(3):
            o = ix[np.argsort(cen[ix, ax], kind="stable")]
            b = cb[o]; w = weight[o].astype(np.float64)
            lo_l = np.minimum.accumulate(b[:, 0::2], axis=0); hi_l = np.maximum.accumulate(b[:, 1::2], axis=0)
            lo_r = np.minimum.accumulate(b[::-1, 0::2], axis=0)[::-1]; hi_r = np.maximum.accumulate(b[::-1, 1::2], axis=0)[::-1]
            dl = hi_l - lo_l; dr = hi_r - lo_r
            al = dl[:, 0] * dl[:, 1] + dl[:, 1] * dl[:, 2] + dl[:, 2] * dl[:, 0]
            ar = dr[:, 0] * dr[:, 1] + dr[:, 1] * dr[:, 2] + dr[:, 2] * dr[:, 0]
            cw = np.cumsum(w)
            cost = al[:-1] * cw[:-1] + ar[1:] * (cw[-1] - cw[:-1])          # split after position k
            k = int(np.argmin(cost))
            if cost[k] < best[0]:
                best = (cost[k], o, k)
        _, o, k = best
        L = build(o[: k + 1]); R = build(o[k + 1:])
        return int(tree.new_nodes(np.array([L]), np.array([R]))[0])

    sys.setrecursionlimit(100000)
    return build(np.arange(len(cid)))


def binned_top2(tree, cid, weight, nbins=16, longest=False, loose=False):
    """binned_top with the simplifications a cheap device pass would make.  longest: only the axis of the largest centroid
    extent is binned.  loose: a child does not recompute its centroid bounds -- it inherits the parent's, cut at the split plane."""
    cb = tree.boxes[cid]
    cen = (0.5 * (cb[:, 0::2] + cb[:, 1::2])).astype(np.float32)

    nb_big = nbins

    def build(ix, lo, hi):
        if len(ix) == 1:
            return int(cid[ix[0]])
        nbins = nb_big if len(ix) > ADAPT[0] else ADAPT[1]
        c = cen[ix]
        if not loose or lo is None:
            lo = c.min(0); hi = c.max(0)
        axes = [int(np.argmax(hi - lo))] if longest else [0, 1, 2]
        best = (np.inf, None)
        for ax in axes:
            ext = hi[ax] - lo[ax]
            if not ext > 0:
                continue
            b = np.clip((nbins * (c[:, ax] - lo[ax]) / ext).astype(np.int64), 0, nbins - 1)
            bl = np.full((nbins, 3), np.inf, np.float32); bh = np.full((nbins, 3), -np.inf, np.float32); bw = np.zeros(nbins)
            np.minimum.at(bl, b, cb[ix][:, 0::2]); np.maximum.at(bh, b, cb[ix][:, 1::2]); np.add.at(bw, b, weight[ix])
            ll = np.minimum.accumulate(bl, 0); lh = np.maximum.accumulate(bh, 0)
            rl = np.minimum.accumulate(bl[::-1], 0)[::-1]; rh = np.maximum.accumulate(bh[::-1], 0)[::-1]
            cw = np.cumsum(bw)
            for k in range(nbins - 1):
                if cw[k] == 0 or cw[-1] - cw[k] == 0:
                    continue
                dl = lh[k] - ll[k]; dr = rh[k + 1] - rl[k + 1]
                cost = (dl[0] * dl[1] + dl[1] * dl[2] + dl[2] * dl[0]) * cw[k] + (dr[0] * dr[1] + dr[1] * dr[2] + dr[2] * dr[0]) * (cw[-1] - cw[k])
                if cost < best[0]:
                    best = (cost, (ax, k))
        if best[1] is None:
            if loose:                                         # the inherited range may hold every centroid in one bin: tighten once
                lo2 = c.min(0); hi2 = c.max(0)
                if (hi2 > lo2).any() and not (np.array_equal(lo2, lo) and np.array_equal(hi2, hi)):
                    return build(ix, lo2, hi2) if False else build_tight(ix)
            h = len(ix) // 2
            left, right, llo, lhi, rlo, rhi = ix[:h], ix[h:], None, None, None, None
        else:
            ax, k = best[1]
            ext = hi[ax] - lo[ax]
            b = np.clip((nbins * (c[:, ax] - lo[ax]) / ext).astype(np.int64), 0, nbins - 1)
            left, right = ix[b <= k], ix[b > k]
            plane = lo[ax] + ext * (k + 1) / nbins
            llo, lhi, rlo, rhi = lo.copy(), hi.copy(), lo.copy(), hi.copy()
            lhi[ax] = plane; rlo[ax] = plane
        L = build(left, llo, lhi); R = build(right, rlo, rhi)
        return int(tree.new_nodes(np.array([L]), np.array([R]))[0])

    def build_tight(ix):
        c = cen[ix]
        # one level with exact bounds, then loose again below
        save = loose
        return build(ix, None, None)

    sys.setrecursionlimit(100000)
    return build(np.arange(len(cid)), None, None)


ADAPT = [0, 16]          # tasks of at most ADAPT[0] members use ADAPT[1] bins per axis


def binned_top(tree, cid, weight, nbins=16):
    """Top-down BINNED SAH over the clusters (nbins bins per axis over the node's centroid bounds): what a level-synchronous
    device builder would do.  A node whose centroids coincide is split in the middle of its (Morton-ordered) member list."""
    cb = tree.boxes[cid]
    cen = (0.5 * (cb[:, 0::2] + cb[:, 1::2])).astype(np.float32)

    nb_big = nbins

    def build(ix):
        if len(ix) == 1:
            return int(cid[ix[0]])
        nbins = nb_big if len(ix) > ADAPT[0] else ADAPT[1]
        c = cen[ix]; lo = c.min(0); hi = c.max(0)
        if nbins == 2 and (hi > lo).any():
            # "2 bins" = no SAH evaluation at all: cut the longest axis of the centroid bounds in the middle
            ax = int(np.argmax(hi - lo))
            mid = lo[ax] + 0.5 * (hi[ax] - lo[ax])
            left, right = ix[c[:, ax] < mid], ix[c[:, ax] >= mid]
            if len(left) and len(right):
                L = build(left); R = build(right)
                return int(tree.new_nodes(np.array([L]), np.array([R]))[0])
        best = (np.inf, None)
        for ax in range(3):
            ext = hi[ax] - lo[ax]
            if not ext > 0:
                continue
            b = np.minimum((nbins * (c[:, ax] - lo[ax]) / ext).astype(np.int64), nbins - 1)
            bl = np.full((nbins, 3), np.inf, np.float32); bh = np.full((nbins, 3), -np.inf, np.float32); bw = np.zeros(nbins)
            np.minimum.at(bl, b, cb[ix][:, 0::2]); np.maximum.at(bh, b, cb[ix][:, 1::2]); np.add.at(bw, b, weight[ix])
            ll = np.minimum.accumulate(bl, 0); lh = np.maximum.accumulate(bh, 0)
            rl = np.minimum.accumulate(bl[::-1], 0)[::-1]; rh = np.maximum.accumulate(bh[::-1], 0)[::-1]
            cw = np.cumsum(bw)
            for k in range(nbins - 1):
                if cw[k] == 0 or cw[-1] - cw[k] == 0:
                    continue
                dl = lh[k] - ll[k]; dr = rh[k + 1] - rl[k + 1]
                cost = (dl[0] * dl[1] + dl[1] * dl[2] + dl[2] * dl[0]) * cw[k] + (dr[0] * dr[1] + dr[1] * dr[2] + dr[2] * dr[0]) * (cw[-1] - cw[k])
                if cost < best[0]:
                    best = (cost, (ax, k, b))
        if best[1] is None:
            h = len(ix) // 2
            left, right = ix[:h], ix[h:]
        else:
            ax, k, _ = best[1]
            ext = hi[ax] - lo[ax]
            b = np.minimum((nbins * (c[:, ax] - lo[ax]) / ext).astype(np.int64), nbins - 1)
            left, right = ix[b <= k], ix[b > k]
        L = build(left); R = build(right)
        return int(tree.new_nodes(np.array([L]), np.array([R]))[0])

    sys.setrecursionlimit(100000)
    return build(np.arange(len(cid)))


def range_sah(tree, cid, weight, step=1):
    """Top-down SAH over RANGES of the Morton order: a node is a contiguous run of the sorted clusters and splits at the position
    that minimises area(left) x weight(left) + area(right) x weight(right) -- no primitive ever moves.  step > 1: only every
    step-th position (from the range start) is a candidate."""
    cb = tree.boxes[cid]
    stack = [(0, len(cid), -1, 0)]
    res = {}
    # iterative: children are created after their subtrees, so process in post-order with an explicit stack
    def build(a, b):
        if b - a == 1:
            return int(cid[a])
        seg = cb[a:b]; w = weight[a:b]
        lo_l = np.minimum.accumulate(seg[:, 0::2], axis=0); hi_l = np.maximum.accumulate(seg[:, 1::2], axis=0)
        lo_r = np.minimum.accumulate(seg[::-1, 0::2], axis=0)[::-1]; hi_r = np.maximum.accumulate(seg[::-1, 1::2], axis=0)[::-1]
        dl = hi_l - lo_l; dr = hi_r - lo_r
        al = dl[:, 0] * dl[:, 1] + dl[:, 1] * dl[:, 2] + dl[:, 2] * dl[:, 0]
        ar = dr[:, 0] * dr[:, 1] + dr[:, 1] * dr[:, 2] + dr[:, 2] * dr[:, 0]
        cw = np.cumsum(w)
        cost = al[:-1] * cw[:-1] + ar[1:] * (cw[-1] - cw[:-1])
        if step > 1 and len(cost) > step:
            cand = np.arange(step - 1, len(cost), step)
            k = int(cand[np.argmin(cost[cand])])
        else:
            k = int(np.argmin(cost))
        L = build(a, a + k + 1); R = build(a + k + 1, b)
        return int(tree.new_nodes(np.array([L]), np.array([R]))[0])
    sys.setrecursionlimit(100000)
    return build(0, len(cid))


def subtree_counts(tree, root):
    cnt = np.zeros(tree.next, np.int64)
    cnt[: tree.n] = 1
    order = []
    stack = [root]
    while stack:
        v = stack.pop()
        if v >= tree.n:
            order.append(v); stack.append(tree.left[v]); stack.append(tree.right[v])
    for v in reversed(order):
        cnt[v] = cnt[tree.left[v]] + cnt[tree.right[v]]
    return cnt, order


def subtree_counts_many(tree, cid):
    cnt = np.zeros(tree.next, np.int64)
    cnt[: tree.n] = 1
    for v in range(tree.n, tree.next):                          # children are always created before their parent
        cnt[v] = cnt[tree.left[v]] + cnt[tree.right[v]]
    return cnt[cid].astype(np.float64)


def rotations(tree, root, passes=1):
    """Kensler-style tree rotations, bottom-up: at every inner node try to swap a child with a grandchild on the other side
    when that shrinks the area of the rebuilt child."""
    total = 0
    for _ in range(passes):
        _, order = subtree_counts(tree, root)
        changed = 0
        for v in reversed(order):
            L, R = tree.left[v], tree.right[v]
            best = (0.0, None)
            for (a, b, side) in ((L, R, 0), (R, L, 1)):          # swap `a` with a grandchild under `b`
                if b < tree.n:
                    continue
                bl, br = tree.left[b], tree.right[b]
                area_b = half_area(tree.boxes[b])
                # a <-> bl : b becomes (a, br)
                g1 = area_b - half_area(union(tree.boxes[a], tree.boxes[br]))
                g2 = area_b - half_area(union(tree.boxes[bl], tree.boxes[a]))
                if g1 > best[0]: best = (g1, (a, b, side, 0))
                if g2 > best[0]: best = (g2, (a, b, side, 1))
            if best[1] is not None:
                a, b, side, which = best[1]
                bl, br = tree.left[b], tree.right[b]
                if which == 0:
                    tree.left[b] = a; up = bl
                else:
                    tree.right[b] = a; up = br
                tree.boxes[b] = union(tree.boxes[tree.left[b]], tree.boxes[tree.right[b]])
                if side == 0: tree.left[v] = up
                else: tree.right[v] = up
                changed += 1
        total += changed
    return total


def collapse_and_emit(tree, root, ct=1.0, ci=1.5, max_leaf=8):
    """SAH leaf collapsing (bvhc_cost_kernel's rule) and the reference layout: nodes (root 0, siblings adjacent, pre-order),
    prim_indices in depth-first leaf order."""
    cnt, order = subtree_counts(tree, root)
    area = half_area(tree.boxes[: tree.next]).astype(np.float64)
    cost = np.zeros(tree.next); cost[: tree.n] = area[: tree.n] * ci
    col = np.zeros(tree.next, bool)
    for v in reversed(order):
        as_leaf = area[v] * cnt[v] * ci
        as_tree = area[v] * ct + cost[tree.left[v]] + cost[tree.right[v]]
        col[v] = v != root and cnt[v] <= max_leaf and as_leaf <= as_tree
        cost[v] = as_leaf if col[v] else as_tree
    nodes = np.zeros(2 * tree.n - 1, NODE_DT)
    prim = []
    nodes[0]["bounds"] = tree.boxes[root]

    def leaves_of(v, out):
        st = [v]
        while st:
            x = st.pop()
            if x < tree.n: out.append(int(tree.leaf_tri[x]))
            else: st.append(tree.right[x]); st.append(tree.left[x])

    used = 1
    stack = [(root, 0)]
    while stack:
        v, at = stack.pop()
        if v < tree.n or col[v]:
            nodes[at]["primitive_count"] = cnt[v]; nodes[at]["first_child_or_primitive"] = len(prim)
            leaves_of(v, prim)
            continue
        nodes[at]["primitive_count"] = 0; nodes[at]["first_child_or_primitive"] = used
        l, r = tree.left[v], tree.right[v]
        nodes[used]["bounds"] = tree.boxes[l]; nodes[used + 1]["bounds"] = tree.boxes[r]
        stack.append((r, used + 1)); stack.append((l, used))
        used += 2
    return nodes[:used].copy(), np.asarray(prim, np.uint64), float(cost[root] / area[root])


def build_variant(name, tris12, radius=8):
    box = tri_boxes(tris12)
    opts = set(name.split("+"))
    bits = 21 if "m63" in opts else 10
    order = morton_order(box, bits)
    tree = Tree(box[order], order)
    cid = np.arange(tree.n)
    top = [int(o[3:]) for o in opts if o.startswith("top")]
    ad = [o for o in opts if o.startswith("ad")]               # adN/B: tasks of at most N members use B bins
    ADAPT[:] = [int(ad[0][2:].split("/")[0]), int(ad[0][2:].split("/")[1])] if ad else [0, 16]
    btop = [o for o in opts if o.startswith("bin")]           # binK or binK/B: binned SAH top over K clusters with B bins
    r = [int(o[1:]) for o in opts if o.startswith("r") and o[1:].isdigit()]
    radius = r[0] if r else radius
    mr = [o for o in opts if o.startswith("mr")]               # mrK or mrK/step: SAH over Morton ranges above K clusters
    if mr:
        K, _, B = mr[0][2:].partition("/")
        cid = ploc(tree, cid, radius, stop_at=int(K)) if int(K) < tree.n else cid
        root = range_sah(tree, cid, subtree_counts_many(tree, cid), int(B) if B else 1)
    elif btop:
        K, _, B = btop[0][3:].partition("/")
        cid = ploc(tree, cid, radius, stop_at=int(K)) if int(K) < tree.n else cid
        cnt_all = subtree_counts_many(tree, cid)
        if "unw" in opts:
            cnt_all = np.ones(len(cid))
        if "longest" in opts or "loose" in opts:
            root = binned_top2(tree, cid, cnt_all, int(B) if B else 16, "longest" in opts, "loose" in opts)
        else:
            root = binned_top(tree, cid, cnt_all, int(B) if B else 16)
    elif top:
        cid = ploc(tree, cid, radius, stop_at=top[0])
        cnt, _ = (None, None)
        w = np.array([subtree_counts(tree, c)[0][c] for c in cid])
        root = sah_top(tree, cid, w)
    else:
        root = int(ploc(tree, cid, radius)[0])
    rot = [int(o[3:]) for o in opts if o.startswith("rot")]
    if rot:
        n = rotations(tree, root, rot[0])
        print(f"   [{name}] rotations applied: {n}")
    return collapse_and_emit(tree, root)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, nargs=2, default=[960, 540])
    ap.add_argument("--variants", default="base,m63,top4096,top512,rot2,top4096+rot2")
    a = ap.parse_args()
    import ceres_raytracer_b200.configs as configs
    with np.load(os.path.join(ROOT, "tests", "golden", "dragon_scene.npz")) as z:
        pod = {k: z[k] for k in z.files}
    W, H = a.size
    cam = configs.DRAGON_CAMERA.as_array(); sun = np.array([-50, -20, 0], np.float32)

    ref = [None, None]

    def judge(tag, p, extra=""):
        st = port.render(p, cam, sun, W, H, 0, want_pixels=False)["stats"]
        S = st["S_miss"] + st["S_hit"] + st["S_shadow"]; T = st["T_miss"] + st["T_hit"] + st["T_shadow"]
        if ref[0] is None:
            ref[:] = [S, T]
        print(f"{tag:18s} nodes {len(p['nodes']):6d}  S {S/1e6:7.3f} M x{S/ref[0]:.3f}  T {T/1e6:6.3f} M x{T/ref[1]:.3f}  (S primary {(st['S_miss']+st['S_hit'])/1e6:.3f}, shadow "
              f"{st['S_shadow']/1e6:.3f}) hits {st['hits']} {extra}", flush=True)
        return S, T

    judge("SweepSah (ref)", pod)
    for v in a.variants.split(","):
        t0 = time.time()
        nodes, prim, sah = build_variant(v, pod["triangles"])
        p = dict(pod); p["nodes"] = nodes; p["prim_indices"] = prim
        judge(v, p, f"SAH {sah:.2f} [{time.time()-t0:.0f}s]")


if __name__ == "__main__":
    main()
