"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product package.

Edge-list builders, PyG-style collate and seeded synthetic SiO2-like point clouds (numpy, CPU).

Follows (reference file:line):
  * parts/train_per_iretation.py:308-313, make_dataset.py:131-136 -- fully connected directed edges, row-major
    ``for i: for j: if i != j`` ; int64 [2,E]
  * something.ipynb:6340-6365 -- the only EDGE cutoff in the reference: same double loop, strict ``<`` on
    ``torch.norm(pos[i]-pos[j])`` in fp32, open boundaries, fallback to the closest pair when no edge exists
  * make_dataset.py:79-92 -- 27-image replication, the model for the periodic builder.  Periodic minimum image
    is NOT in the reference: it is a new feature and this file DEFINES its semantics:
        d_k = fl( dx_k - fl( L_k * rint( fl(dx_k / L_k) ) ) )   (fp32, round-half-even, no FMA contraction)
  * PyG DataLoader collate (third-party, version unpinned): concatenate node tensors, offset edge_index by the
    cumulative node count, ``batch`` = graph id per node.
fp32 distance (pinned against torch.norm on CPU in make_golden.py -- ATen accumulates with fused multiply-add):
        dist = sqrtf( fmaf(dz, dz, fmaf(dy, dy, fl(dx*dx))) )
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def fc_edges(n, offset=0):
    """Row-major fully connected directed edges of one n-atom graph; int64 [2, n(n-1)]."""
    i = np.repeat(np.arange(n, dtype=np.int64), n)
    j = np.tile(np.arange(n, dtype=np.int64), n)
    keep = i != j
    return np.stack((i[keep], j[keep])) + offset


def fc_edges_batch(sizes):
    out, off = [], 0
    for n in sizes:
        out.append(fc_edges(n, off))
        off += n
    return np.concatenate(out, axis=1) if out else np.zeros((2, 0), np.int64)


def _fma32(a, b, c):
    """Correctly rounded fp32 fused multiply-add RN32(a*b+c) for float32 arrays (numpy has no fmaf).

    a*b is exact in float64; s=RN64(p+c) with the TwoSum error term e gives the exact sum s+e.  RN32(s) can
    only be wrong when s sits exactly on a midpoint of the fp32 grid, where the sign of e breaks the tie.
    """
    p = a.astype(np.float64) * b.astype(np.float64)
    c = c.astype(np.float64)
    s = p + c
    bb = s - p
    e = (p - (s - bb)) + (c - bb)
    r = s.astype(F32)
    lo = np.where(r.astype(np.float64) > s, np.nextafter(r, F32(-np.inf)), r)
    hi = np.nextafter(lo, F32(np.inf))
    tie = (lo.astype(np.float64) + hi.astype(np.float64)) == 2.0 * s
    fix = tie & (e != 0) & np.isfinite(s)
    return np.where(fix, np.where(e > 0, hi, lo), r).astype(F32)


def _dist_f32(d):
    """d: float32 [...,3] displacement -> float32 distance with the pinned operation order."""
    d = np.asarray(d, dtype=F32)
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    s = _fma32(z, z, _fma32(y, y, (x * x).astype(F32)))
    return np.sqrt(s, dtype=F32)


def min_image(d, box):
    """Orthorhombic minimum image of float32 displacements d [...,3] for box lengths box [3] (fp32 ops)."""
    d = d.astype(F32)
    box = np.asarray(box, dtype=F32)
    q = (d / box).astype(F32)
    return (d - (box * np.rint(q).astype(F32)).astype(F32)).astype(F32)


def cutoff_edges(pos, r_c, box=None, offset=0, fallback=True):
    """Directed edges (i,j), i!=j, dist(i,j) < r_c (strict), lexicographic order; int64 [2,E].

    box=None: open boundaries (something.ipynb:6340-6365).  box=[Lx,Ly,Lz]: periodic minimum image.
    fallback: when no pair is inside the cutoff, emit the single closest ordered pair found by the
    reference's scan (first strict minimum in row-major order among pairs with dist > r_c).
    """
    pos = np.asarray(pos, dtype=F32)
    n = pos.shape[0]
    d = (pos[:, None, :] - pos[None, :, :]).astype(F32)
    if box is not None:
        d = min_image(d, box)
    dist = _dist_f32(d)
    mask = dist < F32(r_c)
    np.fill_diagonal(mask, False)
    i, j = np.nonzero(mask)  # row-major == lexicographic
    if i.size == 0 and fallback and n > 1:
        cand = np.where(dist > F32(r_c), dist, F32(np.inf))
        np.fill_diagonal(cand, F32(np.inf))
        cand = np.minimum(cand, F32(100.0))  # reference starts from min_norm = 100
        k = int(np.argmin(cand))  # first occurrence of the minimum in row-major order
        if cand.flat[k] < F32(100.0):
            i, j = np.array([k // n]), np.array([k % n])
    return np.stack((i.astype(np.int64), j.astype(np.int64))) + offset


def cutoff_edges_batch(pos, graph_ptr, r_c, boxes=None):
    out = []
    for g in range(len(graph_ptr) - 1):
        a, b = int(graph_ptr[g]), int(graph_ptr[g + 1])
        box = None if boxes is None else boxes[g]
        out.append(cutoff_edges(pos[a:b], r_c, box, offset=a, fallback=False))
    return np.concatenate(out, axis=1) if out else np.zeros((2, 0), np.int64)


def collate(graphs):
    """PyG-DataLoader-style batching of dicts {x:[n,A], pos:[n,3], edge_index:[2,e]} (numpy)."""
    xs, ps, es, bs, off = [], [], [], [], 0
    for g, d in enumerate(graphs):
        n = d["pos"].shape[0]
        xs.append(d["x"]); ps.append(d["pos"]); es.append(d["edge_index"] + off)
        bs.append(np.full(n, g, np.int64)); off += n
    return dict(x=np.concatenate(xs), pos=np.concatenate(ps), edge_index=np.concatenate(es, axis=1),
                batch=np.concatenate(bs), num_graphs=len(graphs))


# ---- synthetic SiO2-like clouds -------------------------------------------------------
SIO2_DENSITY = 0.0662  # atoms / A^3


def sio2_cell(n, rng, min_dist=1.4):
    """n atoms (n/3 Si, 2n/3 O) uniform in a cubic cell of side (n/0.0662)^(1/3) A with min-distance rejection.

    Returns (pos float32 [n,3] shifted to zero COM, onehot float32 [n,2] with [1,0]=O,[0,1]=Si
    (make_dataset.py:112-117), side).  Rejection uses periodic minimum image so the cell tiles cleanly.
    """
    side = float((n / SIO2_DENSITY) ** (1.0 / 3.0))
    pts = np.zeros((n, 3), np.float64)
    k = 0
    while k < n:
        c = rng.uniform(0.0, side, size=3)
        if k:
            d = pts[:k] - c
            d -= side * np.rint(d / side)
            if np.min(np.einsum("ij,ij->i", d, d)) < min_dist * min_dist:
                continue
        pts[k] = c
        k += 1
    pos = (pts - pts.mean(axis=0)).astype(F32)
    onehot = np.zeros((n, 2), F32)
    n_si = n // 3
    species = np.zeros(n, np.int64)
    species[rng.permutation(n)[:n_si]] = 1
    onehot[np.arange(n), species] = 1.0
    return pos, onehot, side


def make_batch(num_graphs, n_atoms, seed, edges="fc", r_c=5.0):
    """Seeded synthetic batch for configs 1-4: dict(x,pos,edge_index,batch,num_graphs,graph_ptr,boxes)."""
    rng = np.random.default_rng(seed)
    graphs, boxes = [], []
    for _ in range(num_graphs):
        pos, onehot, side = sio2_cell(n_atoms, rng)
        box = np.array([side] * 3, F32)
        if edges == "fc":
            ei = fc_edges(n_atoms)
        elif edges == "cutoff_open":
            ei = cutoff_edges(pos, r_c, None)
        elif edges == "cutoff_periodic":
            ei = cutoff_edges(pos, r_c, box)
        else:
            raise ValueError(edges)
        graphs.append(dict(x=onehot, pos=pos, edge_index=ei))
        boxes.append(box)
    b = collate(graphs)
    b["graph_ptr"] = np.arange(num_graphs + 1, dtype=np.int64) * n_atoms
    b["boxes"] = np.stack(boxes)
    return b
