"""Drop-ins for the reference's structure-evaluation functions, batched over structures on the device
(SURVEY.md 8f rank 2): ``evaluate_RDF.{length_from_exO, RDF, cos_similarity}``,
``CN2_evaluate.{calculate_angle_for_CN2, calculate_bond_length_for_CN2}``, the site selection of
``evaluate_Si-O-Si.py:28-39`` and ``evaluate_rmsd_for_pos_generate.kabsch_torch``.

The single-structure functions keep the reference's names, arguments and return types; the ``*_batch`` forms take
S structures at once (a [S,n,3] tensor, or a list of [n_s,3] tensors) and launch ONE kernel with one CTA per structure
instead of the reference's python loops (500 bins x n distances per structure, evaluate_RDF.py:53-56).
CUDA tensors only: CPU tensors are moved to the device first; there is no CPU implementation here.
"""
from __future__ import annotations

import numpy as np
import torch

from ._lib import call, ptr, stream_ptr

I32 = torch.int32


def _pack(positions, device="cuda"):
    """-> (pos [N,3] fp32 contiguous on the device, node_ptr int32 [S+1], S)"""
    if isinstance(positions, torch.Tensor) and positions.dim() == 3:
        S, n = positions.shape[0], positions.shape[1]
        pos = positions.to(device=device, dtype=torch.float32).reshape(S * n, 3).contiguous()
        node_ptr = (torch.arange(S + 1, device=pos.device) * n).to(I32)
        return pos, node_ptr, S
    if isinstance(positions, torch.Tensor):
        positions = [positions]
    sizes = [int(p.shape[0]) for p in positions]
    pos = torch.cat([torch.as_tensor(p).to(device=device, dtype=torch.float32).reshape(-1, 3) for p in positions]).contiguous()
    node_ptr = torch.tensor(np.concatenate(([0], np.cumsum(sizes))), dtype=I32, device=pos.device)
    return pos, node_ptr, len(sizes)


def _gaussian_weights(sigma, truncate=4.0):
    """scipy.ndimage.gaussian_filter1d's kernel (order 0): radius = int(truncate*sigma + 0.5), normalised exp(-x^2/2s^2)"""
    sd = float(sigma)
    radius = int(truncate * sd + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sd * sd) * x ** 2)
    return phi / phi.sum(), radius


_TABLES = {}


def _rdf_tables(R, dR, sigma, device):
    key = (float(R), float(dR), float(sigma), str(device))
    tab = _TABLES.get(key)
    if tab is None:
        r = np.arange(0 + dR, R + dR, dR)  # evaluate_RDF.py:53
        lo, hi = r.astype(np.float32), (r + dR).astype(np.float32)  # the reference compares in fp32 (eval_ref.rdf_edges)
        r2 = np.array([float(v ** 2) for v in r])  # scalar powers, as :56 evaluates them
        w, radius = _gaussian_weights(sigma)
        vol = 4 / 3 * np.pi * R ** 3  # (:51)
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        tab = dict(nb=len(r), lo=dev(lo), hi=dev(hi), r2=dev(r2), w=dev(w), radius=radius, vol=float(vol), r=r)
        _TABLES[key] = tab
    return tab


def RDF_batch(positions, sigma=5, R=5.0, dR=0.01, Normalize=False, return_counts=False, return_dists=False):
    """Local RDF around atom 0 of every structure -> float64 tensor [S, nbins] (evaluate_RDF.py:48-60 per row).
    Optionally also the integer histogram [S, nbins] and the distances |p_i - p_0| [N]."""
    pos, node_ptr, S = _pack(positions)
    tab = _rdf_tables(R, dR, sigma, pos.device)
    nb = tab["nb"]
    counts = torch.empty(S, nb, dtype=I32, device=pos.device)
    rdf = torch.empty(S, nb, dtype=torch.float64, device=pos.device)
    dists = torch.empty(pos.shape[0], device=pos.device) if return_dists else None
    call("egnn_rdf_local", ptr(pos), ptr(node_ptr), S, ptr(tab["lo"]), ptr(tab["hi"]), ptr(tab["r2"]), nb, float(dR),
         tab["vol"], ptr(tab["w"]), tab["radius"], ptr(counts), ptr(rdf), ptr(dists, allow_none=True), stream_ptr(pos.device))
    if Normalize:
        rdf = rdf / rdf.max(dim=1, keepdim=True).values
    out = (rdf,)
    if return_counts:
        out += (counts,)
    if return_dists:
        out += (dists,)
    return out[0] if len(out) == 1 else out


def RDF(position, sigma=5, R=5.0, dR=0.01, Normalize=False):
    """evaluate_RDF.py:48-60 for one [n,3] structure -> numpy float64 [nbins] like the reference"""
    return RDF_batch([position], sigma, R, dR, Normalize)[0].cpu().numpy()


def length_from_exO(position):
    """evaluate_RDF.py:39-45: the distances of atoms 1.. from atom 0, as a list of 0-dim tensors"""
    _, d = RDF_batch([position], return_dists=True)
    return list(d[1:].unbind(0))


def cos_similarity(a, b):
    """evaluate_RDF.py:62-63 (two 1-D curves; host arithmetic on a few hundred numbers)"""
    a, b = np.asarray(a), np.asarray(b)
    return np.dot(a, b) / (np.linalg.norm(a) * np.linalg.norm(b))


def cn2_sites(positions, radius=2.0):
    """Site of atom 0 in every structure: dict(count [S] int32, idx [S,2] int32, angle [S] float64 degrees,
    bonds [S,2] fp32).  radius > 0: the reference's rule (evaluate_Si-O-Si.py:28-39; keep rows with count == 2);
    radius <= 0 / None: the two nearest atoms."""
    pos, node_ptr, S = _pack(positions)
    count = torch.empty(S, dtype=I32, device=pos.device)
    idx = torch.empty(S, 2, dtype=I32, device=pos.device)
    angle = torch.empty(S, dtype=torch.float64, device=pos.device)
    bonds = torch.empty(S, 2, device=pos.device)
    call("egnn_cn2_sites", ptr(pos), ptr(node_ptr), S, float(radius) if radius else 0.0, ptr(count), ptr(idx), ptr(angle),
         ptr(bonds), stream_ptr(pos.device))
    return dict(count=count, idx=idx, angle=angle, bonds=bonds)


def calculate_angle_for_CN2(coords_tensor):
    """CN2_evaluate.py:12-16: angle in degrees at row 0 of a [3,3] tensor (python float)"""
    return float(cn2_sites([coords_tensor[:3]], radius=None)["angle"][0])


def calculate_bond_length_for_CN2(coords_tensor):
    """CN2_evaluate.py:18-21 -> (|c1 - c0|, |c2 - c0|)"""
    b = cn2_sites([coords_tensor[:3]], radius=None)["bonds"][0].cpu()
    return b[0].item(), b[1].item()


def kabsch_batch(P, Q):
    """Kabsch alignment per structure -> (R [S,3,3], t [S,3], rmsd [S]) (evaluate_rmsd_for_pos_generate.py:11-51)"""
    p, node_ptr, S = _pack(P)
    q, node_ptr_q, Sq = _pack(Q)
    assert p.shape == q.shape and S == Sq and torch.equal(node_ptr, node_ptr_q), "Matrix dimensions must match"
    R = torch.empty(S, 3, 3, device=p.device)
    t = torch.empty(S, 3, device=p.device)
    rmsd = torch.empty(S, device=p.device)
    call("egnn_kabsch", ptr(p), ptr(q), ptr(node_ptr), S, ptr(R), ptr(t), ptr(rmsd), stream_ptr(p.device))
    return R, t, rmsd


def kabsch_torch(P, Q):
    """evaluate_rmsd_for_pos_generate.py:11-51 for one pair of [n,3] point sets"""
    assert P.shape == Q.shape, "Matrix dimensions must match"
    R, t, rmsd = kabsch_batch([P], [Q])
    return R[0], t[0], rmsd[0]
