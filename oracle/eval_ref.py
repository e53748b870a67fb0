"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's structure-evaluation functions used to compare sampled
structures ("Tolerances"): the local radial distribution function around atom 0 and the angle / bond
lengths of atom 0 with two neighbours.  Pinned against the unmodified reference functions by oracle/make_eval_golden.py
(tests/golden/eval_kat.npz).  Never imported by the product package.

Follows: evaluate_RDF.py:39-45 (length_from_exO), :48-60 (RDF), :62-63 (cos_similarity), :13-24 (Wasserstein);
CN2_evaluate.py:12-21 (calculate_angle_for_CN2, calculate_bond_length_for_CN2);
evaluate_Si-O-Si.py:28-33 (neighbour selection: atoms within a radius of atom 0);
evaluate_rmsd_for_pos_generate.py:11-51 (kabsch_torch).
"""
import numpy as np
import torch
from scipy.ndimage import gaussian_filter1d
from scipy.stats import wasserstein_distance


def length_from_exO(position):
    """distances of atoms 1.. from atom 0 (evaluate_RDF.py:39-45), fp32 torch.norm like the reference"""
    position = torch.as_tensor(position, dtype=torch.float32)
    return [torch.norm(position[i] - position[0]) for i in range(1, len(position))]


def rdf_edges(R=5.0, dR=0.01):
    """bin edges of evaluate_RDF.py:53-56: r over np.arange(dR, R+dR, dR) (float64) and r + dR.  ``r < d < r + dR``
    there compares a python/numpy float with a 0-dim fp32 torch tensor, which torch evaluates in FP32 (the scalar is
    cast to the tensor's dtype): the effective edges are fp32(r) and fp32(r + dR).  Adjacent bins can therefore overlap
    or leave a gap of one ulp; a distance exactly equal to fp32(r) is in neither bin."""
    r = np.arange(0 + dR, R + dR, dR)
    return r, r.astype(np.float32), (r + dR).astype(np.float32)


def rdf_counts(position, R=5.0, dR=0.01):
    """integer histogram of evaluate_RDF.py:55 (the numerator), one count per bin"""
    lengths = np.array([float(d) for d in length_from_exO(position)], dtype=np.float32)
    _, lo, hi = rdf_edges(R, dR)
    return np.array([np.count_nonzero((lo[k] < lengths) & (lengths < hi[k])) for k in range(len(lo))], dtype=np.int64)


def rdf(position, sigma=5, R=5.0, dR=0.01, Normalize=False):
    """evaluate_RDF.py:48-60: histogram of the distances from atom 0 with STRICT bin edges r < d < r + dR over
    r = arange(dR, R + dR, dR), normalised by 4 pi rho r^2 dR with rho = n / (4/3 pi R^3), then gaussian_filter1d."""
    num_atom = len(position)
    ro = num_atom / (4 / 3 * np.pi * R ** 3)
    edges, _, _ = rdf_edges(R, dR)
    out = rdf_counts(position, R, dR) / (4 * np.pi * ro * edges ** 2 * dR)
    out = gaussian_filter1d(out, sigma)
    if Normalize:
        out = out / np.max(out)
    return out


def cos_similarity(a, b):
    return float(np.dot(a, b) / (np.linalg.norm(a) * np.linalg.norm(b)))


def wasserstein(rdf1, rdf2):
    return float(wasserstein_distance(rdf1, rdf2))


def angle_cn2(coords):
    """CN2_evaluate.py:12-16: angle (degrees) at atom 0 of a [3,3] tensor"""
    coords = torch.as_tensor(coords, dtype=torch.float32)
    v1, v2 = coords[1] - coords[0], coords[2] - coords[0]
    cos = torch.dot(v1, v2) / (torch.norm(v1) * torch.norm(v2))
    return float(np.degrees(torch.acos(cos).item()))


def bond_lengths_cn2(coords):
    """CN2_evaluate.py:18-21"""
    coords = torch.as_tensor(coords, dtype=torch.float32)
    return torch.norm(coords[1] - coords[0]).item(), torch.norm(coords[2] - coords[0]).item()


def neighbours_within(position, radius):
    """evaluate_Si-O-Si.py:28-33: indices i >= 1 with |pos_i - pos_0| < radius, in index order"""
    position = torch.as_tensor(position, dtype=torch.float32)
    return [i for i in range(1, position.shape[0]) if torch.norm(position[i] - position[0]) < radius]


def two_nearest(position):
    """the two atoms closest to atom 0 -- what the radius rule selects for a well-formed exO site; used for UNTRAINED
    samplers, whose structures have no 2-Angstrom shell, so that the same angle / bond functions still apply"""
    position = torch.as_tensor(position, dtype=torch.float32)
    d = torch.norm(position[1:] - position[0], dim=1)
    idx = torch.argsort(d)[:2] + 1
    return [0] + sorted(int(i) for i in idx)


def structure_set_stats(positions, R, dR, sigma=5):
    """per-set statistics of a list of [n,3] structures: mean RDF curve, all distances from atom 0, CN2 angles and
    mean bond lengths of atom 0 with its two nearest atoms"""
    curves, dists, angles, bonds = [], [], [], []
    for p in positions:
        p = torch.as_tensor(p, dtype=torch.float32)
        curves.append(rdf(p, sigma=sigma, R=R, dR=dR))
        dists.extend(float(d) for d in length_from_exO(p))
        tri = p[two_nearest(p)]
        angles.append(angle_cn2(tri))
        b1, b2 = bond_lengths_cn2(tri)
        bonds.append(0.5 * (b1 + b2))
    return dict(rdf=np.mean(curves, axis=0), dists=np.array(dists), angles=np.array(angles), bonds=np.array(bonds))


def kabsch(P, Q):
    """evaluate_rmsd_for_pos_generate.py:11-51 (kabsch_torch): rotation, translation and RMSD aligning P onto Q.  Kept
    quirk (:42-45): for a left-handed V U^T the reference negates the last COLUMN of Vt (the commented-out line negates
    the last row), i.e. R = diag(1,1,-1) V U^T -- an improper matrix times a mirror, not the optimal proper rotation."""
    P, Q = torch.as_tensor(P, dtype=torch.float32), torch.as_tensor(Q, dtype=torch.float32)
    cP, cQ = torch.mean(P, dim=0), torch.mean(Q, dim=0)
    t = cQ - cP
    p, q = P - cP, Q - cQ
    H = torch.matmul(p.transpose(0, 1), q)
    U, S, Vt = torch.linalg.svd(H)
    if torch.det(torch.matmul(Vt.transpose(0, 1), U.transpose(0, 1))) < 0.0:
        Vt = Vt.clone()
        Vt[:, -1] *= -1.0
    R = torch.matmul(Vt.transpose(0, 1), U.transpose(0, 1))
    rmsd = torch.sqrt(torch.sum(torch.square(torch.matmul(p, R.transpose(0, 1)) - q)) / P.shape[0])
    return R, t, rmsd


def cn2_site(position, radius=2.0):
    """evaluate_Si-O-Si.py:28-39 for one structure: neighbours of atom 0 within ``radius``; when there are exactly two,
    the angle (degrees) and the two bond lengths of CN2_evaluate.py:12-21, else None for those."""
    position = torch.as_tensor(position, dtype=torch.float32)
    nb = neighbours_within(position, radius)
    if len(nb) != 2:
        return nb, None, None
    tri = position[[0] + nb]
    return nb, angle_cn2(tri), bond_lengths_cn2(tri)
