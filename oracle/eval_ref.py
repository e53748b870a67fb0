"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's structure-evaluation functions used to compare sampled
structures ("Tolerances"): the local radial distribution function around atom 0 and the angle / bond
lengths of atom 0 with two neighbours.  Pinned against the unmodified reference functions by oracle/make_eval_golden.py
(tests/golden/eval_kat.npz).  Never imported by the product package.

Follows: evaluate_RDF.py:39-45 (length_from_exO), :48-60 (RDF), :62-63 (cos_similarity), :13-24 (Wasserstein);
CN2_evaluate.py:12-21 (calculate_angle_for_CN2, calculate_bond_length_for_CN2);
evaluate_Si-O-Si.py:28-33 (neighbour selection: atoms within a radius of atom 0).
"""
import numpy as np
import torch
from scipy.ndimage import gaussian_filter1d
from scipy.stats import wasserstein_distance


def length_from_exO(position):
    """distances of atoms 1.. from atom 0 (evaluate_RDF.py:39-45), fp32 torch.norm like the reference"""
    position = torch.as_tensor(position, dtype=torch.float32)
    return [torch.norm(position[i] - position[0]) for i in range(1, len(position))]


def rdf(position, sigma=5, R=5.0, dR=0.01, Normalize=False):
    """evaluate_RDF.py:48-60: histogram of the distances from atom 0 with STRICT bin edges r < d < r + dR over
    r = arange(dR, R + dR, dR), normalised by 4 pi rho r^2 dR with rho = n / (4/3 pi R^3), then gaussian_filter1d."""
    lengths = np.array([float(d) for d in length_from_exO(position)], dtype=np.float64)
    num_atom = len(position)
    ro = num_atom / (4 / 3 * np.pi * R ** 3)
    edges = np.arange(0 + dR, R + dR, dR)
    out = np.empty(len(edges))
    for k, r in enumerate(edges):
        out[k] = np.count_nonzero((r < lengths) & (lengths < r + dR)) / (4 * np.pi * ro * r ** 2 * dR)
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
