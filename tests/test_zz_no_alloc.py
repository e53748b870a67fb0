"""egnn_graph_prepare_noalloc -- the allocation-free variant of egnn_graph_prepare (scratch in a static __device__ array
inside the library; no counterpart in the reference).

Called through the C ABI (ctypes) and compared with the existing entry point on the same edge lists: integer work, so the
bar is BIT-EXACT on every output (recv, send, row_ptr, col_ptr, col_edge, flags).  Two calls in a row on different graphs
-- the second one smaller, so the shared array still holds the first call's leftovers -- plus a malformed list (flags) and
the size limit.  Collected last on purpose; passes alone and after the other tests.
"""
import ctypes as C
import re

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
I32 = torch.int32


def _max_ints():
    import os
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "egnn_b200.h")).read()
    m = re.search(r"#define\s+EGNN_GRAPH_PREPARE_NOALLOC_MAX_INTS\s+\(1\s*<<\s*(\d+)\)", hdr)
    return 1 << int(m.group(1))


def _outputs(N, E):
    return dict(recv=torch.full((max(E, 1),), -7, dtype=I32, device=DEV), send=torch.full((max(E, 1),), -7, dtype=I32, device=DEV),
                row_ptr=torch.full((N + 1,), -7, dtype=I32, device=DEV), col_ptr=torch.full((N + 1,), -7, dtype=I32, device=DEV),
                col_edge=torch.full((max(E, 1),), -7, dtype=I32, device=DEV), flags=torch.full((2,), -7, dtype=I32, device=DEV))


def _prepare(lib, ei, N, noalloc):
    from diffusion_model_b200._lib import ptr, stream_ptr
    E = int(ei.shape[1])
    o = _outputs(N, E)
    st = stream_ptr(ei.device)
    if noalloc:
        rc = lib.egnn_graph_prepare_noalloc(ptr(ei), N, E, ptr(o["recv"]), ptr(o["send"]), ptr(o["row_ptr"]), ptr(o["col_ptr"]),
                                            ptr(o["col_edge"]), ptr(o["flags"]), st)
    else:
        scratch = torch.empty(N + 1 + E, dtype=I32, device=DEV)
        rc = lib.egnn_graph_prepare(ptr(ei), N, E, ptr(o["recv"]), ptr(o["send"]), ptr(o["row_ptr"]), ptr(o["col_ptr"]),
                                    ptr(o["col_edge"]), ptr(scratch), ptr(o["flags"]), st)
    assert rc == 0, lib.egnn_b200_last_error().decode()
    torch.cuda.synchronize()
    return {k: v.cpu() for k, v in o.items()}


def _cutoff_graph(G, n, seed):
    """periodic cutoff graphs of the benchmarked kind (r_c = 5 A), receiver-sorted int64 [2,E] on the device"""
    import bench
    from diffusion_model_b200.graph import build_edges_cutoff
    pos, _, box = bench.synth_batch(G, n, seed)
    node_ptr = (torch.arange(G + 1) * n).to(I32).to(DEV)
    return build_edges_cutoff(pos.to(DEV), 5.0, node_ptr=node_ptr, box=box.to(DEV)), G * n


def test_graph_prepare_noalloc_matches_graph_prepare(lib):
    from diffusion_model_b200.graph import build_edges_fc
    big, Nb = _cutoff_graph(6, 192, 31)            # ~39k edges
    small = build_edges_fc([24, 17, 5], DEV)       # 866 edges: runs second, over the first call's leftovers
    Ns = 46
    assert Nb + 1 + big.shape[1] <= _max_ints()
    for ei, N in ((big, Nb), (small, Ns), (big, Nb)):   # calls in a row on ONE shared array
        want = _prepare(lib, ei, N, noalloc=False)
        got = _prepare(lib, ei, N, noalloc=True)
        for k in want:
            assert torch.equal(got[k], want[k]), k
        assert got["flags"].tolist() == [0, 0]
        # and the structures are right in their own terms: receiver CSR and sender CSR of the edge list
        recv, send = ei[0].cpu(), ei[1].cpu()
        assert torch.equal(got["recv"].long(), recv) and torch.equal(got["send"].long(), send)
        assert torch.equal(got["row_ptr"].long(), torch.searchsorted(recv, torch.arange(N + 1)))
        order = torch.sort(send, stable=True).indices    # edges by sender, ascending edge id inside a sender
        assert torch.equal(got["col_edge"].long(), order)
        assert torch.equal(got["col_ptr"].long(), torch.searchsorted(send[order], torch.arange(N + 1)))


def test_graph_prepare_noalloc_flags_and_empty(lib):
    # unsorted receivers and an out-of-range id raise the same flags as the existing entry point
    bad = torch.tensor([[3, 1, 2, 2], [0, 2, 1, 9]], dtype=torch.int64, device=DEV)
    want, got = _prepare(lib, bad, 5, noalloc=False), _prepare(lib, bad, 5, noalloc=True)
    assert got["flags"].tolist() == want["flags"].tolist() == [1, 1]
    empty = torch.zeros(2, 0, dtype=torch.int64, device=DEV)
    want, got = _prepare(lib, empty, 7, noalloc=False), _prepare(lib, empty, 7, noalloc=True)
    for k in ("row_ptr", "col_ptr", "flags"):
        assert torch.equal(got[k], want[k]), k


def test_graph_prepare_noalloc_rejects_sizes_beyond_the_static_scratch(lib):
    from diffusion_model_b200._lib import ptr, stream_ptr
    mx = _max_ints()
    o = _outputs(4, 4)
    ei = torch.zeros(2, 4, dtype=torch.int64, device=DEV)
    # the size check comes before any launch: the (too short) buffers are never touched
    rc = lib.egnn_graph_prepare_noalloc(ptr(ei), 4, C.c_int64(mx - 4), ptr(o["recv"]), ptr(o["send"]), ptr(o["row_ptr"]),
                                        ptr(o["col_ptr"]), ptr(o["col_edge"]), ptr(o["flags"]), stream_ptr(ei.device))
    assert rc == -1  # EGNN_ERR_ARG
    msg = lib.egnn_b200_last_error().decode()
    assert "exceeds the static scratch" in msg and str(mx) in msg, msg
    torch.cuda.synchronize()
    assert o["flags"].tolist() == [-7, -7]
