"""TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Minimal stand-in for ``torch_geometric.nn.MessagePassing`` so that the reference file
``EquivariantGraphNeuralNetwork.py`` (which does ``from torch_geometric.nn import MessagePassing``,
reference EquivariantGraphNeuralNetwork.py:3) can be imported UNMODIFIED in this container, where
torch_geometric is not installed and cannot be installed (no network; version unpinned by the
reference: no requirements file).

The stub restates only the arithmetic PyG contributes to the path:
  * ``propagate(edge_index=..., **kw)`` gathers every ``message()`` argument named ``*_i`` from
    ``edge_index[0]`` and every ``*_j`` from ``edge_index[1]`` when ``flow='target_to_source'``
    (swapped for ``'source_to_target'``), passes other kwargs through unchanged,
  * sum-aggregates the message onto ``edge_index[0]`` (resp. ``[1]``) with ``scatter_add_`` and
    ``dim_size = number of nodes``.
It is used by ``oracle/make_golden.py`` (run in the build container, where /root/reference exists)
to check the independent restatement in ``oracle/egnn_ref.py`` and to generate tests/golden/*.
"""
import inspect
import sys
import types

import torch


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr="sum", flow="source_to_target", **kwargs):
        super().__init__()
        if aggr not in ("sum", "add"):
            raise NotImplementedError("stub supports sum aggregation only")
        self.aggr = aggr
        self.flow = flow

    def propagate(self, edge_index, size=None, **kwargs):
        if self.flow == "target_to_source":
            idx_i, idx_j = edge_index[0], edge_index[1]
        else:
            idx_i, idx_j = edge_index[1], edge_index[0]
        num_nodes = None
        call = {}
        for name in inspect.signature(self.message).parameters:
            if name.endswith("_i") or name.endswith("_j"):
                src = kwargs[name[:-2]]
                num_nodes = src.size(0)
                call[name] = src.index_select(0, idx_i if name.endswith("_i") else idx_j)
            else:
                call[name] = kwargs[name]
        out = self.message(**call)
        agg = out.new_zeros((num_nodes,) + tuple(out.shape[1:]))
        index = idx_i.view(-1, *([1] * (out.dim() - 1))).expand_as(out)
        return agg.scatter_add_(0, index, out)


def install():
    """Register the stub as ``torch_geometric.nn.MessagePassing`` (idempotent)."""
    if "torch_geometric" in sys.modules and not getattr(sys.modules["torch_geometric"], "_b200_stub", False):
        return  # a real torch_geometric is present: use it
    tg = types.ModuleType("torch_geometric")
    tg._b200_stub = True
    tgnn = types.ModuleType("torch_geometric.nn")
    tgnn.MessagePassing = MessagePassing
    tg.nn = tgnn
    sys.modules["torch_geometric"] = tg
    sys.modules["torch_geometric.nn"] = tgnn
