"""PyG-free containers with the attribute surface the reference's loops use on ``torch_geometric.data.Data`` /
``Batch`` / ``DataLoader`` (parts/train_per_iretation.py:122-129, 288-316): ``x, pos, edge_index, batch, num_nodes,
spectrum, exO, id`` and ``.to(device)``.  Batching follows PyG's collate: node tensors concatenated, ``edge_index``
offset by the cumulative node count, ``batch`` = graph id per node.  Real PyG objects work with the loops as well.
"""
from __future__ import annotations

import random

import torch


class Data:
    def __init__(self, x=None, edge_index=None, pos=None, **kw):
        self.x, self.edge_index, self.pos = x, edge_index, pos
        for k, v in kw.items():
            setattr(self, k, v)

    @property
    def num_nodes(self):
        for t in (self.pos, self.x):
            if t is not None:
                return int(t.shape[0])
        return 0

    def keys(self):
        return [k for k, v in self.__dict__.items() if v is not None and not k.startswith("_")]

    def to(self, device):
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                setattr(self, k, v.to(device))
        return self

    def clone(self):
        return type(self)(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})

    def __repr__(self):
        parts = [f"{k}={list(v.shape) if torch.is_tensor(v) else v}" for k, v in self.__dict__.items() if v is not None]
        return f"{type(self).__name__}({', '.join(parts)})"


class Batch(Data):
    @classmethod
    def from_data_list(cls, data_list):
        out = cls()
        keys = [k for k in data_list[0].keys()]
        sizes = [d.num_nodes for d in data_list]
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + n)
        for k in keys:
            vals = [getattr(d, k) for d in data_list]
            if k == "edge_index":
                out.edge_index = torch.cat([v + o for v, o in zip(vals, offs)], dim=1)
            elif torch.is_tensor(vals[0]) and vals[0].dim() >= 1 and vals[0].shape[0] == sizes[0]:
                setattr(out, k, torch.cat(vals, dim=0))
            elif torch.is_tensor(vals[0]):
                setattr(out, k, torch.stack(vals))
            else:
                setattr(out, k, vals)
        dev = data_list[0].pos.device if data_list[0].pos is not None else data_list[0].x.device
        out.batch = torch.repeat_interleave(torch.arange(len(sizes), device=dev), torch.tensor(sizes, device=dev))
        out.node_ptr = torch.tensor(offs, dtype=torch.int32, device=dev)
        out.num_graphs = len(sizes)
        return out


class DataLoader:
    def __init__(self, dataset, batch_size=1, shuffle=False, drop_last=False):
        self.dataset, self.batch_size, self.shuffle, self.drop_last = list(dataset), batch_size, shuffle, drop_last

    def __len__(self):
        n = len(self.dataset)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        idx = list(range(len(self.dataset)))
        if self.shuffle:
            random.shuffle(idx)
        for s in range(0, len(idx), self.batch_size):
            chunk = idx[s:s + self.batch_size]
            if self.drop_last and len(chunk) < self.batch_size:
                return
            yield Batch.from_data_list([self.dataset[i] for i in chunk])


class DevicePrefetcher:
    """Iterates ``loader`` one batch ahead: batch i+1 is copied host->device on a side stream (from pinned memory when
    ``pin`` is set or the tensors already are pinned) while the caller's stream still works on batch i; the caller's
    stream waits on the copy's event before the batch is handed out, so no host synchronisation is involved.  The
    reference's loop does ``data.to('cuda')`` in line (parts/train_per_iretation.py:123); this is the same copy, moved
    off the critical path.  CUDA only."""

    def __init__(self, loader, device="cuda", pin=True):
        self.loader, self.device, self.pin = loader, torch.device(device), pin
        if self.device.type != "cuda":
            raise ValueError("DevicePrefetcher copies to a CUDA device")
        self._stream = None

    def __len__(self):
        return len(self.loader)

    def _upload(self, batch):
        if not isinstance(batch, Data):  # foreign containers (real PyG batches): plain in-line copy, no prefetch
            return batch.to(self.device), None
        if self._stream is None:
            self._stream = torch.cuda.Stream(self.device)
        out = batch.__class__.__new__(batch.__class__)
        out.__dict__.update(batch.__dict__)
        with torch.cuda.stream(self._stream):
            for k, v in batch.__dict__.items():
                if torch.is_tensor(v) and not v.is_cuda:
                    if self.pin and not v.is_pinned():
                        v = v.pin_memory()
                    out.__dict__[k] = v.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._stream)
        return out, ev

    def _hand_out(self, item):
        out, ev = item
        if ev is None:
            return out
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for v in out.__dict__.values():
            if torch.is_tensor(v) and v.is_cuda:
                v.record_stream(cur)  # allocated on the copy stream, consumed on the caller's
        return out

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._upload(next(it))
        except StopIteration:
            return
        for batch in it:
            cur, nxt = nxt, self._upload(batch)
            yield self._hand_out(cur)
        yield self._hand_out(nxt)
