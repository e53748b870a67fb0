"""One process per GPU (torchrun): data-parallel training with ONE collective per step -- an all-reduce (sum) of the
flat fp32 gradient buffer over NCCL/NVLink -- and communication-free batch-sharded sampling.

The reference is single-process (main.py:125); nothing here replaces reference code, it shards its two loops:
  * training: whole graphs of the minibatch are split across ranks; each rank's loss is sum/num_graph over ITS graphs
    (parts/train_per_iretation.py:169), so averaging the summed gradients by the world size reproduces a single-GPU
    step over the concatenated batch, except that EGCL's global normaliser (EquivariantGraphNeuralNetwork.py:64) is
    evaluated per rank (use norm_scope='graph' for exact equivalence).
  * sampling: structures [r*S/W, (r+1)*S/W) go to rank r; no data-path collective; results are gathered at the end.
"""
from __future__ import annotations

import contextlib
import os
import sys

import torch
import torch.distributed as dist

from .optim import flat_grad_of, scatter_flat_grad


@contextlib.contextmanager
def _stdout_to_stderr():
    """NCCL prints its version banner on file descriptor 1 when the first communicator is created; programs whose
    stdout is a machine-readable result line (bench.py) create the communicator inside this redirect."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        yield
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)


def init_from_env(backend=None):
    """Initialise torch.distributed from torchrun's environment.  Returns (rank, world_size, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            with _stdout_to_stderr():
                dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                        device_id=torch.device("cuda", local))
                warm = torch.zeros(1, device=torch.device("cuda", local))
                dist.all_reduce(warm)  # creates the communicator now (banner, topology search) rather than in a timed step
                torch.cuda.synchronize()
        else:
            dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def shard_range(total, rank, world):
    """Contiguous, balanced [lo, hi) share of `total` independent units for `rank` (first `total % world` ranks get +1)."""
    base, rem = divmod(int(total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_flat(flat, world=None, group=None):
    """In-place sum over ranks of one flat tensor (the only collective of a training step)."""
    if dist.is_initialized() and (world or dist.get_world_size(group)) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def allreduce_gradients(module, group=None):
    """Sum the module's parameter gradients over ranks as ONE buffer.  Returns the reduced flat gradient; the caller
    divides by the world size (FlatAdam.step(grad_scale=1/world)) to average."""
    g, zero_copy = flat_grad_of(module)
    allreduce_flat(g, group=group)
    if not zero_copy:
        scatter_flat_grad(module, g)
    return g


class OverlappedGradSync:
    """The training step's only collective -- the all-reduce (sum) of the flat fp32 gradient -- split into a few buckets of
    whole layers and issued on a communication stream WHILE backward is still running.

    Backward walks the layers L-1 .. 0 and finishes layer l's slice of the flat gradient with its partial-row reduction;
    ``egnn_net_backward_events`` records a CUDA event there.  Bucket b (layers [lo_b, hi_b)) is all-reduced as soon as the
    event of its lowest layer has fired, so only the last bucket's all-reduce is exposed; ``finish()`` (called by
    FlatAdam.step, or by hand before any other optimizer) makes the compute stream wait for it.  Everything is stream
    ordered -- no host synchronisation -- and capturable in a CUDA graph (the side stream joins the capture through the
    events).  Needs ``zero_grad(set_to_none=True)`` every step: the per-parameter ``.grad`` tensors must be the views of the
    flat buffer autograd was handed, not accumulated copies."""

    def __init__(self, module, buckets=3, group=None):
        import ctypes as C
        self.module, self.group = module, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        L = module._net.L
        nb = max(1, min(int(buckets), L))
        # bucket b = layers [bounds[b], bounds[b+1]); backward reaches the HIGHEST bucket first
        self.bounds = [round(b * L / nb) for b in range(nb + 1)]
        self.pending = False
        self._C = C
        self._arr = (C.c_void_p * L)()
        if self.world > 1:
            dev = module.flat_parameters().device
            self.comm = torch.cuda.Stream(device=dev)
            self.events = []
            for b in range(nb):
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(dev))  # torch creates the cudaEvent_t lazily, at the first record
                self.events.append(ev)
                h = ev.cuda_event                          # raw cudaEvent_t (int or ctypes.c_void_p, by torch version)
                self._arr[self.bounds[b]] = int(getattr(h, "value", h) or 0)  # recorded after layer bounds[b], the bucket's last
                if not self._arr[self.bounds[b]]:
                    raise RuntimeError("torch.cuda.Event has no native handle after record()")
            self.done = torch.cuda.Event()
        module._grad_sync = self

    def active(self):
        return self.world > 1

    def event_array(self):
        return self._arr

    def launch(self, gflat):
        """Called by the autograd backward right after the backward kernels were enqueued."""
        L = self.module._net.L
        P = gflat.numel() // L
        gflat.record_stream(self.comm)
        with torch.cuda.stream(self.comm):
            for b in reversed(range(len(self.events))):
                self.comm.wait_event(self.events[b])
                dist.all_reduce(gflat[self.bounds[b] * P:self.bounds[b + 1] * P], op=dist.ReduceOp.SUM, group=self.group)
            self.done.record(self.comm)
        self.pending = True

    def finish(self):
        """Make the current stream wait for the reduced gradient (no-op when nothing is in flight)."""
        if self.pending:
            torch.cuda.current_stream(self.comm.device).wait_event(self.done)
            self.pending = False
            g, zero_copy = flat_grad_of(self.module)
            if not zero_copy:
                raise RuntimeError("OverlappedGradSync needs optimizer.zero_grad(set_to_none=True) before every backward: "
                                   "the parameter .grad tensors are not the views of the reduced flat buffer")


def broadcast_parameters(module, src=0, group=None):
    """Make every rank start from rank `src`'s weights (one broadcast of the flat parameter buffer)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(module.flat_parameters(), src=src, group=group)
        from .EquivariantGraphNeuralNetwork import bump_param_epoch
        bump_param_epoch()


def sample_sharded(egnn, dp, num_atoms, num_structures, seed=2024, gather=True, **kw):
    """Batch-sharded sampling: this rank draws its share with generator seed `seed + rank`; optionally all-gathers.

    Returns the dict of sampler.sample_structures for the local share, plus (when gather) 'pos_all' [S,n,3],
    'atoms_all' [S,n,A], 'finite_all' [S] on every rank."""
    from .sampler import sample_structures
    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    lo, hi = shard_range(num_structures, rank, world)
    dev = torch.device(kw.pop("device", "cuda"))
    gen = torch.Generator(device=dev)
    gen.manual_seed(seed + rank)
    noises = kw.pop("noises", None)
    if noises is not None:  # injected draws for the WHOLE job (parity runs): this rank takes the rows of its structures
        n = int(num_atoms)
        noises = {k: (v[:, lo * n:hi * n] if k.startswith("step") else v[lo * n:hi * n]) for k, v in noises.items()}
    out = sample_structures(egnn, dp, num_atoms, hi - lo, generator=gen, device=dev, noises=noises, **kw)
    out["range"] = (lo, hi)
    if gather and world > 1:
        sizes = [shard_range(num_structures, r, world) for r in range(world)]
        mx = max(b - a for a, b in sizes)

        def gather_cat(t):
            pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            pad[: t.shape[0]] = t
            bufs = [torch.empty_like(pad) for _ in range(world)]
            dist.all_gather(bufs, pad)
            return torch.cat([b[: hi_ - lo_] for b, (lo_, hi_) in zip(bufs, sizes)])

        out["pos_all"] = gather_cat(out["pos"].contiguous())
        out["atoms_all"] = gather_cat(out["atoms"].contiguous())
        out["finite_all"] = gather_cat(out["finite"].to(torch.int32)).bool()
    elif gather:
        out["pos_all"], out["atoms_all"], out["finite_all"] = out["pos"], out["atoms"], out["finite"]
    return out
