"""Drop-in for the reference's parts/train_per_iretation.py: ``EarlyStopping``, ``diffuse_as_batch``, ``train_epoch``,
``eval_epoch`` and ``generate`` with the reference's signatures and return values, running on libegnn_b200 kernels.

What changes against the reference (and why the results stay the same):
  * diffuse_as_batch (:36-92): the per-graph python loop with boolean masks becomes ONE kernel launch (one CTA per
    graph).  Diffusion times still come from python ``random.choice`` -- one draw per graph in graph order, the
    reference's RNG stream (:56) -- and the noise from torch's generator on the device (one [N,3] and one [N,A] draw
    instead of 2G small ones: same distribution, different stream).
  * train_epoch / eval_epoch (:99-262): noising, h assembly, COM projection, summed squared error and its gradient are
    fused kernels; the loss value is accumulated on the device and read once per epoch instead of once per step (:179).
  * generate (:264-444): the ``gen_num_per_spectrum`` samples of one test graph run as ONE batched trajectory with the
    normaliser of EGCL.message evaluated per structure (what the reference computes one structure at a time); failed
    (non-finite) structures are redrawn, 10 failures abort like :379-381.
"""
from __future__ import annotations

import random

import torch

from . import _lib
from .graph import check_pending, graph_index_for
from ._lib import call, ptr, stream_ptr
from .data import Data, DevicePrefetcher
from .diffusion_x_h import _node_ptr_of, remove_mean  # noqa: F401  (re-exported like the reference module)

I32 = torch.int32


class EarlyStopping:
    """parts/train_per_iretation.py:19-34."""

    def __init__(self, patience=0):
        self._step = 0
        self._loss = float("inf")
        self._patience = patience

    def validate(self, loss):
        if self._loss < loss:
            self._step += 1
            if self._step > self._patience:
                return True
        else:
            self._step = 0
            self._loss = loss
        return False


def _graph_boundaries(batch_data, graph_index=None):
    """(node_ptr int32 [G+1] on the device, G).  Uses a precomputed ``node_ptr`` when the batch carries one."""
    dev = batch_data.pos.device
    np_ = getattr(batch_data, "node_ptr", None)
    if np_ is not None:
        np_ = np_.to(device=dev, dtype=I32)
        return np_, np_.numel() - 1
    if graph_index is None:
        graph_index = getattr(batch_data, "batch", None)
    if graph_index is None:
        return torch.tensor([0, batch_data.pos.shape[0]], dtype=I32, device=dev), 1
    return _node_ptr_of(graph_index.to(dev))


def _diffuse(batch_data, node_ptr, G, dp, t=None, noise_pos=None, noise_h=None, cond=None):
    pos = batch_data.pos.float().contiguous()
    x0 = batch_data.x.float().contiguous()
    if not pos.is_cuda:
        raise _lib.EgnnError("diffuse_as_batch runs on CUDA tensors only: call batch.to('cuda') first")
    dev, N, A = pos.device, pos.shape[0], x0.shape[1]
    T = dp.num_diffusion_timestep
    if t is None:
        time_list = [i for i in range(1, T + 1)]
        t = [random.choice(time_list) for _ in range(G)]  # one draw per graph, in graph order (:56)
    if torch.is_tensor(t) and t.is_cuda:  # already on the device (a graphed step's static buffer): no host copy
        t_dev = t.to(dtype=I32).contiguous()
    else:
        t_dev = torch.as_tensor(t, dtype=torch.int64).to(device=dev, dtype=I32).contiguous()
    if noise_pos is None:
        noise_pos = torch.empty(N, 3, device=dev).normal_(mean=0, std=1)
    if noise_h is None:
        noise_h = torch.empty(N, A, device=dev).normal_(mean=0, std=1)
    noise_pos, noise_h = noise_pos.float().contiguous(), noise_h.float().contiguous()
    cdet = None if cond is None else cond.detach().float().contiguous()
    Cw = 0 if cdet is None else cdet.shape[1]
    alpha, sigma, _ = dp.device_tables(dev)
    z_x = torch.empty_like(pos)
    eps_x = torch.empty_like(pos)
    h_in = torch.empty(N, A + Cw + 1, device=dev)
    call("egnn_diffuse_batch", ptr(pos), ptr(x0), ptr(noise_pos), ptr(noise_h), ptr(cdet, allow_none=True), ptr(t_dev),
         ptr(alpha), ptr(sigma), T, ptr(node_ptr), G, A, Cw, ptr(z_x), ptr(eps_x), ptr(h_in), stream_ptr(dev))
    if cond is not None and cond.requires_grad:  # keep the graph to the conditioning network (spectrum compressor)
        h_in = torch.cat((h_in[:, :A], cond.float(), h_in[:, A + Cw:]), dim=1)
    return dict(pos=pos, x0=x0, z_x=z_x, eps_x=eps_x, eps_h=noise_h, h_in=h_in, t=t_dev, A=A)


def diffuse_as_batch(batch_data, graph_index, params, diffusion_process, t=None, noise_pos=None, noise_h=None):
    """parts/train_per_iretation.py:36-92 -- attaches pos_at_t, h_at_t, pos_at_0, h_at_0, y_for_noise_pos, y_for_noise_h,
    each_time_list to ``batch_data`` and returns it."""
    node_ptr, G = _graph_boundaries(batch_data, graph_index)
    d = _diffuse(batch_data, node_ptr, G, diffusion_process, t, noise_pos, noise_h)
    A = d["A"]
    counts = (node_ptr[1:] - node_ptr[:-1]).to(torch.int64)
    batch_data.pos_at_t = d["z_x"]
    batch_data.h_at_t = d["h_in"][:, :A].contiguous()
    batch_data.pos_at_0 = d["pos"]
    batch_data.h_at_0 = d["x0"]
    batch_data.y_for_noise_pos = d["eps_x"]
    batch_data.y_for_noise_h = d["eps_h"]
    batch_data.each_time_list = torch.repeat_interleave(d["t"].to(torch.int64), counts).view(-1, 1)
    return batch_data


class _EpsLoss(torch.autograd.Function):
    """loss = (||P(x' - z_x) - eps_x||^2 + ||h'[:, :A] - eps_h||^2) / G with its gradient from the same kernel."""

    @staticmethod
    def forward(ctx, x_out, h_out, z_x, eps_x, eps_h, node_ptr, G, A):
        x_out, h_out = x_out.contiguous(), h_out.contiguous()
        loss_g = torch.empty(G, device=x_out.device)
        gx = torch.empty_like(x_out)
        gh = torch.empty_like(h_out)
        call("egnn_loss_fwd_bwd", ptr(x_out), ptr(h_out), ptr(z_x), ptr(eps_x), ptr(eps_h), ptr(node_ptr), G,
             h_out.shape[1], A, ptr(loss_g), ptr(gx), ptr(gh), stream_ptr(x_out.device))
        ctx.save_for_backward(gx, gh)
        return loss_g.sum()

    @staticmethod
    def backward(ctx, g):
        gx, gh = ctx.saved_tensors
        return gx * g, gh * g, None, None, None, None, None, None


def _conditioning(batch_data, params, nn_dict):
    """[compressed spectrum | spectrum] [, exO] block of h (parts/train_per_iretation.py:136-144), or None."""
    parts = []
    if params.get("conditional", False):
        spec = batch_data.spectrum.to(batch_data.pos.device)
        if params.get("to_compress_spectrum", False):
            spec = nn_dict["spectrum_compressor"](spec)
        parts.append(spec)
    if params.get("give_exO", False):
        parts.append(batch_data.exO.to(batch_data.pos.device))
    if not parts:
        return None
    return torch.cat(parts, dim=1) if len(parts) > 1 else parts[0]


def batch_loss(egnn, batch_data, diffusion_process, params, t=None, noise_pos=None, noise_h=None, cond=None,
               graph_index=None, graph=None):
    """One minibatch of train_epoch up to ``loss = loss / num_graph`` (:130-169).  Returns (loss, num_graph, aux).
    ``graph``: a prebuilt graph.GraphIndex of batch_data.edge_index (graphed steps keep one and refresh it in place)."""
    node_ptr, G = _graph_boundaries(batch_data, graph_index)
    d = _diffuse(batch_data, node_ptr, G, diffusion_process, t, noise_pos, noise_h, cond)
    # edge-list validation is deferred to the next host sync (graph.check_pending): no pipeline stall per minibatch
    gi = graph if graph is not None else graph_index_for(batch_data.edge_index.to(d["z_x"].device), d["z_x"].shape[0],
                                                         node_ptr=node_ptr, validate="deferred")
    h, x = egnn(batch_data.edge_index, d["h_in"], d["z_x"], node_ptr=node_ptr,
                norm_scope=params.get("norm_scope", "call"), graph=gi)
    A = params.get("atom_type_size", d["A"])
    loss = _EpsLoss.apply(x, h, d["z_x"], d["eps_x"], d["eps_h"], node_ptr, G, A)
    return loss, G, dict(d, h_out=h, x_out=x, graph=gi)


def _run_epoch(nn_dict, loader, params, diffusion_process, optimizer, train):
    egnn = nn_dict["egnn"]
    egnn.train(train)
    egnn.to("cuda")
    if params.get("optimizer") == "RAdamScheduleFree":
        optimizer.train() if train else optimizer.eval()
    if params.get("to_compress_spectrum", False):
        nn_dict["spectrum_compressor"].train(train)
    epoch_loss = torch.zeros((), device="cuda")
    total_nodes = 0
    for data in DevicePrefetcher(loader, "cuda"):  # next minibatch's H2D copy overlaps this one's kernels
        total_nodes += data.num_nodes
        cond = _conditioning(data, params, nn_dict)
        if train:
            optimizer.zero_grad()
            loss, G, aux = batch_loss(egnn, data, diffusion_process, params, cond=cond)
            loss.backward()
            # the minibatch's edge-list validation has landed long before the host gets here (it was enqueued ahead of
            # the forward); reading it BEFORE the update keeps a malformed list from ever reaching optimizer.step()
            try:
                aux["graph"].check(wait=True)
            except _lib.UnsortedEdgeIndex:
                # any edge order is legal for the reference's scatter-add: sort by receiver (stable) and redo the batch
                data.edge_index = data.edge_index[:, torch.sort(data.edge_index[0], stable=True).indices].contiguous()
                optimizer.zero_grad()
                loss, G, aux = batch_loss(egnn, data, diffusion_process, params, cond=cond)
                loss.backward()
                aux["graph"].check(wait=True)
            optimizer.step()
        else:
            with torch.no_grad():
                loss, G, aux = batch_loss(egnn, data, diffusion_process, params, cond=cond)
                try:
                    aux["graph"].check(wait=True)
                except _lib.UnsortedEdgeIndex:
                    data.edge_index = data.edge_index[:, torch.sort(data.edge_index[0], stable=True).indices].contiguous()
                    loss, G, aux = batch_loss(egnn, data, diffusion_process, params, cond=cond)
        epoch_loss += loss.detach() * G  # (:178) summed squared error of the batch
    out = epoch_loss.item() / max(total_nodes, 1)  # per-node average (:181); the only host sync of the epoch
    check_pending()
    return out


def train_epoch(nn_dict, train_loader, params, diffusion_process, optimizer):
    """parts/train_per_iretation.py:99-183."""
    return _run_epoch(nn_dict, train_loader, params, diffusion_process, optimizer, True)


def eval_epoch(nn_dict, eval_loader, params, diffusion_process, optimizer):
    """parts/train_per_iretation.py:185-262."""
    return _run_epoch(nn_dict, eval_loader, params, diffusion_process, optimizer, False)


def generate(nn_dict, test_data, params, diffusion_process, gen_num_per_spectrum=5):
    """parts/train_per_iretation.py:264-444 -> (original_graph_list, generated_graph_list)."""
    from .sampler import sample_structures
    egnn = nn_dict["egnn"]
    egnn.eval()
    egnn.to("cuda")
    if params.get("to_compress_spectrum", False):
        nn_dict["spectrum_compressor"].eval().to("cuda")
    A = params["atom_type_size"]
    s = params.get("onehot_scaling_factor", 1.0)
    original_graph_list, generated_graph_list = [], []
    with torch.no_grad():
        for k in range(len(test_data)):
            data = test_data[k].to("cuda")
            n = data.x.shape[0]
            cond = _conditioning(data, params, nn_dict)
            need, nan_count = gen_num_per_spectrum, 0
            while need > 0:
                out = sample_structures(egnn, diffusion_process, n, need, atom_type_size=A, onehot_scale=s, cond=cond)
                for q in range(need):
                    if not bool(out["finite"][q]):
                        nan_count += 1
                        if nan_count == 10:  # (:379-381)
                            print("too much nan was generated")
                            raise SystemExit
                        continue
                    pos = out["pos"][q]
                    if torch.any(pos > 1000):  # (:434) unrealistic coordinates are dropped and redrawn
                        continue
                    graph = Data(x=out["atoms"][q], edge_index=out["edge_index_single"], pos=pos, h=out["h"][q])
                    if params.get("conditional", False):
                        graph.spectrum = data.spectrum
                    if params.get("give_exO", False):
                        graph.exO = data.exO
                    T = diffusion_process.num_diffusion_timestep
                    # the reference appends the SAME Data object every 100 steps and once more at t=0 (:336-337, :432)
                    generated_graph_list.append([graph] * (T // 100 + 1))
                    original_graph_list.append(data if params.get("conditional", False) else -1)
                    need -= 1
    return original_graph_list, generated_graph_list


class GraphedTrainStep:
    """One minibatch of train_epoch (:122-179) -- diffuse_as_batch, denoiser forward, loss, backward, gradient all-reduce,
    optimizer step -- captured ONCE in a CUDA graph and replayed (a new minibatch's edge-list preparation runs in front of
    the replay): one graph launch per step instead
    of ~190 kernel launches enqueued through Python (host enqueue was 1.7-1.9 ms against 4.2 ms of GPU time at cfg2).

    For minibatches of a FIXED shape (same node count, edge count and graph boundaries as the example batch): the content
    of the step's inputs changes, their shapes cannot.  ``step(batch)`` copies a (host or device) minibatch into the graph's
    static buffers -- staged on a copy stream, so the host-to-device copy of step i+1 overlaps the replay of step i when
    ``prefetch(batch)`` is called first -- draws the diffusion times with python ``random.choice`` like the reference (:56),
    and replays.  Noise is drawn inside the graph from torch's CUDA generator (graph-safe philox offsets).  The optimizer
    must be capturable (optim.FlatAdam(capturable=True)); its state and the weights are restored after the warm-up steps
    capture needs, so constructing this object does not train."""

    def __init__(self, egnn, diffusion_process, params, optimizer, example_batch, grad_scale=1.0, device="cuda", warmup=2,
                 static_noise=False):
        from .graph import GraphIndex
        self.egnn, self.dp, self.params, self.opt, self.grad_scale = egnn, diffusion_process, dict(params), optimizer, grad_scale
        dev = torch.device(device)
        self.dev = dev
        if not getattr(optimizer, "capturable", False):
            raise ValueError("GraphedTrainStep needs a capturable optimizer: optim.FlatAdam(..., capturable=True)")
        b = example_batch
        node_ptr, G = _graph_boundaries(b)
        self.G = G
        self.T = diffusion_process.num_diffusion_timestep
        # static inputs of the captured step, and one staging copy of each (filled on the copy stream)
        self.static = dict(pos=b.pos.to(dev).float().contiguous().clone(), x=b.x.to(dev).float().contiguous().clone(),
                           edge_index=b.edge_index.to(dev).contiguous().clone())
        self.node_ptr = node_ptr.to(dev).clone()
        self.stage = {k: torch.empty_like(v) for k, v in self.static.items()}
        self.t_dev = torch.ones(G, dtype=I32, device=dev)
        self.t_stage = torch.ones(G, dtype=I32, device=dev)
        self.t_pin = [torch.ones(G, dtype=I32).pin_memory() for _ in range(2)]
        self.t_pin_ev = [None, None]  # the H2D copy that last read each pinned buffer
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.staged_ev, self.consumed_ev, self.flip, self.staged_batch = None, None, 0, False
        self.batch = Data(x=self.static["x"], pos=self.static["pos"], edge_index=self.static["edge_index"])
        self.batch.node_ptr = self.node_ptr
        self.gi = GraphIndex(self.static["edge_index"], self.static["pos"].shape[0], node_ptr=self.node_ptr, validate=True)
        self.loss = torch.zeros((), device=dev)
        # static_noise: the step reads its noise from self.noise_pos / self.noise_h (the caller fills them before step():
        # injected draws for parity tests) instead of drawing it inside the graph
        N, A = self.static["pos"].shape[0], self.static["x"].shape[1]
        self.noise_pos = torch.zeros(N, 3, device=dev) if static_noise else None
        self.noise_h = torch.zeros(N, A, device=dev) if static_noise else None
        self.flags_pin = torch.zeros(2, dtype=I32).pin_memory()
        # warm-up (allocator, autograd, NCCL) on a side stream, then capture; weights / optimizer state are put back
        flat = egnn.flat_parameters()
        keep_w = flat.clone()
        keep_opt = None if optimizer.exp_avg is None else (optimizer.exp_avg.clone(), optimizer.exp_avg_sq.clone())
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()
        torch.cuda.synchronize(dev)
        with torch.no_grad():
            flat.copy_(keep_w)
        from .EquivariantGraphNeuralNetwork import bump_param_epoch
        bump_param_epoch()
        if keep_opt is None:
            optimizer.exp_avg.zero_()
            optimizer.exp_avg_sq.zero_()
        else:
            optimizer.exp_avg.copy_(keep_opt[0])
            optimizer.exp_avg_sq.copy_(keep_opt[1])
        optimizer.step_dev.fill_(optimizer.step_count)

    def _body(self):
        self.opt.zero_grad()
        loss, _, _ = batch_loss(self.egnn, self.batch, self.dp, self.params, t=self.t_dev, graph=self.gi,
                                noise_pos=self.noise_pos, noise_h=self.noise_h)
        loss.backward()
        self.opt.step(grad_scale=self.grad_scale)   # waits for the overlapped all-reduce when data-parallel
        self.loss.copy_(loss.detach())

    def prefetch(self, batch=None, t=None):
        """Stage the NEXT step's inputs on the copy stream (it overlaps the replay in flight): a minibatch of host or device
        tensors, or with batch=None only fresh diffusion times (the static buffers keep their minibatch)."""
        if t is None:
            time_list = [i for i in range(1, self.T + 1)]
            t = [random.choice(time_list) for _ in range(self.G)]  # one draw per graph, in graph order (:56)
        if batch is not None:
            for k, dst in self.stage.items():
                if tuple(getattr(batch, k).shape) != tuple(dst.shape):
                    raise _lib.EgnnError(f"GraphedTrainStep: minibatch field '{k}' has shape {tuple(getattr(batch, k).shape)}, "
                                         f"the captured step was built for {tuple(dst.shape)}")
        k = self.flip
        self.flip ^= 1
        if self.t_pin_ev[k] is not None:
            self.t_pin_ev[k].synchronize()  # two prefetches ago: long done unless the host runs that far ahead
        self.t_pin[k].copy_(torch.as_tensor(t, dtype=I32))
        with torch.cuda.stream(self.copy_stream):
            if self.consumed_ev is not None:
                self.copy_stream.wait_event(self.consumed_ev)  # the previous staging copy has been moved to the static buffers
            if batch is not None:
                for name, dst in self.stage.items():
                    dst.copy_(getattr(batch, name), non_blocking=True)
            self.t_stage.copy_(self.t_pin[k], non_blocking=True)
            self.t_pin_ev[k] = torch.cuda.Event()
            self.t_pin_ev[k].record(self.copy_stream)
            self.staged_ev = self.t_pin_ev[k]
            self.staged_batch = batch is not None

    def step(self, batch=None, t=None):
        """Train on `batch` (default: whatever prefetch() staged; with neither, the current static buffers again).
        Returns the loss as a device scalar that the next replay overwrites."""
        if batch is not None or self.staged_ev is None:
            self.prefetch(batch, t)               # nothing staged: at least new diffusion times, like every reference step
        cur = torch.cuda.current_stream(self.dev)
        if self.staged_ev is not None:
            cur.wait_event(self.staged_ev)
            if self.staged_batch:
                for k, dst in self.static.items():
                    dst.copy_(self.stage[k])      # device-to-device, microseconds
                # a new edge list: rebuild the index structures in place (6 small launches, ~0.15 ms of GPU time at cfg2;
                # kept out of the captured graph so that a step on unchanged inputs does not pay for it)
                self.gi.refresh(self.static["edge_index"])
            self.t_dev.copy_(self.t_stage)
            self.consumed_ev = torch.cuda.Event()
            self.consumed_ev.record(cur)
            self.staged_ev = None
        self.graph.replay()
        return self.loss

    def check(self):
        """Read the edge-list validation flags of the last replay (one small D2H copy + sync): raises like GraphIndex."""
        self.flags_pin.copy_(self.gi.flags, non_blocking=False)
        self.gi._raise_on(self.flags_pin.tolist())
