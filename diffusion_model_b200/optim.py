"""Fused Adam on the flat parameter / gradient buffers of a kernel-backed EGNN (one launch per step instead of one per
tensor).  Arithmetic of torch.optim.Adam, which the reference selects with optimizer: 'Adam' (parts/def_for_main.py:133-134).
Any torch optimizer built from ``egnn.parameters()`` keeps working too; this one is the fast path."""
from __future__ import annotations

import torch

from ._lib import call, ptr, stream_ptr


def bump_param_epoch():
    from .EquivariantGraphNeuralNetwork import bump_param_epoch as _b
    _b()


def flat_grad_of(module):
    """The gradient of every parameter of `module` as ONE contiguous fp32 tensor in flat-parameter order.

    The backward kernel writes all parameter gradients into one buffer and hands autograd views of it; when autograd
    kept those views as ``p.grad`` the buffer itself is returned (zero copy), otherwise the grads are concatenated."""
    fp = module._flat
    flat = fp.ensure()
    g = module._last_grad_flat
    if g is not None and g.numel() == flat.numel():
        off, ok = 0, True
        for p in fp.params:
            if p.grad is None or p.grad.data_ptr() != g.data_ptr() + 4 * off:
                ok = False
                break
            off += p.numel()
        if ok:
            return g, True
    parts = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in fp.params]
    return torch.cat(parts), False


def scatter_flat_grad(module, g):
    """Write a flat gradient back into the per-parameter ``.grad`` tensors (used after an all-reduce of a copy)."""
    off = 0
    for p in module._flat.params:
        n = p.numel()
        if p.grad is None:
            p.grad = g[off:off + n].view(p.shape)
        else:
            p.grad.copy_(g[off:off + n].view(p.shape))
        off += n


class FlatAdam:
    def __init__(self, module, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, capturable=False):
        """capturable=True keeps the step counter in device memory (egnn_adam_step_dev), so that ``step()`` can be captured
        in a CUDA graph and replayed with the right bias corrections (same meaning as torch.optim.Adam(capturable=True))."""
        self.module, self.lr, self.betas, self.eps, self.weight_decay = module, lr, betas, eps, weight_decay
        self.step_count = 0
        self.exp_avg = self.exp_avg_sq = None
        self.capturable, self.step_dev = capturable, None

    def zero_grad(self, set_to_none=True):
        for p in self.module._flat.params:
            p.grad = None
        self.module._last_grad_flat = None

    def train(self):
        pass

    def eval(self):
        pass

    def step(self, grad_scale=1.0):
        flat = self.module.flat_parameters()
        g, _ = flat_grad_of(self.module)
        if self.exp_avg is None or self.exp_avg.numel() != flat.numel() or self.exp_avg.device != flat.device:
            self.exp_avg = torch.zeros_like(flat)
            self.exp_avg_sq = torch.zeros_like(flat)
        sync = getattr(self.module, "_grad_sync", None)
        if sync is not None:
            sync.finish()  # the update waits for the (overlapped) all-reduce of the last bucket
        if self.capturable:
            if self.step_dev is None or self.step_dev.device != flat.device:
                self.step_dev = torch.full((1,), self.step_count, dtype=torch.int32, device=flat.device)
            call("egnn_counter_add", ptr(self.step_dev), 1, stream_ptr(flat.device))
            call("egnn_adam_step_dev", ptr(flat), ptr(g), ptr(self.exp_avg), ptr(self.exp_avg_sq), flat.numel(),
                 float(self.lr), float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
                 ptr(self.step_dev), float(grad_scale), stream_ptr(flat.device))
        else:
            self.step_count += 1
            call("egnn_adam_step", ptr(flat), ptr(g), ptr(self.exp_avg), ptr(self.exp_avg_sq), flat.numel(), float(self.lr),
                 float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay), self.step_count,
                 float(grad_scale), stream_ptr(flat.device))
        bump_param_epoch()  # parameters changed behind autograd's back: every packed weight image is stale

    def state_dict(self):
        if self.capturable and self.step_dev is not None:
            self.step_count = int(self.step_dev.item())
        return dict(step=self.step_count, exp_avg=self.exp_avg, exp_avg_sq=self.exp_avg_sq, lr=self.lr, betas=self.betas,
                    eps=self.eps, weight_decay=self.weight_decay)

    def load_state_dict(self, sd):
        self.step_count, self.exp_avg, self.exp_avg_sq = sd["step"], sd["exp_avg"], sd["exp_avg_sq"]
        if self.step_dev is not None:
            self.step_dev.fill_(self.step_count)
        self.lr, self.betas, self.eps, self.weight_decay = sd["lr"], sd["betas"], sd["eps"], sd["weight_decay"]


class RAdamScheduleFree:
    """Drop-in for ``schedulefree.RAdamScheduleFree`` (parts/def_for_main.py:137-138, ``optimizer: RAdamScheduleFree`` in
    parameters.yaml): same constructor arguments, ``train()`` / ``eval()`` / ``step()`` / ``zero_grad()`` and
    ``param_groups[0]`` bookkeeping keys (k, train_mode, weight_sum, lr_max, scheduled_lr), with the update fused into ONE
    kernel per contiguous parameter range (the whole EGNN is one range: its parameters are views of one flat buffer).

    The package is absent from the reference checkout and from this image, so the arithmetic is restated from the
    published algorithm (facebookresearch/schedule_free, radam_schedulefree.py) -- "parity unpinned" for this optimizer;
    tests compare against the CPU restatement oracle/radam_sf_ref.py.  Parameters outside a kernel-backed module (the
    spectrum compressor) are handled by the same kernel, one launch per contiguous run."""

    def __init__(self, params, lr=0.0025, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, r=0.0, weight_lr_power=2.0,
                 foreach=None, silent_sgd_phase=True):
        params = [p for p in params]
        if not params:
            raise ValueError("optimizer got an empty parameter list")
        self.param_groups = [dict(params=params, lr=lr, betas=betas, eps=eps, r=r, k=0, train_mode=False, weight_sum=0.0,
                                  lr_max=-1.0, scheduled_lr=0.0, weight_lr_power=weight_lr_power,
                                  weight_decay=weight_decay, silent_sgd_phase=silent_sgd_phase)]

    # -- contiguous runs ---------------------------------------------------------------------------------------
    @staticmethod
    def _runs(params, need_grad):
        """Maximal runs of parameters that are adjacent in memory (and, with need_grad, all have a gradient)."""
        runs, cur = [], []
        for p in params:
            if need_grad and p.grad is None:
                if cur:
                    runs.append(cur)
                cur = []
                continue
            ok = (cur and p.is_contiguous() and p.dtype == torch.float32 and p.device == cur[-1].device
                  and p.untyped_storage().data_ptr() == cur[-1].untyped_storage().data_ptr()
                  and p.data_ptr() == cur[-1].data_ptr() + 4 * cur[-1].numel())
            if ok:
                cur.append(p)
            else:
                if cur:
                    runs.append(cur)
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise ValueError("RAdamScheduleFree (libegnn_b200) needs contiguous fp32 CUDA parameters")
                cur = [p]
        if cur:
            runs.append(cur)
        return runs

    @staticmethod
    def _flat_view(tensors):
        t0 = tensors[0]
        n = sum(t.numel() for t in tensors)
        return t0.detach().as_strided((n,), (1,), t0.storage_offset())

    @staticmethod
    def _flat_grad(run):
        g0, off, ok = run[0].grad, 0, True
        for p in run:
            g = p.grad
            if (not g.is_contiguous() or g.dtype != torch.float32
                    or g.untyped_storage().data_ptr() != g0.untyped_storage().data_ptr()
                    or g.data_ptr() != g0.data_ptr() + 4 * off):
                ok = False
                break
            off += p.numel()
        if ok:
            return g0.as_strided((off,), (1,), g0.storage_offset())
        return torch.cat([p.grad.reshape(-1).float() for p in run])

    def _ensure_state(self, params):
        """One flat z / exp_avg_sq buffer per maximal memory run of ALL parameters (allocated at the first step)."""
        if getattr(self, "_layout", None) is not None:
            ok = all(p.data_ptr() == ptr_ for p, ptr_ in self._layout)
            if ok:
                return
            # parameter storage moved (module.to(), re-flattening): carry the state over by parameter
            old = {p: (self._z[p], self._v[p]) for p, _ in self._layout}
        else:
            old = {}
        self._layout = [(p, p.data_ptr()) for p in params]
        self._z, self._v, self._all_runs = {}, {}, []
        for run in self._runs(params, need_grad=False):
            n = sum(p.numel() for p in run)
            flat = self._flat_view(run)
            z, v = flat.clone(), torch.zeros_like(flat)
            off = 0
            for p in run:
                m = p.numel()
                if p in old:
                    z[off:off + m].copy_(old[p][0].reshape(-1))
                    v[off:off + m].copy_(old[p][1].reshape(-1))
                self._z[p], self._v[p] = z[off:off + m], v[off:off + m]
                off += m
            self._all_runs.append((run, z, v))

    # -- mode switches (y <-> x) ---------------------------------------------------------------------------------
    def _switch(self, weight):
        from .EquivariantGraphNeuralNetwork import bump_param_epoch
        if getattr(self, "_layout", None) is None:
            return  # no step yet: x == y == z
        self._ensure_state(self.param_groups[0]["params"])
        for run, z, _ in self._all_runs:
            flat = self._flat_view(run)
            call("egnn_lerp", ptr(flat), ptr(z), flat.numel(), float(weight), stream_ptr(flat.device))
        bump_param_epoch()

    def eval(self):
        g = self.param_groups[0]
        if g["train_mode"]:
            self._switch(1 - 1 / g["betas"][0])  # p <- x
            g["train_mode"] = False

    def train(self):
        g = self.param_groups[0]
        if not g["train_mode"]:
            self._switch(1 - g["betas"][0])  # p <- y
            g["train_mode"] = True

    def zero_grad(self, set_to_none=True):
        for p in self.param_groups[0]["params"]:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    def step(self, closure=None, grad_scale=1.0):
        from .EquivariantGraphNeuralNetwork import bump_param_epoch
        g = self.param_groups[0]
        if not g["train_mode"]:
            raise Exception("Optimizer was not in train mode when step is called. Please insert .train() and .eval() calls "
                            "on the optimizer. See documentation for details.")
        loss = closure() if closure is not None else None
        beta1, beta2 = g["betas"]
        step = g["k"] + 1
        beta2_t = beta2 ** step
        bias_correction2 = 1 - beta2_t
        rho_inf = 2 / (1 - beta2) - 1
        rho_t = rho_inf - 2 * step * beta2_t / bias_correction2
        rect = (((rho_t - 4) * (rho_t - 2) * rho_inf / ((rho_inf - 4) * (rho_inf - 2) * rho_t)) ** 0.5 if rho_t > 4.0
                else float(not g["silent_sgd_phase"]))
        lr = g["lr"] * rect
        g["scheduled_lr"] = lr
        lr_max = g["lr_max"] = max(lr, g["lr_max"])
        weight = (step ** g["r"]) * (lr_max ** g["weight_lr_power"])
        weight_sum = g["weight_sum"] = g["weight_sum"] + weight
        try:
            ckp1 = weight / weight_sum
        except ZeroDivisionError:
            ckp1 = 0
        adaptive_y_lr = lr * (beta1 * (1 - ckp1) - 1)
        params = g["params"]
        self._ensure_state(params)
        for run in self._runs(params, need_grad=True):
            y = self._flat_view(run)
            grad = self._flat_grad(run)
            z0, v0 = self._z[run[0]], self._v[run[0]]
            n = y.numel()
            z = z0.as_strided((n,), (1,), z0.storage_offset())  # runs with gradients are sub-ranges of the state runs
            v = v0.as_strided((n,), (1,), v0.storage_offset())
            call("egnn_radam_sf_step", ptr(y), ptr(grad), ptr(z), ptr(v), n, float(lr), float(beta2), float(g["eps"]),
                 float(g["weight_decay"]), float(ckp1), float(adaptive_y_lr), float(bias_correction2),
                 1 if rho_t > 4.0 else 0, float(grad_scale), stream_ptr(y.device))
        g["k"] = step
        bump_param_epoch()
        return loss

    def state_dict(self):
        g = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        params = self.param_groups[0]["params"]
        has = getattr(self, "_layout", None) is not None
        return dict(group=g, z=[self._z[p].clone() for p in params] if has else None,
                    exp_avg_sq=[self._v[p].clone() for p in params] if has else None)

    def load_state_dict(self, sd):
        self.param_groups[0].update(sd["group"])
        if sd["z"] is not None:
            params = self.param_groups[0]["params"]
            self._layout = None
            self._ensure_state(params)
            for p, z, v in zip(params, sd["z"], sd["exp_avg_sq"]):
                self._z[p].copy_(z.to(p.device).reshape(-1))
                self._v[p].copy_(v.to(p.device).reshape(-1))
