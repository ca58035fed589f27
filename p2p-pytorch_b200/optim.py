"""FusedAdam: torch.optim.Adam as the reference configures it (train.py:241-242 -- lr 2e-4, betas (0.5, 0.999),
eps 1e-8, no weight decay, no amsgrad) as ONE kernel launch per network.

Parameters are re-homed into one flat fp32 buffer (each `p.data` becomes a view of it), gradients live in a
second flat buffer (each `p.grad` is a view, kept across steps: `zero_grad()` zeroes in place), and the
moments in two more.  The backward kernels accumulate straight into the gradient views, the data-parallel
all-reduce runs on slices of the flat gradient buffer, and `step()` is p2p_adam_step over the whole buffer:
28 bytes of HBM traffic per parameter.
"""
from __future__ import annotations

import torch

from . import _lib as L
from ._lib import call, ptr


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        params = [p for p in params]
        if any(isinstance(p, dict) for p in params):
            raise ValueError("FusedAdam takes a flat list of parameters (one group)")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self._ps = [p for p in self.param_groups[0]["params"] if p.requires_grad]
        if not self._ps:
            raise ValueError("no trainable parameters")
        dev = self._ps[0].device
        for p in self._ps:
            L.require_cuda(p, "parameter")
            if p.dtype != torch.float32 or p.device != dev:
                raise ValueError("FusedAdam needs fp32 parameters on one device")
        # every parameter starts on a 32-byte boundary of the flat buffer (vector loads in the kernels that read
        # weights in place, e.g. the spectral-norm matvec); the gaps stay zero in p, g, m and v forever
        self._offsets, n = [], 0
        for p in self._ps:
            self._offsets.append(n)
            n += (p.numel() + 7) // 8 * 8
        self.numel = n
        self.flat_p = torch.empty(n, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_p.zero_()
        with torch.no_grad():
            for p, off in zip(self._ps, self._offsets):
                k = p.numel()
                view = self.flat_p[off:off + k].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_g[off:off + k].view(p.shape)
        self._steps = 0
        self._step_t = None     # device-side step count (capturable mode)
        self._hyper_t = None    # device-side [lr, grad_scale] (capturable mode) and the host values it holds
        self._hyper_host = None
        self.grad_scale = 1.0   # set to 1/world_size by the data-parallel trainer (all-reduce is a sum)

    @property
    def steps(self) -> int:
        return int(self._step_t.item()) if self._step_t is not None else self._steps

    @steps.setter
    def steps(self, n: int):
        self._steps = int(n)
        if self._step_t is not None:
            self._step_t.fill_(int(n))

    def make_capturable(self):
        """Keep the step count in device memory (torch.optim.Adam's capturable=True): `step()` then makes no
        host-dependent launch and can be recorded into, and replayed from, a CUDA graph."""
        if self._step_t is None:
            self._step_t = torch.full((1,), self._steps, dtype=torch.int32, device=self.flat_p.device)
            self._hyper_t = torch.zeros(2, dtype=torch.float32, device=self.flat_p.device)
            self.sync_hyper()

    def sync_hyper(self):
        """Capturable mode: bring the device copy of (lr, grad_scale) up to date with `param_groups[0]['lr']` and
        `grad_scale` -- what torch.optim.lr_scheduler (networks.py:104-125, stepped at train.py:441-443) changes
        between iterations.  A tiny host-to-device copy, issued only when a value changed; the owner of a captured
        graph calls this before every replay (Pix2PixTrainer.step does)."""
        if self._hyper_t is None:
            return
        cur = (float(self.param_groups[0]["lr"]), float(self.grad_scale))
        if cur != self._hyper_host:
            self._hyper_t.copy_(torch.tensor(cur, dtype=torch.float32), non_blocking=False)
            self._hyper_host = cur

    def _rebind_grads(self):
        """Make sure every p.grad is still the flat view (a caller may have used set_to_none or assigned one)."""
        for p, off in zip(self._ps, self._offsets):
            view_ptr = self.flat_g.data_ptr() + 4 * off
            g = p.grad
            if g is not None and g.data_ptr() == view_ptr:
                continue
            view = self.flat_g[off:off + p.numel()].view(p.shape)
            if g is None:
                view.zero_()
            else:
                view.copy_(g)
            p.grad = view

    def zero_grad(self, set_to_none: bool = False):
        # gradients are persistent views of one buffer: zero it in place (one memset) and keep the views
        self._rebind_grads()
        self.flat_g.zero_()

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("closures are not supported")
        self._rebind_grads()
        g = self.param_groups[0]
        if self._step_t is not None:
            if not torch.cuda.is_current_stream_capturing():
                self.sync_hyper()          # (inside a capture the recorded kernel only keeps the POINTER)
            self._step_t.add_(1)
            call("p2p_adam_step_graph", ptr(self.flat_p), ptr(self.flat_g), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                 self.numel, float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), ptr(self._step_t),
                 ptr(self._hyper_t))
            return
        self._steps += 1
        call("p2p_adam_step", ptr(self.flat_p), ptr(self.flat_g), ptr(self.exp_avg), ptr(self.exp_avg_sq), self.numel,
             float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), self._steps,
             float(self.grad_scale))

    def state_dict(self):
        return {"steps": self.steps, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}]}

    def load_state_dict(self, sd):
        self.steps = int(sd["steps"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for k, v in sd["param_groups"][0].items():
            self.param_groups[0][k] = v
