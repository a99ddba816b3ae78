"""One CUDA graph per (shape, mode) for a whole `Bottleneck` training step (forward + backward).

The extension only ENQUEUES work on the current stream and never reads device data on the host (include/b200vq.h,
"Conventions"), so a step is capturable as is.  The one thing that had to move is the sampler's RNG state: the
reference draws from the torch CUDA generator (torch.multinomial, models/bottleneck.py:280) and so does the eager
path here (functional.philox_state), on the HOST -- a replayed graph would repeat the captured (seed, offset) forever.
`SimpleVectorQuantizer.enable_device_rng()` keeps (seed, offset) in a device tensor that the kernels read and advance,
which makes replays draw exactly what the eager path would draw from the same generator state.

Why bother: enqueueing one eager step costs the host ~0.4 ms of Python / autograd / ctypes (tools/host_probe.py)
against ~0.1 ms of GPU work at the reference's real per-GPU batch (8-16 clips); a replay costs the host one call.
"""
from __future__ import annotations

from typing import Optional

import torch


class GraphedBottleneckStep:
    """Captures `out = module(x); backward([out['output'], out['loss_q']], [grad_output, loss_q_grad])` on static
    buffers.  `replay(x, grad_output)` copies new inputs in (optional) and launches the graph; results are the
    static tensors `self.out` (the module's output dict), `self.stats` (code-usage statistics), `self.x.grad` and the
    parameters' `.grad` (overwritten, not accumulated, on every replay)."""

    def __init__(self, module, x: torch.Tensor, grad_output: torch.Tensor, loss_q_grad: float = 0.1, warmup: int = 3,
                 autocast_dtype: Optional[torch.dtype] = None, post_backward=None):
        if not x.is_cuda:
            raise RuntimeError("GraphedBottleneckStep needs CUDA tensors")
        self.module = module
        self.autocast_dtype = autocast_dtype   # bf16 / fp16: the step runs inside torch.autocast (grad_output in that dtype)
        # optional callable run right after backward INSIDE the captured step, e.g. the data-parallel gradient sync
        # (sharded.token_sharded_sync: NCCL collectives are capturable), so that a replay is the whole training step
        self.post_backward = post_backward
        self.reg = module.regularizer
        self.x = x.detach().clone().requires_grad_(True)
        self.g = grad_output.detach().clone()
        self.gq = torch.tensor(float(loss_q_grad), device=x.device)
        self.reg.enable_device_rng(x.device)
        cur = torch.cuda.current_stream(x.device)
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):          # warm-up off the capture: lazy initialisations, workspace growth
            for _ in range(max(1, warmup)):
                self._step()
        cur.wait_stream(side)
        torch.cuda.synchronize(x.device)
        for p in module.parameters():
            p.grad = None                      # backward then creates the .grad tensors inside the graph's pool
        self.x.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out, self.stats = self._step()
        self.reg.enable_device_rng(x.device)   # re-synchronise with the generator: the warm-up advanced the device state

    def _step(self):
        if self.autocast_dtype is not None:
            with torch.autocast("cuda", dtype=self.autocast_dtype):
                out = self.module(self.x)
        else:
            out = self.module(self.x)
        stats = self.reg.code_usage() if hasattr(self.reg, "code_usage") else None
        torch.autograd.backward([out["output"], out["loss_q"]], [self.g, self.gq])
        if self.post_backward is not None:
            self.post_backward()
        return out, stats

    def release(self):
        """Drop the captured graph (and its memory pool).  Call before torch.distributed.destroy_process_group() when
        `post_backward` captured a collective: a live graph that holds one keeps the communicator busy at teardown."""
        self.graph = None
        self.out = self.stats = None

    def sync_rng(self):
        """take (seed, offset) from the torch CUDA generator again (e.g. after torch.cuda.manual_seed)"""
        self.reg.enable_device_rng(self.x.device)

    def replay(self, x: Optional[torch.Tensor] = None, grad_output: Optional[torch.Tensor] = None):
        if x is not None:
            self.x.data.copy_(x, non_blocking=True)
        if grad_output is not None:
            self.g.copy_(grad_output, non_blocking=True)
        self.graph.replay()
        return self.out
