#!/usr/bin/env python
"""Benchmark of the quantization bottleneck hot path (BASELINE.json metric:
quantized tokens/sec, fwd+bwd).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step = one forward+backward of the drop-in ``Bottleneck`` (in_linear -> fused VQ ->
out_linear, plus the trainer's code-usage statistics) over one synthetic batch shaped like
``cfgs/larp_tokenizer.yaml``'s bottleneck block (C=768, D=24, K=8192, l2-normalised,
stochastic T=0.03, train mode), B clips x 1024 tokens per GPU.  Token-sharded across
GPUs (replicated codebook, one all-reduce of [codebook/linear grads | histograms]).

Prints ONE JSON line (rank 0).  ``--impl reference`` times the CPU oracle port (the
reference's own torch ops, all host threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

CFG = dict(C=768, D=24, K=8192, n=1024, T=0.03, beta=0.25, w_cb=1.0, loss_q_weight=0.1)
METRIC = "quantized_tokens_per_sec_fwd_bwd"
UNIT = "tokens/s"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=30)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=64, help="clips per GPU per step (x1024 tokens)")
    p.add_argument("--mode", default="stochastic", choices=["stochastic", "deterministic"])
    p.add_argument("--cpu-batch", type=int, default=4, help="clips in the bounded CPU sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--dim", type=int, default=CFG["D"], choices=[8, 16, 24, 32],
                   help="bottleneck_dim: 24 = cfgs/larp_tokenizer.yaml (the headline), 16 = larp_tokenizerf256t1024.yaml / "
                        "larp_tokenizer_large.yaml (BASELINE configs[2], [3])")
    args = p.parse_args()
    CFG["D"] = args.dim
    return args


def workload_name(args, batch):
    yaml = "cfgs/larp_tokenizer.yaml" if CFG["D"] == 24 else ("cfgs/larp_tokenizerf256t1024.yaml" if CFG["D"] == 16 else "custom")
    return (f"Bottleneck fwd+bwd, {yaml} block (C={CFG['C']}, D={CFG['D']}, K={CFG['K']}, "
            f"l2, {'stochastic T=0.03 train' if args.mode == 'stochastic' else 'deterministic argmin'}), "
            f"{batch}x{CFG['n']} tokens/GPU/step, fp32")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sustained=float(d.get(
            "bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback")


# ------------------------------------------------------------------------------------------
# clocks sampler (NVML), runs during the timed regions
# ------------------------------------------------------------------------------------------
class ClockSampler:
    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
    NOTE = {"sw_power_cap": 0x4, "hw_power_brake": 0x80}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._active = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            if self._active.is_set():
                try:
                    self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                    r = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                    for k, bit in {**self.BAD, **self.NOTE}.items():
                        if r & bit:
                            self.reasons.add(k)
                except Exception:
                    pass
            time.sleep(0.004)

    def start(self):
        self._active.set()

    def pause(self):
        self._active.clear()

    def result(self):
        self._stop.set()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------
# CPU reference leg (oracle port of the reference's torch ops; bench.py may execute oracle/
# only here)
# ------------------------------------------------------------------------------------------
def cpu_reference_step_fn(args, batch):
    from oracle import vq_oracle as O
    torch.manual_seed(1234)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    C, D, K, n = CFG["C"], CFG["D"], CFG["K"], CFG["n"]
    cfg = O.VQConfig(dim=D, codebook_size=K, commitment_loss_weight=CFG["beta"], codebook_loss_weight=CFG["w_cb"],
                     l2_normalized=True, stochastic=args.mode == "stochastic", stochastic_temperature=CFG["T"])
    lin_in = torch.nn.Linear(C, D)
    lin_out = torch.nn.Linear(D, C)
    W = torch.empty(K, D)
    torch.nn.init.kaiming_uniform_(W)
    params = [lin_in.weight, lin_in.bias, lin_out.weight, lin_out.bias, W.requires_grad_(True)]
    x = torch.randn(batch, n, C)
    G = torch.randn(batch, n, C)

    def step():
        for p in params:
            p.grad = None
        xr = x.clone().requires_grad_(True)
        out = O.bottleneck_forward(xr, params[0], params[1], params[2], params[3], params[4], cfg, training=True)
        O.code_usage_stats(out["bottleneck_rep"], K)
        torch.autograd.backward([out["output"], out["loss_q"]], [G, torch.tensor(CFG["loss_q_weight"])])
        return float(out["loss_q"])

    return step, batch * n, cores


def time_cpu(step, warmup, steps):
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    return (time.perf_counter() - t0) / steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    step, tokens, cores = cpu_reference_step_fn(args, args.cpu_batch)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    sec = time_cpu(step, warmup, steps)
    val = tokens / sec
    sample = f"{args.cpu_batch}x{CFG['n']} tokens per step, {steps} steps after {warmup} warm-up, all host threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(args, args.cpu_batch),
                                                       "l2_policy": "n/a (CPU)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist

    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import _lib, functional as VF, sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback; use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = V.load_library()
    C, D, K, n = CFG["C"], CFG["D"], CFG["K"], CFG["n"]
    B = args.batch
    N = B * n

    torch.manual_seed(1234)  # same weights on every rank (replicated codebook)
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "commitment_loss_weight": CFG["beta"], "codebook_loss_weight": CFG["w_cb"],
        "entropy_loss_weight": 0.0, "l2_normalized": True, "stochastic": args.mode == "stochastic",
        "stochastic_temperature": CFG["T"]}}}}
    m = V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(dev)
    m.train()
    torch.manual_seed(1234 + rank)  # per-rank data and per-rank Philox stream (reference: seed + rank)
    torch.cuda.manual_seed(1234 + rank)
    x = torch.randn(B, n, C, device=dev, requires_grad=True)
    G = torch.randn(B, n, C, device=dev)
    gq = torch.tensor(CFG["loss_q_weight"], device=dev)

    def step(xin=x, gin=G):
        for p in m.parameters():
            p.grad = None
        xin.grad = None
        out = m(xin)
        stats = m.regularizer.code_usage()
        torch.autograd.backward([out["output"], out["loss_q"]], [gin, gq])
        if world > 1:
            sharded.token_sharded_sync(m)
        return out, stats

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None

    # ---- device-resident throughput ("value") ------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    l0 = lib.b200vq_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler:
        sampler.start()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    if sampler:
        sampler.pause()
    launches = int(lib.b200vq_launch_count() - l0)
    ms = e0.elapsed_time(e1) / args.steps
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item())
    value = world * N / (ms_step * 1e-3)

    # ---- end to end through the public module API with HOST buffers ---------------------------
    xh = torch.randn(B, n, C).pin_memory()
    gh = torch.randn(B, n, C).pin_memory()
    # two device buffer sets: the copies of step i+1 run on a copy stream while step i computes (every
    # step's inputs still cross PCIe from pinned host memory inside the timed region)
    xd = [torch.empty(B, n, C, device=dev, requires_grad=True) for _ in range(2)]
    gd = [torch.empty(B, n, C, device=dev) for _ in range(2)]
    idx_h = torch.empty(B, n, dtype=torch.int64).pin_memory()
    sc_h = torch.empty(8, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    state = {"i": 0, "primed": False}

    def issue_copies(slot):
        with torch.cuda.stream(copy_stream), torch.no_grad():
            copy_stream.wait_event(consumed[slot])      # the step that last used this slot has finished
            xd[slot].copy_(xh, non_blocking=True)
            gd[slot].copy_(gh, non_blocking=True)
            ready[slot].record(copy_stream)

    def e2e_step():
        cur = state["i"] & 1
        if not state["primed"]:
            issue_copies(cur)
            state["primed"] = True
        issue_copies(cur ^ 1)                            # prefetch the next step's inputs
        torch.cuda.current_stream().wait_event(ready[cur])
        out, stats = step(xd[cur], gd[cur])
        consumed[cur].record()
        idx_h.copy_(out["bottleneck_rep"], non_blocking=True)
        sc_h[:3].copy_(torch.stack([out["loss_q"].detach(), out["loss_commit"].detach(), out["loss_codebook"].detach()]),
                       non_blocking=True)
        sc_h[3:7].copy_(stats, non_blocking=True)
        torch.cuda.synchronize()  # the caller reads the step's result on the host
        state["i"] += 1
        return float(sc_h[0])

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        e2e_step()
    barrier()
    if sampler:
        sampler.start()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    if sampler:
        sampler.pause()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = world * N / float(t.item())
    h2d = 2 * B * n * C * 4
    d2h = B * n * 8 + 7 * 4

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- scoring kernels, timed alone with CUDA events on their launch stream -------------------------
    pk = peaks()
    mode = {"stochastic": _lib.MODE_SAMPLE, "deterministic": _lib.MODE_ARGMIN_DIST}[args.mode]
    with torch.no_grad():
        z = m.project_in(x).contiguous()
        emb, e2, _ = VF.prepare_codebook(m.regularizer.embedding.weight, True)
        image = VF.prepare_codebook_tc(emb)
    keys = torch.zeros(N, dtype=torch.int64, device=dev)
    second = torch.zeros(2 * N + 4, dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def scores_exact():
        keys.zero_()
        _lib.check(lib.b200vq_scores_partial(z.data_ptr(), 0, D, N, D, emb.data_ptr(), e2.data_ptr(), K, 0, mode, 1,
                                             1.0 / CFG["T"], 1, 4, 0, keys.data_ptr(), st), "scores")

    def scores_tc():
        keys.zero_()      # the step does the same resets (b200vq_forward zeroes its workspace)
        second.zero_()
        _lib.check(lib.b200vq_scores_tc(z.data_ptr(), 0, D, N, D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(), K, 0,
                                        mode, 1, 1.0 / CFG["T"], 1, 4, 0, keys.data_ptr(), second.data_ptr(), None, st),
                   "scores_tc")

    def time_kernel(fn, reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if sampler:
            sampler.pause()
        return e0.elapsed_time(e1) / reps

    tc_ms = time_kernel(scores_tc, 20)       # includes the exact re-scoring launch in deterministic mode
    ex_ms = time_kernel(scores_exact, 5)
    flops = 2.0 * K * D * N                  # algorithmic: one fp32 dot product of length D per (token, code)
    kc = (3 * D + 15) // 16 * 16             # issued on the tensor cores: split-bf16, contraction length kc
    alg_bytes = N * (4 * D + 8) + K * kc * 2
    tc_tflops_issued = 2.0 * K * kc * N / (tc_ms * 1e-3) / 1e12
    fma_peak = 148 * 128 * 2 * 1.965e9 / 1e12
    traffic = None   # dram bytes per launch of the same kernel from the committed ncu --set full capture
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01c_traffic.json")))["tc_scores_kernel"]
        if N == 65536 and D == 24 and args.mode == "stochastic":
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except Exception:
        traffic = None
    roofline = {"kernel": "tc_scores_kernel (tcgen05 split-bf16 UMMA, TMEM accumulators, fused arg-max / Gumbel-max)",
                "bound": "tensor", "achieved": tc_tflops_issued, "peak": pk["bf16"], "unit": "TFLOP/s",
                "frac": tc_tflops_issued / pk["bf16"], "traffic": traffic, "algorithmic_bytes": alg_bytes,
                "peak_source": pk["source"],
                "kernel_ms": tc_ms, "share_of_step": tc_ms / ms_step,
                "algorithmic_tflops_fp32_equiv": flops / (tc_ms * 1e-3) / 1e12,
                "hbm": {"achieved_gbs": alg_bytes / (tc_ms * 1e-3) / 1e9, "peak_gbs": pk["hbm"],
                        "frac": alg_bytes / (tc_ms * 1e-3) / 1e9 / pk["hbm"]},
                "note": "issued flops = 2*K*kc*N with kc = 3*D padded to 16 (split bf16 emulating fp32)",
                "exact_fp32_kernel": {"kernel_ms": ex_ms, "achieved_tflops": flops / (ex_ms * 1e-3) / 1e12,
                                      "peak_tflops_nominal": fma_peak, "frac": flops / (ex_ms * 1e-3) / 1e12 / fma_peak}}

    # ---- CPU baseline (oracle port, bounded sample) ------------------------------------------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cstep, ctok, cores = cpu_reference_step_fn(args, args.cpu_batch)
        csec = time_cpu(cstep, 1, 3)
        cpu = {"value": ctok / csec, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{args.cpu_batch}x{n} tokens per step, 3 steps after 1 warm-up, same config"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(args, B), "tokens_per_gpu_per_step": N, "parallelism": f"token-sharded x{world}",
                   "l2_policy": f"inputs larger than L2 (x+grad = {2 * B * n * C * 4 / 2**20:.0f} MiB per step vs 126 MB L2)"},
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps},
        "gpu_launches": launches,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "clocks": sampler.result() if sampler else None,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    # stdout carries exactly ONE JSON line: anything libraries print there (e.g. NCCL's version banner)
    # is sent to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
