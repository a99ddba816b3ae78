#!/usr/bin/env python
"""Benchmark of the quantization bottleneck hot path (BASELINE.json metric:
quantized tokens/sec, fwd+bwd).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step = one forward+backward of the drop-in ``Bottleneck`` (in_linear -> fused VQ ->
out_linear, plus the trainer's code-usage statistics) over one synthetic batch shaped like
``cfgs/larp_tokenizer.yaml``'s bottleneck block (C=768, D=24, K=8192, l2-normalised,
stochastic T=0.03, train mode), B clips x 1024 tokens per GPU.  Token-sharded across
GPUs (replicated codebook, one all-reduce of [codebook/linear grads | histograms]).

Prints ONE JSON line (rank 0):
  value        device-resident tokens/s of the step replayed as ONE CUDA graph (GraphedBottleneckStep; the same
               kernels as the eager module call, whose rate is reported beside it as `eager`)
  e2e          the same step through the module API with pinned HOST buffers copied in / results copied out
  roofline     the dominant kernel (the tcgen05 stochastic scoring kernel) timed alone with CUDA events:
               frac = ALGORITHMIC flops 2*K*D*N / time / measured bf16 peak; `step_hbm` = the whole step's
               algorithmic bytes / step time / measured HBM peak
  small_batch  tokens/s at the reference's real per-GPU batches (8 and 16 clips), graph-replayed and eager
  cpu_baseline the reference bottleneck itself (baseline/_ref, kind "reference"; the oracle port when the staged
               reference files are absent) on the host cores over a bounded sample of the SAME workload
  tokenizer    BASELINE configs[1]: the unmodified LARPTokenizer encode -> quantize -> decode, fwd+bwd on this GPU,
               with the drop-in installed vs with the reference's own bottleneck (torch eager on the same GPU)
``--impl reference`` times the reference's CPU implementation on the same config (bounded sample per step).
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
CPU_CHUNK = 4          # clips per row chunk of the CPU arm (the N x K matrices of the reference: 2 x 128 MiB per chunk)
CPU_BUDGET_S = 150.0   # wall-clock target of a whole --impl reference run


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=30)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--batch", type=int, default=64, help="clips per GPU per step (x1024 tokens)")
    p.add_argument("--mode", default="stochastic", choices=["stochastic", "deterministic"])
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the small-batch, tokenizer and codebook-sharded legs")
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


def config_dict(args, world):
    """identical for both arms: the driver compares the two lines' configs"""
    B, n, C = args.batch, CFG["n"], CFG["C"]
    return {"workload": workload_name(args, B), "tokens_per_gpu_per_step": B * n, "parallelism": f"token-sharded x{world}",
            "l2_policy": f"inputs larger than L2 (x+grad = {2 * B * n * C * 4 / 2**20:.0f} MiB per step vs 126 MB L2)"}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sustained=float(d.get(
            "bf16_tflops_sustained", d["bf16_tflops"])), source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback")


def step_bytes_per_token():
    """algorithmic HBM bytes per token of the whole Bottleneck fwd+bwd (SURVEY.md 8d Boundary B):
    five passes over (N, C) fp32 activations (x, output; grad_output, x again, grad_x) + the (N, D) tensors"""
    C, D = CFG["C"], CFG["D"]
    fwd = 4 * C + 4 * C + 4 * D + (4 * D + 4 * D + 4 * D + 8 + 4)          # x, output, projected_z | z read, znorm, ste, idx, zlen
    bwd = 4 * C + 4 * C + 4 * C + (4 * D + 4 * D + 8 + 4 + 4 * D) + 3 * 4 * D  # grad_output, x, grad_x | quantizer bwd | ga, gz r/w
    return fwd + bwd


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
# CPU reference leg.  The reference module itself when its files are staged under baseline/_ref (or
# /root/reference in the build container), else the oracle port of its torch ops.  bench.py may
# execute oracle/ only here.
# ------------------------------------------------------------------------------------------
def cpu_reference_step_fn(args):
    """-> (chunk_step(i) running fwd+bwd on row chunk i of the 64 x 1024-token batch, tokens per chunk,
    chunks per full step, cores, kind)"""
    from oracle import vq_oracle as O
    torch.manual_seed(1234)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    C, D, K, n = CFG["C"], CFG["D"], CFG["K"], CFG["n"]
    chunks = max(1, args.batch // CPU_CHUNK)
    kind, ref = "port", None
    try:
        from oracle import ref_loader
        bott, registry = ref_loader.load_reference()
        spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
            "codebook_size": K, "commitment_loss_weight": CFG["beta"], "codebook_loss_weight": CFG["w_cb"],
            "entropy_loss_weight": 0.0, "entropy_loss_temperature": 0.01, "l2_normalized": True,
            "stochastic": args.mode == "stochastic", "stochastic_temperature": CFG["T"]}}}}
        ref = registry.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C})
        ref.train()
        kind = "reference"
    except Exception:
        ref = None
    x = torch.randn(CPU_CHUNK, n, C)
    G = torch.randn(CPU_CHUNK, n, C)
    gq = torch.tensor(CFG["loss_q_weight"])
    if ref is not None:
        params = list(ref.parameters())

        def chunk_step(i):
            if i == 0:
                for p in params:
                    p.grad = None
            xr = x.clone().requires_grad_(True)
            out = ref(xr)                                   # models/bottleneck.py:170-188, unmodified
            O.code_usage_stats(out["bottleneck_rep"], K)    # the trainer's statistics (larp_tokenizer_trainer.py:335-352)
            torch.autograd.backward([out["output"], out["loss_q"]], [G, gq])
            return float(out["loss_q"].detach())
    else:
        cfg = O.VQConfig(dim=D, codebook_size=K, commitment_loss_weight=CFG["beta"], codebook_loss_weight=CFG["w_cb"],
                         l2_normalized=True, stochastic=args.mode == "stochastic", stochastic_temperature=CFG["T"])
        lin_in = torch.nn.Linear(C, D)
        lin_out = torch.nn.Linear(D, C)
        W = torch.empty(K, D)
        torch.nn.init.kaiming_uniform_(W)
        params = [lin_in.weight, lin_in.bias, lin_out.weight, lin_out.bias, W.requires_grad_(True)]

        def chunk_step(i):
            if i == 0:
                for p in params:
                    p.grad = None
            xr = x.clone().requires_grad_(True)
            out = O.bottleneck_forward(xr, params[0], params[1], params[2], params[3], params[4], cfg, training=True)
            O.code_usage_stats(out["bottleneck_rep"], K)
            torch.autograd.backward([out["output"], out["loss_q"]], [G, gq])
            return float(out["loss_q"].detach())

    return chunk_step, CPU_CHUNK * n, chunks, cores, kind


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    chunk_step, ctok, chunks, cores, kind = cpu_reference_step_fn(args)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    chunk_step(0)                                   # untimed: first-touch allocations
    t0 = time.perf_counter()
    chunk_step(0)
    t_chunk = time.perf_counter() - t0
    # a step = the first `per_step` row chunks of the 64 x 1024-token batch, sized so that the run ends in ~CPU_BUDGET_S
    per_step = int(max(1, min(chunks, CPU_BUDGET_S / max(1e-6, (steps + warmup) * t_chunk))))

    def step():
        for i in range(per_step):
            chunk_step(i)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    sec = (time.perf_counter() - t0) / steps
    val = per_step * ctok / sec
    what = "reference module (models/bottleneck.py, unmodified)" if kind == "reference" else "oracle port of the reference's torch ops"
    sample = (f"{what}: {per_step} of {chunks} row chunks of {CPU_CHUNK}x{CFG['n']} tokens per step "
              f"({per_step * ctok} tokens), {steps} steps after {warmup} warm-up, {cores} host threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": config_dict(args, max(1, args.gpus)),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------
def make_module(V, args, dev, n=None):
    C, D, K = CFG["C"], CFG["D"], CFG["K"]
    n = n or CFG["n"]
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "commitment_loss_weight": CFG["beta"], "codebook_loss_weight": CFG["w_cb"],
        "entropy_loss_weight": 0.0, "l2_normalized": True, "stochastic": args.mode == "stochastic",
        "stochastic_temperature": CFG["T"]}}}}
    m = V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(dev)
    m.train()
    return m


def time_events(fn, reps, sampler=None, sync=None):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    (sync or torch.cuda.synchronize)()
    if sampler:
        sampler.start()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    (sync or torch.cuda.synchronize)()
    if sampler:
        sampler.pause()
    return e0.elapsed_time(e1) / reps


def small_batch_leg(V, args, dev, sampler):
    """tokens/s at the reference's real per-GPU batch sizes (scripts/train_larp_tokenizer_reproduce.sh: -b 8 / 16),
    where one eager step costs the host more than the GPU: graph replay vs eager module call"""
    out = {}
    for B in (8, 16):
        m = make_module(V, args, dev)
        x = torch.randn(B, CFG["n"], CFG["C"], device=dev)
        G = torch.randn(B, CFG["n"], CFG["C"], device=dev)
        gq = torch.tensor(CFG["loss_q_weight"], device=dev)
        xr = x.clone().requires_grad_(True)

        def eager():
            for p in m.parameters():
                p.grad = None
            xr.grad = None
            o = m(xr)
            m.regularizer.code_usage()
            torch.autograd.backward([o["output"], o["loss_q"]], [G, gq])

        for _ in range(5):
            eager()
        t_e = time_events(eager, 30, sampler)
        step = V.GraphedBottleneckStep(m, x, G, loss_q_grad=CFG["loss_q_weight"])
        for _ in range(5):
            step.replay()
        t_g = time_events(step.replay, 50, sampler)
        N = B * CFG["n"]
        out[f"batch{B}"] = {"tokens_per_step": N, "graph_ms": t_g, "graph_tokens_per_s": N / (t_g * 1e-3),
                            "eager_ms": t_e, "eager_tokens_per_s": N / (t_e * 1e-3)}
        m.regularizer.disable_device_rng()
    return out


def amp_leg(V, args, dev, sampler, ms_fp32):
    """BASELINE configs[3]: the same step under bf16 autocast (projected_z / output / grad_output bf16, everything else
    fp32) on the library's own kernels, graph-replayed"""
    B, n, C = args.batch, CFG["n"], CFG["C"]
    m = make_module(V, args, dev)
    x = torch.randn(B, n, C, device=dev)
    G = torch.randn(B, n, C, device=dev).to(torch.bfloat16)
    step = V.GraphedBottleneckStep(m, x, G, loss_q_grad=CFG["loss_q_weight"], autocast_dtype=torch.bfloat16)
    for _ in range(5):
        step.replay()
    t = time_events(step.replay, 30, sampler)
    ok = step.out["output"].dtype is torch.bfloat16 and step.out["projected_z"].dtype is torch.bfloat16
    m.regularizer.disable_device_rng()
    N = B * n
    bpt = step_bytes_per_token() - 2 * 2 * C - 2 * CFG["D"] * 2     # output and grad_output halve, projected_z too
    return {"dtype": "bf16 activations (autocast), fp32 accumulate / quantizer / gradients", "ms_per_step": t,
            "tokens_per_s": N / (t * 1e-3), "vs_fp32": ms_fp32 / t, "activation_dtypes_ok": bool(ok),
            "step_hbm": {"bytes_per_token": bpt, "achieved_gbs": bpt * N / (t * 1e-3) / 1e9}}


def tokenizer_leg(V, dev):
    """BASELINE configs[1]: unmodified LARPTokenizer (cfgs/larp_tokenizer.yaml bottleneck block, 16 x 128 x 128 clips,
    B = 8) fwd+bwd on this GPU: drop-in installed vs the reference's own bottleneck (eager torch) in the same tokenizer"""
    try:
        from oracle import ref_tokenizer as RT
        if RT.reference_root() is None:
            return {"unavailable": "reference files not staged (python -m oracle.stage_ref in the build container)"}
        pkg = RT.load_tokenizer_package()
    except Exception as e:   # pragma: no cover
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}
    saved = dict(pkg.models)
    res = {}
    B = 8
    try:
        clip = torch.rand(B, 3, 16, 128, 128, device=dev)
        for name in ("reference_bottleneck", "b200_dropin"):
            if name == "b200_dropin":
                V.install(pkg)
            tok = RT.build_tokenizer(pkg).to(dev)
            tok.train()
            params = [p for p in tok.parameters() if p.requires_grad]

            def step():
                for p in params:
                    p.grad = None
                out = tok(clip)
                (out["pred_frames"].mean() + 0.1 * out["loss_q"]).backward()

            def bott_only():
                for p in tok.bottleneck.parameters():
                    p.grad = None
                xin = torch.randn(B, 1024, 768, device=dev, requires_grad=True)
                out = tok.bottleneck(xin)
                (out["output"].mean() + 0.1 * out["loss_q"]).backward()

            for _ in range(3):
                step()
            t = time_events(step, 5)
            for _ in range(3):
                bott_only()
            tb = time_events(bott_only, 10)
            res[name] = {"ms_per_step": t, "tokens_per_s": B * 1024 / (t * 1e-3), "bottleneck_only_ms": tb}
            del tok
        res["clips"] = B
        res["note"] = ("whole tokenizer step (6+6 timm-style blocks on torch SDPA/cuBLAS, stub deps); only the bottleneck differs "
                       "between the two lines")
    except Exception as e:   # pragma: no cover
        res["error"] = f"{type(e).__name__}: {e}"[:300]
    finally:
        pkg.models.clear()
        pkg.models.update(saved)
    return res


def codebook_sharded_leg(V, dev, world, rank):
    """BASELINE configs[4] at its top end: K = 262 144, D = 16, 65 536 tokens, codes sharded over the ranks (tokens
    replicated), fwd+bwd of the quantizer step; indices asserted equal to the unsharded kernel on this rank's GPU"""
    import torch.distributed as dist
    from video_tokenizer_b200 import _lib, functional as VF, sharded
    K, D, N = 262144, 16, 65536
    g = torch.Generator().manual_seed(1234)
    z = torch.randn(N // 1024, 1024, D, generator=g).to(dev).requires_grad_(True)
    G = torch.randn(N // 1024, 1024, D, generator=g).to(dev)
    gq = torch.tensor(0.1, device=dev)
    torch.manual_seed(0)
    W = torch.empty(K, D)
    torch.nn.init.kaiming_uniform_(W)
    kb, ke = sharded.shard_bounds(K, world, rank, align=256)
    Wl = W[kb:ke].to(dev).requires_grad_(True)
    Wfull = W.to(dev).requires_grad_(True)
    rec = {"K": K, "D": D, "tokens": N, "n_gpus": world}
    for mode_name, mode in (("deterministic", _lib.MODE_ARGMIN_DIST), ("stochastic", _lib.MODE_SAMPLE)):
        def sharded_step():
            z.grad = None
            Wl.grad = None
            zn, ste, idx, losses, hists = sharded.codebook_sharded_quantize(z, Wl, K, kb, mode=mode, l2norm=True,
                                                                           t_inv=1 / 0.03, seed=7, offset=16)
            torch.autograd.backward([ste, losses[0]], [G, gq])
            return idx

        def single_step():
            z.grad = None
            Wfull.grad = None
            zn, emb, ste, idx, losses, h, hf = VF.vq_quantize(z, Wfull, mode=mode, l2norm=True, t_inv=1 / 0.03, seed=7, offset=16)
            torch.autograd.backward([ste, losses[0]], [G, gq])
            return idx

        def barrier():
            dist.barrier()
            torch.cuda.synchronize()

        # the sharded step is the record; it is timed first so that a failure of the one-GPU comparison cannot lose it
        for _ in range(3):
            idx_s = sharded_step()
        t_s = torch.tensor([time_events(sharded_step, 10, sync=barrier)], device=dev)
        dist.all_reduce(t_s, op=dist.ReduceOp.MAX)
        entry = {"sharded_ms": float(t_s), "tokens_per_s": N / (float(t_s) * 1e-3)}

        # comparison: the whole codebook on this rank's GPU alone (every rank runs it; slowest rank counts)
        ok, t_one, equal, err = 1.0, 0.0, False, None
        for attempt in range(2):
            try:
                idx_1 = single_step()
                torch.cuda.synchronize()
                equal = bool(torch.equal(idx_s.reshape(-1), idx_1.reshape(-1)))
                single_step()
                t_one = time_events(single_step, 5)
                ok = 1.0
                break
            except Exception:   # pragma: no cover
                import traceback
                ok = 0.0
                err = (err or "") + f"[attempt {attempt}] " + traceback.format_exc()[-600:]
        flags = torch.tensor([ok, -t_one], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        if float(flags[0]) > 0.5:
            t_1 = -float(flags[1])
            entry.update({"one_gpu_ms": t_1, "speedup_over_one_gpu": t_1 / float(t_s), "indices_equal_to_unsharded": equal})
        if err:
            entry["one_gpu_error"] = err
        rec[mode_name] = entry
    return rec


def _trace(msg):
    if os.environ.get("B200VQ_BENCH_TRACE"):
        print(f"[bench rank {os.environ.get('RANK', '0')}] {msg}", file=sys.stderr, flush=True)


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
    warm = max(args.warmup, 3)

    torch.manual_seed(1234)  # same weights on every rank (replicated codebook)
    m = make_module(V, args, dev)
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

    def max_over_ranks(v):
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local_rank) if rank == 0 else None

    # ---- eager module call, device-resident inputs --------------------------------------------------
    for _ in range(warm):
        step()
    barrier()
    l0 = lib.b200vq_launch_count()
    ms_eager = max_over_ranks(time_events(step, args.steps, sampler, barrier))
    launches = int(lib.b200vq_launch_count() - l0)

    # ---- the same step as ONE CUDA graph per rank (+ the gradient / histogram all-reduce behind it) ----------
    graphed = None
    ms_step, value_kind = ms_eager, "eager module call"
    if args.mode == "stochastic" and os.environ.get("B200VQ_BENCH_NO_GRAPH") is None:
        try:
            sync_in_graph = False
            if world > 1 and os.environ.get("B200VQ_BENCH_SYNC_IN_GRAPH") is not None:   # opt-in
                # the whole data-parallel step in the graph: the all-reduce is captured too and starts behind the
                # backward's grads-ready event, beside the grad_x pass
                try:
                    sharded.enable_grad_overlap(dev)
                    graphed = V.GraphedBottleneckStep(m, x.detach(), G, loss_q_grad=CFG["loss_q_weight"], warmup=2,
                                                      post_backward=lambda: sharded.token_sharded_sync(m))
                    sync_in_graph = True
                except Exception as e:   # pragma: no cover
                    sharded.disable_grad_overlap(dev)
                    print(f"bench.py: rank {rank}: captured sync unavailable ({type(e).__name__}: {e}); syncing after the replay",
                          file=sys.stderr)
                    graphed = None
            if graphed is None:
                graphed = V.GraphedBottleneckStep(m, x.detach(), G, loss_q_grad=CFG["loss_q_weight"], warmup=2)

            _trace(f"graph captured (sync_in_graph={sync_in_graph})")

            def gstep():
                graphed.replay()
                if world > 1 and not sync_in_graph:
                    sharded.token_sharded_sync(m)

            for _ in range(warm):
                gstep()
            _trace("graph warm-up enqueued")
            barrier()
            _trace("graph warm-up done")
            ms_step = max_over_ranks(time_events(gstep, args.steps, sampler, barrier))
            _trace("graph timed")
            value_kind = "CUDA-graph replay of the module's fwd+bwd (GraphedBottleneckStep)"
            if world > 1:
                value_kind += (" with the gradient / histogram all-reduce captured in the graph behind the grads-ready event"
                               if sync_in_graph else " followed by the gradient / histogram all-reduce")
        except Exception as e:   # pragma: no cover -- keep the eager number rather than no number
            graphed = None
            value_kind = f"eager module call (graph capture failed: {type(e).__name__}: {e})"[:200]
        m.regularizer.disable_device_rng()
        if world > 1 and os.environ.get("B200VQ_BENCH_KEEP_OVERLAP") is None:
            sharded.disable_grad_overlap(dev)   # the eager legs below sync in stream order
    value = world * N / (ms_step * 1e-3)

    _trace("device-resident legs done")
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
        _trace(f"e2e step {state['i']}: inputs ready")
        out, stats = step(xd[cur], gd[cur])
        _trace(f"e2e step {state['i']}: step enqueued")
        consumed[cur].record()
        idx_h.copy_(out["bottleneck_rep"], non_blocking=True)
        sc_h[:3].copy_(torch.stack([out["loss_q"].detach(), out["loss_commit"].detach(), out["loss_codebook"].detach()]),
                       non_blocking=True)
        sc_h[3:7].copy_(stats, non_blocking=True)
        torch.cuda.synchronize()  # the caller reads the step's result on the host
        _trace(f"e2e step {state['i']}: synchronized")
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
    e2e_val = world * N / max_over_ranks((time.perf_counter() - t0) / e2e_steps)
    h2d = 2 * B * n * C * 4
    d2h = B * n * 8 + 7 * 4

    # ---- codebook-sharded record (BASELINE configs[4]) rides on the multi-GPU runs --------------------------
    cb_sharded = None
    if world > 1 and not args.no_extras:
        try:
            cb_sharded = codebook_sharded_leg(V, dev, world, rank)
        except Exception as e:   # pragma: no cover
            cb_sharded = {"error": f"{type(e).__name__}: {e}"[:300]}

    if rank != 0:
        if world > 1:
            graphed = None            # a graph that holds a captured collective goes before its communicator
            torch.cuda.synchronize()
            dist.barrier()
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
        keys.zero_()      # the step does the same reset
        if mode != _lib.MODE_SAMPLE:
            second.zero_()
        _lib.check(lib.b200vq_scores_tc(z.data_ptr(), 0, D, N, D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(), K, 0,
                                        mode, 1, 1.0 / CFG["T"], 1, 4, 0, keys.data_ptr(), second.data_ptr(), None, st),
                   "scores_tc")

    for _ in range(3):
        scores_tc()
    tc_ms = time_events(scores_tc, 20, sampler)       # includes the exact re-scoring launches in deterministic mode
    for _ in range(2):
        scores_exact()
    ex_ms = time_events(scores_exact, 5, sampler)
    flops = 2.0 * K * D * N                  # algorithmic: one fp32 dot product of length D per (token, code)
    if mode == _lib.MODE_SAMPLE:
        kc = (D + 15) // 16 * 16             # the sampler filters with ONE bf16 product (contraction length D padded to 16)
        kname = ("tc_sample_kernel (tcgen05 bf16 filter, TMEM accumulators, lazy Gumbel-max with exact fp32 expansion of the "
                 "candidates)")
    else:
        kc = (3 * D + 15) // 16 * 16         # deterministic modes: split-bf16 emulating fp32, contraction length 3*D
        kname = "tc_scores_kernel (tcgen05 split-bf16 UMMA, TMEM accumulators, fused arg-max) + exact resolve"
    alg_bytes = N * (4 * D + 8) + K * kc * 2
    # fp32 FMA peak of this box, measured (not the nominal 128 lanes x 2 x clock)
    probe_out = torch.zeros(1, device=dev)
    import ctypes
    fl = ctypes.c_double(0.0)

    def fma_probe():
        _lib.check(lib.b200vq_fma_probe(4096, probe_out.data_ptr(), ctypes.addressof(fl), st), "fma_probe")

    for _ in range(2):
        fma_probe()
    fma_ms = time_events(fma_probe, 10)
    fma_peak = fl.value / (fma_ms * 1e-3) / 1e12
    traffic, traffic_src = None, None
    try:   # dram bytes per launch of this kernel from the committed ncu --set full capture of THIS library version
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if (tj.get("library_version") == lib.b200vq_version() and N == tj.get("N") and D == tj.get("D")
                and args.mode == tj.get("mode")):
            traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
            traffic_src = tj.get("source")
    except Exception:
        traffic = None
    bpt = step_bytes_per_token()
    step_gbs = bpt * N / (ms_step * 1e-3) / 1e9
    roofline = {"kernel": kname, "bound": "tensor",
                "achieved": flops / (tc_ms * 1e-3) / 1e12, "peak": pk["bf16"], "unit": "TFLOP/s",
                "frac": flops / (tc_ms * 1e-3) / 1e12 / pk["bf16"],
                "note": "achieved = ALGORITHMIC flops 2*K*D*N per launch / CUDA-event time of the kernel alone",
                "tensor_pipe_issued_frac": 2.0 * K * kc * N / (tc_ms * 1e-3) / 1e12 / pk["bf16"],
                "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes": alg_bytes,
                "peak_source": pk["source"], "kernel_ms": tc_ms, "share_of_step": tc_ms / ms_step,
                "step_hbm": {"bytes_per_token": bpt, "achieved_gbs": step_gbs, "peak_gbs": pk["hbm"],
                             "frac": step_gbs / pk["hbm"]},
                "exact_fp32_kernel": {"kernel_ms": ex_ms, "achieved_tflops": flops / (ex_ms * 1e-3) / 1e12,
                                      "peak_tflops_measured": fma_peak, "frac": flops / (ex_ms * 1e-3) / 1e12 / fma_peak}}

    small = tok = amp = None
    if not args.no_extras and world == 1 and args.mode == "stochastic":
        try:
            amp = amp_leg(V, args, dev, sampler, ms_step)
        except Exception as e:   # pragma: no cover
            amp = {"error": f"{type(e).__name__}: {e}"[:300]}
    if not args.no_extras and world == 1:
        try:
            small = small_batch_leg(V, args, dev, sampler)
            for k in small:
                small[k]["vs_full_batch_rate"] = small[k]["graph_tokens_per_s"] / value
        except Exception as e:   # pragma: no cover
            small = {"error": f"{type(e).__name__}: {e}"[:300]}
        tok = tokenizer_leg(V, dev)

    # ---- CPU baseline: the reference module on the host cores, bounded sample of the same workload ---------------
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        chunk_step, ctok, chunks, cores, kind = cpu_reference_step_fn(args)
        chunk_step(0)
        t0 = time.perf_counter()
        done = 0
        while done < chunks and (done < 2 or time.perf_counter() - t0 < 12.0):
            chunk_step(done)
            done += 1
        csec = time.perf_counter() - t0
        what = "reference module (models/bottleneck.py, unmodified)" if kind == "reference" else "oracle port"
        cpu = {"value": done * ctok / csec, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"{what}: {done} of {chunks} row chunks of {CPU_CHUNK}x{n} tokens of the same batch, {csec:.1f} s"}

    cfgd = config_dict(args, world)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": cfgd, "value_is": value_kind,
        "eager": {"ms_per_step": ms_eager, "value": world * N / (ms_eager * 1e-3)},
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps},
        "gpu_launches": launches,
        "roofline": roofline,
        "amp_bf16": amp,
        "small_batch": small,
        "tokenizer": tok,
        "codebook_sharded": cb_sharded,
        "cpu_baseline": cpu,
        "clocks": sampler.result() if sampler else None,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        graphed = None            # a graph that holds a captured collective goes before its communicator
        torch.cuda.synchronize()
        dist.barrier()
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
