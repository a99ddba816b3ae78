"""GPU: a whole stochastic Bottleneck step (forward + backward) captured in ONE CUDA graph.

The reference's contract (SURVEY.md 8b "Threading"): the extension only enqueues on the current stream, never reads
device data on the host.  What capture additionally needs is an RNG state that lives on the device: replays must
draw fresh noise, and the same noise the eager path draws from the same torch generator state
(models/bottleneck.py:275-280 consumes the generator through torch.multinomial)."""
import pytest
import torch

import video_tokenizer_b200 as V
from video_tokenizer_b200 import _lib
from tests.helpers import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _module(D=24, K=8192, C=768, n=1024, stochastic=True):
    torch.manual_seed(0)
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0, "entropy_loss_weight": 0.0,
        "l2_normalized": True, "stochastic": stochastic, "stochastic_temperature": 0.03}}}}
    m = V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(DEV)
    m.train()
    return m


def _eager_steps(m, x, g, steps, seed):
    m.regularizer.disable_device_rng()
    torch.cuda.manual_seed(seed)
    outs = []
    for _ in range(steps):
        for p in m.parameters():
            p.grad = None
        xin = x.detach().clone().requires_grad_(True)
        out = m(xin)
        torch.autograd.backward([out["output"], out["loss_q"]], [g, torch.tensor(0.1, device=DEV)])
        outs.append((out["bottleneck_rep"].clone(), out["output"].detach().clone(), float(out["loss_q"]),
                     xin.grad.clone(), {k: p.grad.clone() for k, p in m.named_parameters()}))
    return outs


@pytest.mark.parametrize("B", [8, 64])
def test_captured_stochastic_step_replays_with_fresh_noise_equal_to_eager(B):
    m = _module()
    torch.manual_seed(1)
    x = torch.randn(B, 1024, 768, device=DEV)
    g = torch.randn(B, 1024, 768, device=DEV)
    eager = _eager_steps(m, x, g, 3, seed=77)

    step = V.GraphedBottleneckStep(m, x, g, loss_q_grad=0.1)
    torch.cuda.manual_seed(77)
    step.sync_rng()
    lib = _lib.load()
    torch.cuda.synchronize()
    calls0 = lib.b200vq_launch_count()
    draws = []
    for i in range(3):
        out = step.replay()
        torch.cuda.synchronize()
        draws.append(out["bottleneck_rep"].clone())
        # (ii) the replay is the eager step taken from the same generator state
        assert torch.equal(draws[i], eager[i][0]), f"replay {i}: {(draws[i] != eager[i][0]).sum()} indices differ from eager"
        assert_close(out["output"], eager[i][1], "output")
        assert float(out["loss_q"]) == pytest.approx(eager[i][2], rel=1e-6)
        assert_close(step.x.grad, eager[i][3], "grad x")
        for k, p in m.named_parameters():
            assert_close(p.grad, eager[i][4][k], f"grad {k}", tol=2e-3 if "embedding" in k else 1e-3)
    # (iii) no host call into the library during the replays
    assert lib.b200vq_launch_count() == calls0
    # (i) every replay drew new noise
    assert not torch.equal(draws[0], draws[1]) and not torch.equal(draws[1], draws[2])
    # new inputs go through the static buffers
    x2 = torch.randn_like(x)
    out = step.replay(x2, g)
    torch.cuda.synchronize()
    assert_close(out["projected_z"], torch.nn.functional.linear(x2, m.in_linear.weight, m.in_linear.bias), "projected_z", tol=1e-5)
