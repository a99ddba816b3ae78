"""BASELINE configs[1]: the drop-in inside the UNMODIFIED reference `LARPTokenizer` (models/larp_tokenizer.py).

The reference tokenizer is imported through `oracle/ref_tokenizer.py` (stub shells for timm / easydict / vjepa2 /
flash_attn; the reference's own registry, bottleneck, transformer and tokenizer files, unmodified).  The drop-in is
switched in with `video_tokenizer_b200.install(models)` -- the three-line binding INTEGRATION.md shows -- so that
`models.make(bottleneck, ...)` at models/larp_tokenizer.py:214-218 builds the B200 modules.

CPU part: registry swap, constructor / state-dict contract, strict checkpoint load.
GPU part (`-m gpu`; needs baseline/_ref, staged by `python -m oracle.stage_ref`): encode -> quantize -> decode,
forward + backward, against the reference bottleneck evaluated on the identical encoder output.
"""
import copy

import pytest
import torch

import video_tokenizer_b200 as V
from oracle import ref_tokenizer as RT
from oracle import vq_oracle as O
from tests.helpers import assert_close, assert_indices_match

HAVE_REF = RT.reference_root() is not None
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="no reference tree (run `python -m oracle.stage_ref` in the build container)")


@pytest.fixture()
def ref_pkg():
    pkg = RT.load_tokenizer_package()
    saved = dict(pkg.models)
    yield pkg
    pkg.models.clear()
    pkg.models.update(saved)   # undo install() for the next test


@needs_ref
def test_install_swaps_the_reference_registry_and_keeps_the_checkpoint_contract(ref_pkg):
    ref_cls = dict(ref_pkg.models)
    ref_tok = RT.build_tokenizer(ref_pkg, depth=1)
    assert type(ref_tok.bottleneck) is ref_cls["bottleneck"]
    V.install(ref_pkg)
    assert ref_pkg.models["vq"] is V.SimpleVectorQuantizer and ref_pkg.models["bottleneck"] is V.Bottleneck
    assert ref_pkg.models["larp_tokenizer"] is ref_cls["larp_tokenizer"]          # nothing else touched
    tok = RT.build_tokenizer(ref_pkg, depth=1)
    assert type(tok.bottleneck) is V.Bottleneck and type(tok.bottleneck.regularizer) is V.SimpleVectorQuantizer
    assert tok.codebook_size == 8192 and tok.bottleneck_dim == 24
    # identical state-dict keys and shapes; a reference-built checkpoint loads strictly (models/larp_tokenizer.py:397)
    sd_ref = ref_tok.state_dict()
    sd = tok.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    assert all(sd[k].shape == sd_ref[k].shape and sd[k].dtype == sd_ref[k].dtype for k in sd)
    tok.load_state_dict(copy.deepcopy(sd_ref), strict=True)
    assert torch.equal(tok.bottleneck.regularizer.embedding.weight, ref_tok.bottleneck.regularizer.embedding.weight)
    # mode switches the tokenizer / trainers use (models/larp_tokenizer.py:333-334)
    tok.set_vq_eval_deterministic(True)
    assert tok.bottleneck.regularizer.eval_deterministic is True
    # the reference's from_checkpoint path: kwargs + strict state dict
    spec = {"name": "bottleneck", "args": RT.bottleneck_spec()["args"], "sd": ref_tok.bottleneck.state_dict()}
    b = ref_pkg.make(spec, args={"token_nums": 1024, "input_dim": 768, "output_dim": 768}, load_sd=True)
    assert type(b) is V.Bottleneck


def _pair(ref_pkg, depth, stochastic_eval_det=True):
    """(reference tokenizer on CPU, drop-in tokenizer on cuda:0) with identical weights"""
    ref_tok = RT.build_tokenizer(ref_pkg, depth=depth)
    V.install(ref_pkg)
    tok = RT.build_tokenizer(ref_pkg, depth=depth)
    tok.load_state_dict(ref_tok.state_dict(), strict=True)
    tok = tok.to("cuda")
    for t in (ref_tok, tok):
        t.eval()                                  # no dropout anywhere; eval + eval_deterministic -> argmax(cos), :275-280
        t.set_vq_eval_deterministic(stochastic_eval_det)
    return ref_tok, tok


@needs_ref
@pytest.mark.gpu
def test_tokenizer_forward_backward_with_dropin_matches_reference_bottleneck(ref_pkg):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref_tok, tok = _pair(ref_pkg, depth=2)
    torch.manual_seed(3)
    clip = torch.rand(2, 3, 16, 128, 128)

    # ---- drop-in tokenizer on the GPU: capture what the encoder hands the bottleneck (a batch-strided slice,
    # models/transformer.py:69) and the gradient that comes back to it
    seen = {}

    def grab(_m, args):
        x = args[0]
        seen["x"] = x
        seen["contiguous"] = x.is_contiguous()
        x.register_hook(lambda g: seen.__setitem__("gx", g))

    h = tok.bottleneck.register_forward_pre_hook(grab)
    out = tok(clip.cuda())
    h.remove()
    assert not seen["contiguous"], "the encoder's slice is expected to reach the bottleneck non-contiguous"
    loss = out["pred_frames"].mean() * 100.0 + 0.1 * out["loss_q"]       # trainer weighting, larp_tokenizer_trainer.py:330-333
    loss.backward()
    assert out["bottleneck_rep"].dtype == torch.int64 and tuple(out["bottleneck_rep"].shape) == (2, 1024)
    assert out["input_norm_first"].numel() == 1 and out["loss_q"].dim() == 0

    # ---- reference bottleneck + decoder on the CPU from the IDENTICAL encoder output
    x_ref = seen["x"].detach().cpu().clone().requires_grad_(True)
    rb = ref_tok.bottleneck(x_ref)
    enc_ref = rb["output"]
    pred_ref = ref_tok.decode(enc_ref)
    loss_ref = pred_ref.mean() * 100.0 + 0.1 * rb["loss_q"]
    loss_ref.backward()

    # indices: bit-exact outside the 1e-6 window of the reference's own ranking (cosine: eval-deterministic mode)
    w = ref_tok.bottleneck.regularizer.embedding.weight.detach()
    cfg = O.VQConfig(dim=24, codebook_size=8192, l2_normalized=True, stochastic=True, stochastic_temperature=0.03)
    ref_again = O.vq_forward(rb["projected_z"].detach(), w, cfg, training=False, eval_deterministic=True, return_scores=True)
    assert torch.equal(ref_again["bottleneck_rep"], rb["bottleneck_rep"])
    flips = assert_indices_match(out["bottleneck_rep"], rb["bottleneck_rep"], ref_again["scores"], True)
    for k in ("projected_z", "unregularized_z", "loss_q", "loss_commit", "loss_codebook"):
        assert_close(out[k].detach().cpu(), rb[k].detach(), k)
    if flips == 0:
        for k in ("regularized_z", "encoded"):
            assert_close(out[k].detach().cpu(), (enc_ref if k == "encoded" else rb[k]).detach(), k)
        assert_close(out["pred_frames"].detach().cpu(), pred_ref.detach(), "pred_frames")
        assert_close(seen["gx"].cpu(), x_ref.grad, "grad wrt the encoder output")
        for (name, p), (_, q) in zip(tok.bottleneck.named_parameters(), ref_tok.bottleneck.named_parameters()):
            assert p.grad is not None, name
            assert_close(p.grad.cpu(), q.grad, f"grad {name}", tol=2e-3 if "embedding" in name else 1e-3)

    # ---- decode_from_bottleneck(indices) (models/larp_tokenizer.py:484-487)
    with torch.no_grad():
        dec = tok.decode_from_bottleneck(out["bottleneck_rep"])
        dec_ref = ref_tok.decode_from_bottleneck(out["bottleneck_rep"].cpu())
    assert_close(dec.cpu(), dec_ref, "decode_from_bottleneck")

    # ---- whole tokenizer end to end on both devices: the encoders differ by float rounding (CPU vs GPU GEMMs), so
    # only report agreement -- the strict comparison is the one above on identical latents
    with torch.no_grad():
        full = ref_tok(clip)
    agree = float((full["bottleneck_rep"] == out["bottleneck_rep"].cpu()).float().mean())
    assert agree > 0.97, f"only {agree:.3f} of the indices agree end to end"


@needs_ref
@pytest.mark.gpu
def test_tokenizer_training_mode_stochastic_step_runs_and_is_distribution_consistent(ref_pkg):
    """train() mode = stochastic SVQ sampling (models/bottleneck.py:275-280): shapes, finite grads on every
    bottleneck parameter, and the sampled codes are plausible under the reference's softmax (mean probability of
    the drawn code far above uniform)."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref_tok, tok = _pair(ref_pkg, depth=1)
    tok.train()
    torch.manual_seed(4)
    torch.cuda.manual_seed(4)
    clip = torch.rand(2, 3, 16, 128, 128, device="cuda")
    out = tok(clip)
    (out["pred_frames"].mean() + 0.1 * out["loss_q"]).backward()
    for name, p in tok.bottleneck.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
    z = out["unregularized_z"].detach().reshape(-1, 24)
    emb = out["emb"].detach()
    probs = torch.softmax(z @ emb.t() / 0.03, dim=-1)
    drawn = probs.gather(1, out["bottleneck_rep"].reshape(-1, 1)).mean()
    assert float(drawn) > 50.0 / 8192
    out2 = tok(clip)
    assert not torch.equal(out["bottleneck_rep"], out2["bottleneck_rep"])   # the generator offset advanced
