"""Model-level parity of the BENCHMARKED arithmetic (``set_mode("fast")``: tcgen05 kind::tf32 operands, fp32
accumulate) through the drop-in module API on cuda:0, against

1. the committed golden vectors produced by the unmodified reference (tests/golden/*.pt, fp32 reference),
2. the float64 CPU oracle at notebook width (r_dim = 128), ragged sizes,
3. the float64 CPU oracle on chunks of the BASELINE.json shapes: 4 tasks of C2 (AttnCNP 512 / 4096), 2 tasks of C3
   (ConvCNP-UNet 512 / 4096) and 1 task of C4 (AttnLNP, 32 z-samples, epsilon injected) -- the sizes where the
   resident-K/V attention path, the persistent multi-tile GEMM loops and the hoisted SetConv with key skipping switch
   on.  The same chunks also run in ``fp32`` mode at the 1e-4 gate.

Norms (stated once, used everywhere; BASELINE.json north_star: "<=1e-3 relative error on the per-target
log-likelihood in bf16/TF32 mode"):

* ``norm``  = max|a - b| / max|b|                      -- the max-normalised error (tests/util_models.rel_err)
* ``point`` = max_t |lp_a - lp_b| / max(|lp_b|, 1)     -- the pointwise relative error of the per-target
  log-likelihood, with the denominator floored at 1 nat (a log-likelihood that crosses zero has no meaningful
  relative error; 1 nat is the scale of one target's contribution).

Gates in fast mode: per-target log-likelihood ``norm`` <= 1e-3 AND ``point`` <= 1e-3; predictive mean / std
``norm`` <= 1e-3; loss relative <= 1e-3; parameter gradients ``norm`` (denominator floored at 1e-3 of the largest
gradient of the model) <= GRAD_TOL.  A pre-activation within tf32 noise of a ReLU kink may land on the other side of
it; on these continuous-valued inputs that moves single gradient entries, not the forward gate.

Every case appends its measured errors to ``gpurun_out/parity_fast.jsonl`` when that directory exists (evidence kept
under profiles/ by the round's GPU pass).
"""
import json
import os

import pytest
import torch

from conftest import GOLDEN_TAGS, ROOT, load_golden
from oracle import npf_oracle as O
from util_models import build_product, chirp_batch, perturb_, rel_err

pytestmark = pytest.mark.gpu

LP_TOL = 1e-3        # per-target log-likelihood, both norms (north_star)
LOC_TOL = 1e-3       # predictive mean / std in fast mode
GRAD_TOL = 5e-3      # parameter gradients in fast mode
FP32_TOL = 1e-4      # fp32 mode on the same full-shape chunks


def _report(case, mode, **vals):
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "parity_fast.jsonl"), "a") as f:
            f.write(json.dumps(dict(case=case, mode=mode, **{k: (float(v) if not isinstance(v, str) else v)
                                                            for k, v in vals.items()})) + "\n")


def lp_point_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float(((a - b).abs() / b.abs().clamp_min(1.0)).max())


def grad_err(named_params, ref_grads):
    gmax = max(float(g.abs().max()) for g in ref_grads.values())
    worst = ("", 0.0)
    for k, p in named_params:
        if k not in ref_grads:
            continue
        g = ref_grads[k]
        assert p.grad is not None, k
        err = float((p.grad.detach().cpu().double() - g.double()).abs().max()) / max(float(g.abs().max()), 1e-3 * gmax)
        worst = max(worst, (k, err), key=lambda t: t[1])
    return worst


@pytest.fixture(autouse=True)
def _modes():
    import npf_gwwaveform_b200 as npf

    yield
    npf.set_mode("fp32")
    torch.cuda.synchronize()
    from npf_gwwaveform_b200._lib import LIB

    assert LIB.load().npfk_tc_abort_code() == 0, "a tcgen05 kernel gave up on an mbarrier wait"


def _loss_class(tag, name):
    import npf_gwwaveform_b200 as npf

    if name in ("LNP", "AttnLNP", "ConvLNP"):
        return npf.NLLLossLNPF if "nll" in tag else npf.ELBOLossLNPF
    return npf.CNPFLoss


# BatchNorm over a handful of tasks makes ConvLNP's gradient ill-conditioned even in fp32 (tests/test_gpu_models.py);
# the forward gates are the same for every tag
GOLDEN_GRAD_TOL = {"convlnp_both_r32": 5e-2, "convlnp_nll_r32": 2e-2}


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_golden_train_step_fast(tag):
    """fast mode against the vectors the unmodified reference produced (fp32 reference arithmetic)"""
    import npf_gwwaveform_b200 as npf

    fix = load_golden(tag)
    m = build_product(fix["name"], fix["r_dim"], nz=fix["n_z"], **fix["kwargs"])
    m.load_state_dict(fix["state_dict"], strict=True)
    m = m.cuda().train()
    if "eps" in fix:
        eps = fix["eps"].cuda()
        m.draw_eps = lambda n, like: eps
    inp = {k: v.cuda() for k, v in fix["inputs"].items()}
    ref = fix["train"]
    crit = _loss_class(tag, fix["name"])().train()
    npf.set_mode("fast")
    out = m(**inp)
    loss = crit(out, inp["Y_trgt"])
    loss.backward()
    p = out[0]
    lp = p.log_prob(inp["Y_trgt"])
    e = dict(loc=rel_err(p.base_dist.loc, ref["loc"]), scale=rel_err(p.base_dist.scale, ref["scale"]),
             lp_norm=rel_err(lp, ref["log_prob"]), lp_point=lp_point_err(lp, ref["log_prob"]),
             loss=abs(float(loss) - float(ref["loss"])) / max(1.0, abs(float(ref["loss"]))))
    gk, ge = grad_err(m.named_parameters(), ref["grads"])
    _report(f"golden:{tag}", "fast", grad=ge, grad_worst=gk, **e)
    assert e["loc"] < LOC_TOL and e["scale"] < LOC_TOL, e
    assert e["lp_norm"] < LP_TOL and e["lp_point"] < LP_TOL, e
    assert e["loss"] < LP_TOL, e
    assert ge < GOLDEN_GRAD_TOL.get(tag, GRAD_TOL), (gk, ge)


ORACLE_KW = {"ConvCNP_unet": dict(unet=True, density_induced=64), "ConvCNP": dict(density_induced=64)}


def _oracle_case(name, B, Nc, Nt, nz, seed, mode, perturb=True):
    """product model in `mode` vs the float64 oracle on the same seeded chirps; returns the error dict"""
    import npf_gwwaveform_b200 as npf

    torch.manual_seed(0)
    m = build_product(name, 128, nz=nz)
    if perturb:
        perturb_(m, 7)
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    m = m.cuda().train()
    inp = chirp_batch(B, Nc, Nt, seed)
    latent = "LNP" in name
    kw = dict(ORACLE_KW.get(name, {}))
    args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"])
    if latent:
        eps = torch.randn(nz, B, 1, 128, generator=torch.Generator().manual_seed(5))
        m.draw_eps = lambda n, like: eps.cuda()
        kw.update(n_z=nz, eps=eps)
        args["y_t"] = inp["Y_trgt"]
    loss_fn = O.elbo_loss if latent else O.cnpf_loss
    ref_out, ref_loss, ref_grads = O.loss_and_grads(O.FORWARDS[name.split("_")[0]], loss_fn, sd, args, inp["Y_trgt"],
                                                    dtype=torch.float64, training=True, **kw)
    ref_lp = O.normal_log_prob(inp["Y_trgt"].double(), ref_out["loc"], ref_out["scale"]).sum(-1)
    crit = (npf.ELBOLossLNPF if latent else npf.CNPFLoss)().train()
    npf.set_mode(mode)
    out = m(**{k: v.cuda() for k, v in inp.items()})
    loss = crit(out, inp["Y_trgt"].cuda())
    loss.backward()
    torch.cuda.synchronize()
    p = out[0]
    lp = p.log_prob(inp["Y_trgt"].cuda())
    e = dict(loc=rel_err(p.base_dist.loc, ref_out["loc"]), scale=rel_err(p.base_dist.scale, ref_out["scale"]),
             lp_norm=rel_err(lp, ref_lp), lp_point=lp_point_err(lp, ref_lp),
             loss=abs(float(loss) - float(ref_loss)) / max(1.0, abs(float(ref_loss))))
    gk, ge = grad_err(m.named_parameters(), ref_grads)
    e.update(grad=ge, grad_worst=gk)
    _report(f"oracle:{name}:B{B}:Nc{Nc}:Nt{Nt}:nz{nz}", mode, **e)
    return e


@pytest.mark.parametrize("name,B,Nc,Nt", [
    ("CNP", 4, 64, 128), ("AttnCNP", 3, 77, 301), ("ConvCNP", 2, 50, 200), ("ConvCNP_unet", 3, 33, 129),
    ("LNP", 2, 10, 40), ("AttnLNP", 2, 40, 150), ("AttnCNP", 2, 200, 1000),
])
def test_against_oracle_notebook_width_fast(name, B, Nc, Nt):
    e = _oracle_case(name, B, Nc, Nt, 3, 99, "fast")
    assert e["loc"] < LOC_TOL and e["scale"] < LOC_TOL, e
    assert e["lp_norm"] < LP_TOL and e["lp_point"] < LP_TOL, e
    assert e["loss"] < LP_TOL, e
    assert e["grad"] < GRAD_TOL, e


# chunks of the BASELINE.json configurations (C2, C3, C4 shapes)
FULL_SHAPE = [("AttnCNP", 4, 512, 4096, 1), ("ConvCNP_unet", 2, 512, 4096, 1), ("AttnLNP", 1, 512, 4096, 32)]


@pytest.mark.parametrize("name,B,Nc,Nt,nz", FULL_SHAPE)
def test_full_shape_chunk_fast(name, B, Nc, Nt, nz):
    e = _oracle_case(name, B, Nc, Nt, nz, 1234, "fast")
    assert e["loc"] < LOC_TOL and e["scale"] < LOC_TOL, e
    assert e["lp_norm"] < LP_TOL and e["lp_point"] < LP_TOL, e
    assert e["loss"] < LP_TOL, e
    assert e["grad"] < GRAD_TOL, e


@pytest.mark.parametrize("name,B,Nc,Nt,nz", FULL_SHAPE)
def test_full_shape_chunk_fp32(name, B, Nc, Nt, nz):
    e = _oracle_case(name, B, Nc, Nt, nz, 1234, "fp32")
    assert e["loc"] < FP32_TOL and e["scale"] < FP32_TOL, e
    assert e["lp_norm"] < FP32_TOL and e["lp_point"] < FP32_TOL, e
    assert e["loss"] < FP32_TOL, e
    assert e["grad"] < 1e-3, e
