"""Where does the fast-mode (tf32) error come from?  For a few golden fixtures / oracle cases prints the forward and the
per-parameter gradient errors against the fp64 CPU oracle of

  A  the product in fp32 mode
  B  the product in fp32 mode with every matrix weight rounded to tf32 (isolates the weight rounding)
  C  the product in fast mode
  D  the oracle itself on the GPU in fp32 with cuBLAS/cuDNN TF32 allowed (what stock PyTorch TF32 gives)

Developer tool (GPU box): python tools/fast_grad_debug.py [case ...]
"""
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from conftest import load_golden  # noqa: E402
from oracle import npf_oracle as O  # noqa: E402
from util_models import build_product, chirp_batch, perturb_, rel_err  # noqa: E402

import npf_gwwaveform_b200 as npf  # noqa: E402

ORACLE_KW = {"ConvCNP_unet": dict(unet=True, density_induced=64), "ConvCNP": dict(density_induced=64)}


def round_tf32_(sd):
    for k, v in sd.items():
        if k.endswith("weight") and v.dim() >= 2 and v.is_floating_point():
            b = v.contiguous().view(torch.int32)
            v.copy_(((b + 0x1000) & ~0x1FFF).view(torch.float32))


def grad_table(grads, ref, top=4):
    gmax = max(float(g.abs().max()) for g in ref.values())
    rows = []
    num = den = 0.0
    for k, g in ref.items():
        if k not in grads:
            continue
        a = grads[k].detach().double().cpu()
        b = g.double().cpu()
        e_max = float((a - b).abs().max()) / max(float(b.abs().max()), 1e-3 * gmax)
        e_l2 = float((a - b).norm()) / max(float(b.norm()), 1e-30)
        num += float((a - b).norm()) ** 2
        den += float(b.norm()) ** 2
        rows.append((e_max, e_l2, float(b.abs().max()) / gmax, k))
    rows.sort(reverse=True)
    return (num / den) ** 0.5, rows[:top]


def product_run(name, nz, sd, inp, eps, mode, latent, kwargs=None):
    m = build_product(name, sd["__r__"], nz=nz, **(kwargs or {}))
    m.load_state_dict({k: v for k, v in sd.items() if k != "__r__"}, strict=True)
    m = m.cuda().train()
    if eps is not None:
        m.draw_eps = lambda n, like: eps.cuda()
    crit = (npf.ELBOLossLNPF if latent else npf.CNPFLoss)().train()
    npf.set_mode(mode)
    out = m(**{k: v.cuda() for k, v in inp.items()})
    loss = crit(out, inp["Y_trgt"].cuda())
    loss.backward()
    torch.cuda.synchronize()
    npf.set_mode("fp32")
    p = out[0]
    return dict(loc=p.base_dist.loc.detach().cpu(), lp=p.log_prob(inp["Y_trgt"].cuda()).detach().cpu(),
                grads={k: v.grad.detach().cpu() for k, v in m.named_parameters() if v.grad is not None})


def oracle_run(name, sd, inp, eps, nz, latent, dtype, device, tf32=False):
    kw = dict(ORACLE_KW.get(name, {}))
    args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"])
    if latent:
        kw.update(n_z=nz, eps=eps.to(device))
        args["y_t"] = inp["Y_trgt"]
    args = {k: v.to(device) for k, v in args.items()}
    sd = {k: v.to(device) for k, v in sd.items() if k != "__r__"}
    torch.backends.cuda.matmul.allow_tf32 = tf32
    torch.backends.cudnn.allow_tf32 = tf32
    out, loss, grads = O.loss_and_grads(O.FORWARDS[name.split("_")[0]], O.elbo_loss if latent else O.cnpf_loss, sd, args,
                                        inp["Y_trgt"].to(device), dtype=dtype, training=True, **kw)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    lp = O.normal_log_prob(inp["Y_trgt"].to(device).to(dtype), out["loc"], out["scale"]).sum(-1)
    return dict(loc=out["loc"].detach().cpu(), lp=lp.detach().cpu(), grads={k: g.cpu() for k, g in grads.items()})


def case(name, B, Nc, Nt, nz, R=128, tag=None, seed=99):
    torch.manual_seed(0)
    latent = "LNP" in name
    kwargs = None
    if tag:
        fix = load_golden(tag)
        sd = dict(fix["state_dict"])
        R, nz, kwargs = fix["r_dim"], fix["n_z"], fix["kwargs"]
        inp = fix["inputs"]
        eps = fix.get("eps")
        name = fix["name"]
        latent = "LNP" in name
    else:
        m = build_product(name, R, nz=nz)
        perturb_(m, 7)
        sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
        inp = chirp_batch(B, Nc, Nt, seed)
        eps = torch.randn(nz, B, 1, R, generator=torch.Generator().manual_seed(5)) if latent else None
    sd["__r__"] = R
    if tag and (kwargs or name not in O.FORWARDS):
        print(f"--- {tag}: skipped (kwargs {kwargs})")
        return
    ref = oracle_run(name, sd, inp, eps, nz, latent, torch.float64, "cpu")
    sd_r = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in sd.items()}
    round_tf32_(sd_r)
    runs = dict(A_fp32=product_run(name, nz, sd, inp, eps, "fp32", latent),
                B_fp32_roundedW=product_run(name, nz, sd_r, inp, eps, "fp32", latent),
                C_fast=product_run(name, nz, sd, inp, eps, "fast", latent),
                D_torch_tf32=oracle_run(name, sd, inp, eps, nz, latent, torch.float32, "cuda", tf32=True))
    print(f"=== {tag or name} B{B} Nc{Nc} Nt{Nt} nz{nz} R{R}")
    for k, r in runs.items():
        g_all, rows = grad_table(r["grads"], ref["grads"])
        lp_pt = float(((r["lp"].double() - ref["lp"]).abs() / ref["lp"].abs().clamp_min(1.0)).max())
        print(f"  {k:16s} loc {rel_err(r['loc'], ref['loc']):.1e} lp {rel_err(r['lp'], ref['lp']):.1e} lp_pt {lp_pt:.1e} "
              f"grad_all_l2 {g_all:.1e}")
        for e_max, e_l2, mag, pk in rows:
            print(f"        max {e_max:.1e} l2 {e_l2:.1e} |g|/gmax {mag:.1e} {pk}")


def main():
    case("CNP", 0, 0, 0, 0, tag="cnp_r32")
    case("AttnCNP", 0, 0, 0, 0, tag="attncnp_r32")
    case("ConvCNP", 0, 0, 0, 0, tag="convcnp_r32")
    case("LNP", 0, 0, 0, 0, tag="lnp_r32")
    case("AttnLNP", 0, 0, 0, 0, tag="attnlnp_r32")
    case("CNP", 4, 64, 128, 3)
    case("AttnCNP", 3, 77, 301, 3)
    case("AttnLNP", 2, 40, 150, 3)
    case("ConvCNP", 2, 50, 200, 3)
    case("ConvCNP_unet", 2, 512, 4096, 1, seed=1234)
    case("AttnCNP", 4, 512, 4096, 1, seed=1234)
    from npf_gwwaveform_b200._lib import LIB
    print("abort code", LIB.load().npfk_tc_abort_code())


if __name__ == "__main__":
    main()
