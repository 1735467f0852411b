"""fast (tf32 tensor-core) mode vs fp32 mode of the same model on the GPU: prints normalised errors."""
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from conftest import load_golden  # noqa: E402
from util_models import build_product, chirp_batch, perturb_, rel_err  # noqa: E402

import npf_gwwaveform_b200 as npf  # noqa: E402


def run(m, inp, crit, mode):
    npf.set_mode(mode)
    m.zero_grad(set_to_none=True)
    out = m(**inp)
    loss = crit(out, inp["Y_trgt"])
    loss.backward()
    p = out[0]
    return dict(loc=p.base_dist.loc.detach(), scale=p.base_dist.scale.detach(),
                lp=p.log_prob(inp["Y_trgt"]).detach(), loss=float(loss),
                grads={k: v.grad.detach().clone() for k, v in m.named_parameters() if v.grad is not None})


def main():
    cases = [("AttnCNP", "attncnp_pretrained", 4, 512, 4096), ("AttnCNP", "attncnp_pretrained", 4, 64, 1024), ("AttnCNP", None, 4, 512, 4096),
             ("CNP", None, 8, 64, 512), ("ConvCNP", "convcnp_pretrained", 4, 512, 4096), ("ConvCNP_unet", None, 4, 512, 4096),
             ("AttnLNP", None, 2, 512, 1024)]
    for name, tag, B, Nc, Nt in cases:
        torch.manual_seed(0)
        m = build_product(name, 128, nz=4)
        if tag:
            m.load_state_dict(load_golden(tag)["state_dict"], strict=True)
        else:
            perturb_(m, 7)
        m = m.cuda().train()
        if "LNP" in name:
            eps = torch.randn(4, B, 1, 128, device="cuda")
            m.draw_eps = lambda n, like: eps
        inp = chirp_batch(B, Nc, Nt, 3, "cuda")
        if tag and "smooth" in sys.argv:
            # in-distribution-like inputs for the pretrained GP-RBF models: smooth random functions
            g = torch.Generator().manual_seed(5)
            x = inp["X_trgt"].cpu()
            y = sum(torch.randn(B, 1, 1, generator=g) * torch.sin(f * x * 3.1416 + 6.28 * torch.rand(B, 1, 1, generator=g))
                    for f in (0.5, 1.0, 1.7, 2.6))
            idx = torch.stack([torch.randperm(Nt, generator=g)[:Nc] for _ in range(B)]).unsqueeze(-1)
            inp = {k: v.cuda() for k, v in dict(X_cntxt=x.gather(1, idx), Y_cntxt=y.gather(1, idx), X_trgt=x, Y_trgt=y).items()}
        crit = (npf.ELBOLossLNPF if "LNP" in name else npf.CNPFLoss)().train()
        a = run(m, inp, crit, "fp32")
        b = run(m, inp, crit, "fast")
        lp_rel = float(((a["lp"] - b["lp"]).abs() / a["lp"].abs().clamp_min(1.0)).max())
        gw = max(((k, rel_err(b["grads"][k], a["grads"][k])) for k in a["grads"]), key=lambda t: t[1])
        print(f"{name:14s} {str(tag):20s} loc {rel_err(b['loc'], a['loc']):.2e} scale {rel_err(b['scale'], a['scale']):.2e} "
              f"lp(norm) {rel_err(b['lp'], a['lp']):.2e} lp(pointwise,|lp|>=1) {lp_rel:.2e} loss {a['loss']:.4f} {b['loss']:.4f} "
              f"worst grad {gw[0]} {gw[1]:.2e}", flush=True)
    torch.cuda.synchronize()
    from npf_gwwaveform_b200._lib import LIB
    print("abort code", LIB.load().npfk_tc_abort_code())


if __name__ == "__main__":
    main()
