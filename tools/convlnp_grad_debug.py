"""Developer probe: per-parameter gradient error of ConvLNP (encoded_path="both") against the fp64 oracle."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import npf_gwwaveform_b200 as npf
from oracle import npf_oracle as O
from util_models import build_product, chirp_batch, perturb_, rel_err
B, Nc, Nt = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (3, 30, 77)
over = dict(encoded_path="both", is_global=False, is_q_zCct=True)
torch.manual_seed(0)
m = build_product("ConvLNP", 128, nz=3, **over); perturb_(m, 7)
sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
m = m.cuda().train()
inp = chirp_batch(B, Nc, Nt, 99)
eps = torch.randn(3, B, 1, 128, generator=torch.Generator().manual_seed(5))
m.draw_eps = lambda n, like: eps.cuda()
kw = dict(density_induced=64, encoded_path="both", is_global=False, is_q_zCct=True, n_z=3, eps=eps)
args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"], y_t=inp["Y_trgt"])
for dt in (torch.float64, torch.float32):
    ref_out, ref_loss, ref_grads = O.loss_and_grads(O.convlnp_forward, O.elbo_loss, sd, args, inp["Y_trgt"], dtype=dt, training=True, **kw)
    if dt == torch.float64: g64 = ref_grads
crit = npf.ELBOLossLNPF().train()
out = m(**{k: v.cuda() for k, v in inp.items()})
loss = crit(out, inp["Y_trgt"].cuda()); loss.backward()
print("loss", float(loss), float(ref_loss), "loc", rel_err(out[0].base_dist.loc, ref_out["loc"]))
gmax = max(float(g.abs().max()) for g in g64.values())
rows = []
for k, p in m.named_parameters():
    g = g64[k]; sc = max(float(g.abs().max()), 1e-3 * gmax)
    e_ours = float((p.grad.cpu().double() - g).abs().max()) / sc
    e_f32 = float((ref_grads[k].double() - g).abs().max()) / sc
    rows.append((e_ours, e_f32, k))
for e, f, k in sorted(rows, reverse=True)[:12]: print(f"{e:.2e} (fp32 oracle {f:.2e})  {k}")
