"""Per-parameter gradient error of ConvCNP_unet against the float64 oracle (run on the GPU box)."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import npf_gwwaveform_b200 as npf
from oracle import npf_oracle as O
from util_models import build_product, chirp_batch, perturb_
from test_gpu_models import ORACLE_KW
name, B, Nc, Nt = "ConvCNP_unet", int(sys.argv[1]) if len(sys.argv) > 1 else 2, 33, 129
torch.manual_seed(0)
m = build_product(name, 128, nz=3); perturb_(m, 7)
sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
m = m.cuda().train()
inp = chirp_batch(B, Nc, Nt, 99)
args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"])
kw = dict(ORACLE_KW.get(name, {}))
res = {}
for dt in (torch.float64, torch.float32):
    res[dt] = O.loss_and_grads(O.FORWARDS["ConvCNP"], O.cnpf_loss, sd, args, inp["Y_trgt"], dtype=dt, training=True, **kw)
out = m(**{k: v.cuda() for k, v in inp.items()})
loss = npf.CNPFLoss().train()(out, inp["Y_trgt"].cuda()); loss.backward()
g64, g32 = res[torch.float64][2], res[torch.float32][2]
gmax = max(float(g.abs().max()) for g in g64.values())
print("loss", float(loss), float(res[torch.float64][1]), "gmax", gmax)
for k, p in m.named_parameters():
    if k in g64:
        g = g64[k]; den = max(float(g.abs().max()), 1e-3 * gmax)
        e = float((p.grad.cpu().double() - g).abs().max()) / den
        e32 = float((g32[k].double() - g).abs().max()) / den
        flag = " <<<" if e > 1e-3 else ""
        print(f"{k:70s} |g|max {float(g.abs().max()):9.3e} ours {e:.2e} torch32 {e32:.2e}{flag}")
