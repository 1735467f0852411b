"""Per-parameter gradient error of a golden fixture against the float64 oracle and the fp32 reference (GPU box)."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import npf_gwwaveform_b200 as npf
from oracle import npf_oracle as O
from test_gpu_models import load_golden, product_from_fixture, loss_class
tag = sys.argv[1] if len(sys.argv) > 1 else "attncnp_pretrained"
fix = load_golden(tag)
m = product_from_fixture(fix).train()
inp = {k: v.cuda() for k, v in fix["inputs"].items()}
crit = loss_class(tag, fix["name"])().train()
loss = crit(m(**inp), inp["Y_trgt"]); loss.backward()
ci = fix["inputs"]
g64 = O.loss_and_grads(O.FORWARDS[fix["name"].split("_")[0]], O.cnpf_loss, fix["state_dict"],
                       dict(x_c=ci["X_cntxt"], y_c=ci["Y_cntxt"], x_t=ci["X_trgt"]), ci["Y_trgt"], dtype=torch.float64, training=True)[2]
ref = fix["train"]["grads"]
gmax = max(float(g.abs().max()) for g in ref.values())
rows = []
for k, p in m.named_parameters():
    g = g64[k]; den = max(float(g.abs().max()), 1e-3 * gmax)
    rows.append((float((p.grad.cpu().double() - g).abs().max()) / den, float((ref[k].double() - g).abs().max()) / den, k))
rows.sort(reverse=True)
for e, er, k in rows[:8]:
    print(f"{k:45s} ours-vs-64 {e:.2e}  ref32-vs-64 {er:.2e}")
