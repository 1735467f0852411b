"""Pin the CPU oracle (oracle/npf_oracle.py) against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  Runs without a GPU."""
import pytest
import torch

from conftest import GOLDEN_TAGS, load_golden
from oracle import npf_oracle as O

# fp32 reference vs fp32 oracle: only re-association noise is allowed
RTOL, ATOL = 2e-5, 2e-6


def oracle_call(fix, training, with_y):
    name, kw = fix["name"], dict(fix["kwargs"])
    inp = fix["inputs"]
    args = dict(x_c=inp["X_cntxt"], y_c=inp["Y_cntxt"], x_t=inp["X_trgt"], training=training)
    extra = {}
    if name in ("LNP", "AttnLNP"):
        extra = dict(n_z=fix["n_z"], eps=fix["eps"])
        args["y_t"] = inp["Y_trgt"] if with_y else None
        if "encoded_path" in kw:
            extra["encoded_path"] = kw["encoded_path"]
    if name == "AttnCNP" and "attention" in kw:
        extra["attention"] = kw["attention"]
    fwd = O.FORWARDS[name if not name.startswith("ConvCNP") else "ConvCNP"]
    if name.startswith("ConvCNP"):
        extra = dict(density_induced=kw.get("density", 64), unet=name == "ConvCNP_unet")
    if name == "ConvLNP":
        extra = dict(n_z=fix["n_z"], eps=fix["eps"], density_induced=kw.get("density", 64),
                     encoded_path=kw.get("encoded_path", "latent"), is_global=kw.get("is_global", True),
                     is_q_zCct=kw.get("is_q_zCct", False))
        args["y_t"] = inp["Y_trgt"] if with_y else None
    return fwd, args, extra


def loss_for(fix):
    if fix["name"] in ("LNP", "AttnLNP", "ConvLNP"):
        return O.nll_loss if "nll" in fix.get("tag", "") else O.elbo_loss
    return O.cnpf_loss


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_oracle_matches_reference_train(tag):
    fix = load_golden(tag)
    fix["tag"] = tag
    fwd, args, extra = oracle_call(fix, training=True, with_y=True)
    loss_fn = loss_for(fix)
    if loss_fn is O.nll_loss:      # importance weights whenever q_zCct was inferred (losses.py:186-196)
        loss_fn = lambda out, y: O.nll_loss(out, y, importance="q_zCct_loc" in out)  # noqa: E731
    new_stats = {}
    if fix["name"].startswith("ConvCNP") or fix["name"] == "ConvLNP":
        extra["new_stats"] = new_stats
    inputs = {k: v for k, v in args.items() if torch.is_tensor(v)}
    rest = {k: v for k, v in args.items() if not torch.is_tensor(v)}
    out, loss, grads = O.loss_and_grads(fwd, loss_fn, fix["state_dict"], inputs, fix["inputs"]["Y_trgt"], **rest, **extra)
    ref = fix["train"]
    torch.testing.assert_close(out["loc"].detach(), ref["loc"], rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(out["scale"].detach(), ref["scale"], rtol=RTOL, atol=ATOL)
    lp = O.normal_log_prob(fix["inputs"]["Y_trgt"].unsqueeze(0), out["loc"], out["scale"]).sum(-1).detach()
    torch.testing.assert_close(lp, ref["log_prob"], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(loss, ref["loss"], rtol=1e-5, atol=1e-4)
    for k in ("z", "q_zCc_loc", "q_zCc_scale", "q_zCct_loc", "q_zCct_scale"):
        if k in ref:
            torch.testing.assert_close(out[k].detach(), ref[k], rtol=RTOL, atol=ATOL)
    assert set(grads) == set(ref["grads"])
    gmax = max(float(g.abs().max()) for g in ref["grads"].values())
    for k, g in ref["grads"].items():
        # gradients that are mathematically zero (e.g. a bias feeding train-mode BatchNorm) are pure noise
        scale = max(float(g.abs().max()), 1e-3 * gmax)
        err = float((grads[k] - g).abs().max()) / scale
        assert err < 5e-4, (k, err)
    for k, v in ref["state_after"].items():
        if k in new_stats:
            torch.testing.assert_close(new_stats[k].to(v.dtype), v, rtol=1e-5, atol=1e-6)
    if fix["name"].startswith("ConvCNP") or fix["name"] == "ConvLNP":
        assert set(new_stats) == set(ref["state_after"])


@pytest.mark.parametrize("tag", GOLDEN_TAGS)
def test_oracle_matches_reference_eval(tag):
    fix = load_golden(tag)
    fwd, args, extra = oracle_call(fix, training=False, with_y=False)
    with torch.no_grad():
        out = fwd(fix["state_dict"], **args, **extra)
        loss = O.nll_loss(out, fix["inputs"]["Y_trgt"], reduction=None)
    ref = fix["eval"]
    torch.testing.assert_close(out["loc"], ref["loc"], rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(out["scale"], ref["scale"], rtol=RTOL, atol=ATOL)
    torch.testing.assert_close(loss, ref["loss_per_task"], rtol=1e-5, atol=1e-4)


def test_oracle_fp64_is_consistent():
    fix = load_golden("attncnp_r32")
    inp = fix["inputs"]
    sd64 = {k: v.double() for k, v in fix["state_dict"].items()}
    out = O.attncnp_forward(sd64, inp["X_cntxt"].double(), inp["Y_cntxt"].double(), inp["X_trgt"].double())
    torch.testing.assert_close(out["loc"].float(), fix["eval"]["loc"], rtol=1e-4, atol=1e-5)


def test_oracle_empty_context():
    # Nc = 0 is legal and yields the zero representation (np.py:97-99, attnnp.py:108-125, convnp.py:146-151)
    for tag in ("cnp_r32", "attncnp_r32", "convcnp_r32"):
        fix = load_golden(tag)
        inp = fix["inputs"]
        B = inp["X_trgt"].shape[0]
        fwd, args, extra = oracle_call(fix, training=False, with_y=False)
        args["x_c"] = torch.zeros(B, 0, 1)
        args["y_c"] = torch.zeros(B, 0, 1)
        out = fwd(fix["state_dict"], **args, **extra)
        assert torch.isfinite(out["loc"]).all() and (out["scale"] > 0.01).all()


def test_oracle_range_check():
    fix = load_golden("cnp_r32")
    inp = fix["inputs"]
    with pytest.raises(ValueError):
        O.cnp_forward(fix["state_dict"], inp["X_cntxt"] * 3, inp["Y_cntxt"], inp["X_trgt"], training=True)
    O.cnp_forward(fix["state_dict"], inp["X_cntxt"] * 3, inp["Y_cntxt"], inp["X_trgt"], training=False)
