"""CUDA-graph training step (FlatAdamTrainer.capture / step_graphed) against the eager step of the same trainer.

The captured step records forward + loss + backward with the weight-gradient kernels adding straight into the flat
gradient buffer (kernels.set_direct_grad); the eager step goes through autograd's AccumulateGrad.  Both must land on
the same losses, gradients, parameters and BatchNorm buffers (up to the order of float atomics)."""
import copy

import pytest
import torch

from util_models import build_product, chirp_batch

pytestmark = pytest.mark.gpu


LR = 1e-6      # Adam turns rounding noise on exactly-zero gradients (a bias in front of BatchNorm) into +-lr steps, so the
               # two runs are kept on (almost) the same trajectory and the gradients themselves are compared


def _run(model, batches, n_eager):
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.training import FlatAdamTrainer

    tr = FlatAdamTrainer(model, npf.CNPFLoss().train(), lr=LR)
    losses = []
    for i, b in enumerate(batches):
        if i == n_eager:
            tr.capture(b)
        if i >= n_eager:
            losses.append(float(tr.step_graphed(b).clone()))
        else:
            losses.append(float(tr.step(b)))
    torch.cuda.synchronize()
    return losses, {n: p.grad.detach().double().clone() for n, p in model.named_parameters()}


@pytest.mark.parametrize("name,R,mode", [("CNP", 32, "fp32"), ("ConvCNP", 32, "fp32"), ("ConvCNP_unet", 32, "fp32"),
                                         ("AttnCNP", 32, "fp32"), ("ConvCNP", 128, "fast"), ("AttnCNP", 128, "fast")])
def test_graphed_step_matches_eager(name, R, mode):
    import npf_gwwaveform_b200 as npf

    npf.set_mode(mode)
    try:
        torch.manual_seed(3)
        m_eager = build_product(name, R).cuda().train()
        m_graph = copy.deepcopy(m_eager)
        batches = [{k: v.cuda() for k, v in chirp_batch(4, 24, 160, s).items()} for s in range(5)]
        l_eager, g_eager = _run(m_eager, batches, n_eager=len(batches))
        l_graph, g_graph = _run(m_graph, batches, n_eager=2)
        tol = 2e-4 if mode == "fp32" else 3e-2      # tf32 operands re-truncate activations that differ in the last bits;
        # a wrong accumulation (double counting, a missed term) shows up at order 1
        for a, b in zip(l_eager, l_graph):
            assert abs(a - b) <= tol * max(1.0, abs(a)), (l_eager, l_graph)
        gmax = max(float(g.abs().max()) for g in g_eager.values())
        assert gmax > 0
        for k in g_eager:          # last step's gradients, relative to the largest gradient entry of the model
            err = float((g_eager[k] - g_graph[k]).abs().max()) / gmax
            assert err < tol, (k, err)
        sd_e, sd_g = m_eager.state_dict(), m_graph.state_dict()
        for k in sd_e:
            a, b = sd_e[k].double(), sd_g[k].double()
            if k.endswith("num_batches_tracked"):
                assert int(a) == int(b) == len(batches), k
                continue
            err = float((a - b).abs().max() / (a.abs().max() + 1e-3))
            assert err < tol, (k, err)
    finally:
        npf.set_mode("fp32")


def test_eager_fast_steps_see_updated_weights():
    """fast mode keeps rounded (tf32) copies of the weights; Adam updates the flat buffer through raw pointers, so the
    copies must be dropped after every step: the trained module's forward has to match a fresh copy of itself."""
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.training import FlatAdamTrainer

    npf.set_mode("fast")
    try:
        torch.manual_seed(5)
        m = build_product("CNP", 128).cuda().train()
        batches = [{k: v.cuda() for k, v in chirp_batch(4, 24, 160, s).items()} for s in range(4)]
        tr = FlatAdamTrainer(m, npf.CNPFLoss().train(), lr=5e-2)      # big steps: stale copies would be far off
        for b in batches[:3]:
            tr.step(b)
        m.eval()
        with torch.no_grad():
            ours = m(**batches[3])[0].mean.clone()
            fresh = copy.deepcopy(m)(**batches[3])[0].mean
        assert float((ours - fresh).abs().max()) <= 1e-5 * max(1.0, float(fresh.abs().max()))
    finally:
        npf.set_mode("fp32")


def test_prefetched_batches_match_direct_copies():
    """step_graphed fed through prefetch() (H2D on a side stream, overlapping the previous step) sees the same data as
    the direct copy: different batches every step, identical losses."""
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200.training import FlatAdamTrainer

    torch.manual_seed(7)
    m0 = build_product("CNP", 32).cuda().train()
    host = [{k: v.pin_memory() for k, v in chirp_batch(4, 24, 160, s).items()} for s in range(6)]
    losses = []
    for use_prefetch in (False, True):
        m = copy.deepcopy(m0)
        tr = FlatAdamTrainer(m, npf.CNPFLoss().train(), lr=LR)
        dev = [{k: v.cuda() for k, v in b.items()} for b in host[:2]]
        tr.step(dev[0])
        tr.step(dev[1])
        tr.capture(dev[1])
        out = []
        if use_prefetch:
            tr.prefetch(host[2])
        for i in range(2, 6):
            loss = tr.step_graphed(host[i])
            if use_prefetch and i + 1 < 6:
                tr.prefetch(host[i + 1])
            out.append(float(loss.clone()))
        losses.append(out)
    assert len(set(losses[0])) > 1                     # the batches really differ
    for a, b in zip(*losses):
        assert abs(a - b) <= 1e-5 * max(1.0, abs(a)), losses
