"""The reference-side binding printed in INTEGRATION.md (section B) is executed LITERALLY: the ```python block that starts
with ``# npf/_npfk.py`` is cut out of the document, pointed at the in-tree libnpfk.so and exec'd, and its
``MultiheadScaledDot`` autograd Function is checked against the oracle's restatement of the reference's multi-head
attention (npf/architectures/attention.py:507-527, 204-220, 129-164) in both arithmetic modes.  If the document and the
C-ABI drift apart, this fails."""
import math
import os
import re

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _stub_namespace():
    with open(os.path.join(ROOT, "INTEGRATION.md")) as f:
        text = f.read()
    blocks = re.findall(r"```python\n(.*?)```", text, flags=re.S)
    stub = next(b for b in blocks if b.lstrip().startswith("# npf/_npfk.py"))
    stub = stub.replace("/path/to/libnpfk.so", os.path.join(ROOT, "npf_gwwaveform_b200", "libnpfk.so"))
    ns = {}
    exec(compile(stub, "INTEGRATION.md:_npfk.py", "exec"), ns)
    return ns


@pytest.mark.parametrize("mode,tol", [(0, 1e-5), (1, 3e-3)])
def test_integration_stub_attention_matches_oracle(mode, tol):
    from oracle import npf_oracle as O

    ns = _stub_namespace()
    g = torch.Generator().manual_seed(3)
    B, Nq, Nk, H, D = 2, 300, 77, 8, 16
    q, k, v = (torch.randn(B, n, H * D, generator=g) for n in (Nq, Nk, Nk))
    dout = torch.randn(B, Nq, H * D, generator=g)
    # oracle: the reference's head split + scaled dot + softmax + bmm + head merge in float64
    qs, ks, vs = (t.double().requires_grad_(True) for t in (q, k, v))
    qh, kh, vh = (t.view(B, -1, H, D).transpose(1, 2) for t in (qs, ks, vs))
    ref = (torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(D), -1) @ vh).transpose(1, 2).reshape(B, Nq, H * D)
    ref.backward(dout.double())
    assert hasattr(O, "multihead_attender")      # the restatement above is the oracle's (oracle/npf_oracle.py:88-112)
    qd, kd, vd = (t.cuda().requires_grad_(True) for t in (q, k, v))
    out = ns["MultiheadScaledDot"].apply(qd, kd, vd, H, mode)
    out.backward(dout.cuda())
    torch.cuda.synchronize()
    rel = lambda a, b: float((a.detach().cpu().double() - b.detach()).norm() / b.detach().norm())  # noqa: E731
    assert rel(out, ref) < tol
    for a, b in ((qd.grad, qs.grad), (kd.grad, ks.grad), (vd.grad, vs.grad)):
        assert rel(a, b) < 2 * tol
