"""Host-side logic of the data-parallel training step on CPU: world_size-2 gloo processes.

The CUDA kernels cannot run here, so the trainer is driven with a small torch module and a torch Adam stand-in;
what is under test is the flat-buffer layout, bucketed all-reduce issued from the autograd hooks, the
calibration re-layout and the equivalence with single-process training on the concatenated batch.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from npf_gwwaveform_b200.training import FlatAdamTrainer, shard_range


class Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(3, 16)
        self.b = torch.nn.Linear(16, 16)
        self.c = torch.nn.Linear(16, 1)

    def forward(self, X_cntxt, Y_cntxt, X_trgt, Y_trgt=None):
        ctx = torch.tanh(self.a(torch.cat((X_cntxt, Y_cntxt, X_cntxt ** 2), -1))).mean(1, keepdim=True)
        h = torch.tanh(self.b(ctx + self.a(torch.cat((X_trgt, X_trgt * 0, X_trgt ** 2), -1))))
        return self.c(h)


def crit(out, y):
    return ((out - y) ** 2).sum((1, 2)).mean(0)


def torch_adam(p, g, m, v, lr, b1, b2, eps, step, grad_scale):
    g = g * grad_scale
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    p.sub_((lr / (1 - b1 ** step)) * m / (v.sqrt() / (1 - b2 ** step) ** 0.5 + eps))


def make_batch(n, seed):
    g = torch.Generator().manual_seed(seed)
    return dict(X_cntxt=torch.rand(n, 5, 1, generator=g), Y_cntxt=torch.randn(n, 5, 1, generator=g),
                X_trgt=torch.rand(n, 7, 1, generator=g), Y_trgt=torch.randn(n, 7, 1, generator=g))


def worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = Toy()
    tr = FlatAdamTrainer(model, crit, lr=1e-2, n_buckets=2, adam_impl=torch_adam)
    full = make_batch(8, 1)
    lo, hi = shard_range(8, rank, world)
    shard = {k: v[lo:hi] for k, v in full.items()}
    for _ in range(4):
        tr.step(shard)
    assert tr._order is not None and len(tr.bucket_slices) == 2
    # every parameter is a view into the flat buffer and its gradient a view into the flat gradient
    for i, p in enumerate(tr.params):
        o, n = tr.offsets[i]
        assert p.data_ptr() == tr.flat_param.data_ptr() + 4 * o
        assert p.grad.data_ptr() == tr.flat_grad.data_ptr() + 4 * o
    if rank == 0:
        ret.put({k: v.detach().numpy().copy() for k, v in model.state_dict().items()})
    dist.barrier()
    dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(120)
def test_two_rank_training_matches_single_process():
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    sd2 = ret.get(timeout=100)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single process, whole batch, plain torch Adam: mean over 8 tasks == mean of the two shard means
    torch.manual_seed(0)
    model = Toy()
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    full = make_batch(8, 1)
    for _ in range(4):
        opt.zero_grad()
        crit(model(**full), full["Y_trgt"]).backward()
        opt.step()
    for k, v in model.state_dict().items():
        torch.testing.assert_close(torch.from_numpy(sd2[k]), v, rtol=1e-4, atol=1e-6)


def test_single_process_trainer_matches_torch_adam():
    torch.manual_seed(0)
    m1, m2 = Toy(), Toy()
    m2.load_state_dict(m1.state_dict())
    tr = FlatAdamTrainer(m1, crit, lr=1e-2, adam_impl=torch_adam)
    opt = torch.optim.Adam(m2.parameters(), lr=1e-2)
    batch = make_batch(4, 3)
    for _ in range(5):
        l1 = tr.step(batch)
        opt.zero_grad()
        l2 = crit(m2(**batch), batch["Y_trgt"])
        l2.backward()
        opt.step()
        torch.testing.assert_close(l1, l2.detach(), rtol=1e-5, atol=1e-6)
    for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-6)


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 8, 1_048_576):
        for w in (1, 2, 3, 8):
            parts = [shard_range(n, r, w) for r in range(w)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1
