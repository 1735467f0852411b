"""``sync_bn`` (BatchNorm row; reference npf/architectures/cnn.py:189,193): two data-parallel
ranks, each with half of the tasks and synchronised BatchNorm statistics, must reproduce ONE process on the whole batch:
predictive mean/std, loss, the summed parameter gradients and the running statistics.  The two ranks share cuda:0 and
talk through gloo (NCCL refuses two ranks on one device); the collective calls are the ones NCCL serves on a real
multi-GPU job.  Without sync_bn the same split gives different numbers (per-replica statistics), which is checked too."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(model, crit, batch):
    model.zero_grad(set_to_none=True)
    out = model(**batch)
    loss = crit(out, batch["Y_trgt"])
    loss.backward()
    return out[0].base_dist.loc.detach(), out[0].base_dist.scale.detach(), loss.detach()


def _worker(rank, world, port, sync, ret):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import npf_gwwaveform_b200 as npf
    import sys

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from util_models import build_product, chirp_batch

    torch.manual_seed(0)
    model = build_product("ConvCNP_unet", 32, density=16, kernel_size=5).cuda().train()
    crit = npf.CNPFLoss().train()
    full = {k: v.cuda() for k, v in chirp_batch(4, 10, 24, 5).items()}
    lo, hi = rank * 2, rank * 2 + 2
    shard = {k: v[lo:hi].contiguous() for k, v in full.items()}
    npf.set_sync_bn(sync)
    loc, scale, loss = _run(model, crit, shard)
    grads = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    dist.all_reduce(grads)                      # data-parallel mean of the shard gradients
    grads /= world
    lsum = loss.clone()
    dist.all_reduce(lsum)
    stats = torch.cat([b.float().reshape(-1) for n, b in model.named_buffers() if "running" in n])
    ret[rank] = dict(loc=loc.cpu(), scale=scale.cpu(), loss=float(lsum) / world, grads=grads.cpu(), stats=stats.cpu())
    npf.set_sync_bn(False)
    dist.destroy_process_group()


@pytest.mark.parametrize("sync", [True, False])
def test_sync_bn_two_ranks_match_one_process(sync):
    import torch.multiprocessing as mp

    import npf_gwwaveform_b200 as npf
    from util_models import build_product, chirp_batch, rel_err

    torch.manual_seed(0)
    model = build_product("ConvCNP_unet", 32, density=16, kernel_size=5).cuda().train()
    crit = npf.CNPFLoss().train()
    full = {k: v.cuda() for k, v in chirp_batch(4, 10, 24, 5).items()}
    loc, scale, loss = _run(model, crit, full)
    grads = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).cpu()
    stats = torch.cat([b.float().reshape(-1) for n, b in model.named_buffers() if "running" in n]).cpu()

    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, sync, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    got_loc = torch.cat([ret[0]["loc"], ret[1]["loc"]], dim=1)
    got_scale = torch.cat([ret[0]["scale"], ret[1]["scale"]], dim=1)
    errs = dict(loc=rel_err(got_loc, loc.cpu()), scale=rel_err(got_scale, scale.cpu()),
                loss=abs(ret[0]["loss"] - float(loss)) / max(1.0, abs(float(loss))),
                grads=float((ret[0]["grads"] - grads).norm() / grads.norm()),
                stats=rel_err(ret[0]["stats"], stats))
    if sync:
        assert max(errs.values()) < 2e-4, errs
        assert torch.equal(ret[0]["stats"], ret[1]["stats"])
    else:
        assert errs["loc"] > 1e-3 and errs["stats"] > 1e-3, errs      # per-replica statistics: a different model state
