import os, sys, socket, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
def worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import npf_gwwaveform_b200 as npf
    from npf_gwwaveform_b200 import kernels as Kn
    g = torch.Generator().manual_seed(1)
    B, L, C, ks = 4, 24, 32, 5
    x = torch.randn(B, L, C, generator=g).cuda(); w = torch.randn(C, 1, ks, generator=g).cuda(); gamma = (1 + 0.1*torch.randn(C, generator=g)).cuda(); beta = (0.1*torch.randn(C, generator=g)).cuda()
    dy = torch.randn(B, L, C, generator=g).cuda()
    rm, rv = torch.zeros(C).cuda(), torch.ones(C).cuda()
    xs = x[rank*2:rank*2+2].clone().requires_grad_(True); gm = gamma.clone().requires_grad_(True); bt = beta.clone().requires_grad_(True); ww = w.clone().requires_grad_(True)
    npf.set_sync_bn(True)
    y = Kn.norm_act_dwconv(xs, gm, bt, ww, None, None, rm, rv, True, 1e-5, 0.1)
    y.backward(dy[rank*2:rank*2+2])
    gg = torch.cat([gm.grad, bt.grad, ww.grad.reshape(-1)]); dist.all_reduce(gg)
    ret[rank] = dict(y=y.detach().cpu(), dx=xs.grad.cpu(), gg=gg.cpu(), rm=rm.cpu(), rv=rv.cpu())
    dist.destroy_process_group()
if __name__ == "__main__":
    import torch.multiprocessing as mp
    from npf_gwwaveform_b200 import kernels as Kn
    g = torch.Generator().manual_seed(1)
    B, L, C, ks = 4, 24, 32, 5
    x = torch.randn(B, L, C, generator=g).cuda(); w = torch.randn(C, 1, ks, generator=g).cuda(); gamma = (1 + 0.1*torch.randn(C, generator=g)).cuda(); beta = (0.1*torch.randn(C, generator=g)).cuda()
    dy = torch.randn(B, L, C, generator=g).cuda()
    rm, rv = torch.zeros(C).cuda(), torch.ones(C).cuda()
    xs = x.clone().requires_grad_(True); gm = gamma.clone().requires_grad_(True); bt = beta.clone().requires_grad_(True); ww = w.clone().requires_grad_(True)
    y = Kn.norm_act_dwconv(xs, gm, bt, ww, None, None, rm, rv, True, 1e-5, 0.1)
    y.backward(dy)
    gg = torch.cat([gm.grad, bt.grad, ww.grad.reshape(-1)]).cpu()
    ctx = mp.get_context("spawn"); ret = ctx.Manager().dict()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ps = [ctx.Process(target=worker, args=(r, 2, port, ret)) for r in range(2)]
    [p.start() for p in ps]; [p.join(120) for p in ps]
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    print("y", rel(torch.cat([ret[0]["y"], ret[1]["y"]]), y.detach().cpu()))
    print("dx", rel(torch.cat([ret[0]["dx"], ret[1]["dx"]]), xs.grad.cpu()))
    print("gg", rel(ret[0]["gg"], gg), "dgamma", rel(ret[0]["gg"][:C], gg[:C]), "dbeta", rel(ret[0]["gg"][C:2*C], gg[C:2*C]), "dw", rel(ret[0]["gg"][2*C:], gg[2*C:]))
    print("rm", rel(ret[0]["rm"], rm.cpu()), "rv", rel(ret[0]["rv"], rv.cpu()))
