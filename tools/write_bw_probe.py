import torch
x = torch.empty(1048576, 128, device="cuda")
y = torch.empty_like(x)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
us = t(lambda: x.fill_(1.0)); print(f"fill 537MB: {us:.1f} us = {537e6/us/1e6:.2f} TB/s")
us = t(lambda: y.copy_(x)); print(f"copy 537MB: {us:.1f} us = {2*537e6/us/1e6:.2f} TB/s")
us = t(lambda: torch.cuda._sleep(0) or x.zero_()); print(f"zero (memset) 537MB: {us:.1f} us = {537e6/us/1e6:.2f} TB/s")
