import torch, time
x = torch.empty(715*1024*1024, dtype=torch.uint8, device='cuda')
h = torch.empty_like(x, device='cpu').pin_memory()
for name, a, b in (("d2h", h, x), ("h2d", x, h)):
    a.copy_(b, non_blocking=True); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): a.copy_(b, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(name, 5*x.numel()/e0.elapsed_time(e1)/1e6, "GB/s")
