import torch
x=torch.empty(1<<30, dtype=torch.float32, device='cuda')  # 4 GB
y=torch.empty(1<<30, dtype=torch.float32, device='cuda')
for name,fn,bytes_ in (("fill",lambda: x.fill_(1.0),4*(1<<30)),("copy",lambda: y.copy_(x),8*(1<<30)),("sum(read)",lambda: x.sum(),4*(1<<30))):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/10
    print(name, f"{ms:.3f} ms", f"{bytes_/ms/1e6:.0f} GB/s")
