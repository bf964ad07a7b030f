import torch, time
x = torch.empty(64*640*480*3, dtype=torch.uint8).pin_memory()
d = torch.empty_like(x, device='cuda')
for n in (1,4):
    for _ in range(3): d.copy_(x, non_blocking=True)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        for c in range(n):
            a=c*x.numel()//n; b=(c+1)*x.numel()//n
            d[a:b].copy_(x[a:b], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/10
    print(n,'chunks', ms,'ms', x.numel()/ms/1e6,'GB/s')
y = torch.empty(6*1024*1024, dtype=torch.uint8).pin_memory()
dd = torch.empty_like(y, device='cuda')
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): y.copy_(dd, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print('d2h 6MB', e0.elapsed_time(e1)/10, 'ms')
