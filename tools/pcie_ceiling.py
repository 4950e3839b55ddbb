import torch, time
x=torch.empty(64*720*1280,dtype=torch.uint8).pin_memory(); d=torch.empty_like(x,device='cuda')
y=torch.empty(17750784,dtype=torch.uint8).pin_memory(); dy=torch.empty(17750784,dtype=torch.uint8,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
for both in (0,1):
    torch.cuda.synchronize(); t=time.perf_counter()
    for i in range(20):
        with torch.cuda.stream(s1): d.copy_(x,non_blocking=True)
        if both:
            with torch.cuda.stream(s2): y.copy_(dy,non_blocking=True)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t)/20
    print("both" if both else "h2d only", "%.3f ms per 64-frame batch -> %.1f GB/s h2d, ceiling %.0f frames/s"%(dt*1e3, x.numel()/dt/1e9, 64/dt))
