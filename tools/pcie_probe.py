import torch,time
for mb in (0.2, 4, 6.16, 64):
    n=int(mb*1e6)
    h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
    for dirn in ('h2d','d2h'):
        for _ in range(3): (d.copy_(h,non_blocking=True) if dirn=='h2d' else h.copy_(d,non_blocking=True))
        torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): (d.copy_(h,non_blocking=True) if dirn=='h2d' else h.copy_(d,non_blocking=True))
        e1.record(); torch.cuda.synchronize()
        us=e0.elapsed_time(e1)*1e3/20
        print(f"{dirn} {mb} MB: {us:.1f} us  {n/us/1e3:.1f} GB/s")
