import torch, time
n=16384
obs_d=torch.empty(n,87,device='cuda'); obs_h=torch.empty(n,87).pin_memory()
act_h=torch.empty(n,21).pin_memory(); act_d=torch.empty(n,21,device='cuda')
for name,(dst,src) in {'D2H obs 5.7MB':(obs_h,obs_d),'H2D act 1.4MB':(act_d,act_h)}.items():
    for _ in range(3): dst.copy_(src,non_blocking=True)
    torch.cuda.synchronize(); a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): dst.copy_(src,non_blocking=True)
    b.record(); torch.cuda.synchronize(); ms=a.elapsed_time(b)/20
    print(name, f"{ms*1e3:.1f} us", f"{src.numel()*4/ms/1e6:.1f} GB/s")
big_d=torch.empty(256<<20,dtype=torch.uint8,device='cuda'); big_h=torch.empty(256<<20,dtype=torch.uint8).pin_memory()
a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
a.record(); big_h.copy_(big_d,non_blocking=True); b.record(); torch.cuda.synchronize(); print('D2H 256MB', 256/1024/(a.elapsed_time(b)/1e3), 'GB/s')
