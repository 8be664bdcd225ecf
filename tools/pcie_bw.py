import torch, time
n=680*1024*1024
h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
for name,(a,b) in {'h2d':(d,h),'d2h':(h,d)}.items():
    for _ in range(2): a.copy_(b,non_blocking=True); torch.cuda.synchronize()
    t=time.perf_counter()
    for _ in range(5): a.copy_(b,non_blocking=True)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t)/5
    print(name, round(n/dt/1e9,1),'GB/s')
# both directions at once
s1,s2=torch.cuda.Stream(),torch.cuda.Stream(); d2=torch.empty(n,dtype=torch.uint8,device='cuda'); h2=torch.empty(n,dtype=torch.uint8).pin_memory()
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(5):
    with torch.cuda.stream(s1): d.copy_(h,non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
torch.cuda.synchronize(); dt=(time.perf_counter()-t)/5
print('bidir each', round(n/dt/1e9,1),'GB/s')
