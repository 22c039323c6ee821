import torch,time
n=185*1024*1024
h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
h2=torch.empty(n//6,dtype=torch.uint8).pin_memory(); d2=torch.empty(n//6,dtype=torch.uint8,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
for _ in range(3): d.copy_(h,non_blocking=True)
torch.cuda.synchronize()
t=time.perf_counter()
for _ in range(10): d.copy_(h,non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t
print("H2D GB/s",10*n/dt/1e9)
t=time.perf_counter()
for _ in range(10): h.copy_(d,non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t
print("D2H GB/s",10*n/dt/1e9)
t=time.perf_counter()
for _ in range(10):
    with torch.cuda.stream(s1): d.copy_(h,non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t
print("both: H2D GB/s",10*n/dt/1e9, "total", 10*(n+n//6)/dt/1e9)
