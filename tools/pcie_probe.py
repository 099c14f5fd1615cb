import torch, time
n=436*1024*1024//4
h1=torch.empty(n,dtype=torch.float32).pin_memory(); h2=torch.empty(n,dtype=torch.float32).pin_memory()
d1=torch.empty(n,dtype=torch.float32,device='cuda'); d2=torch.empty(n,dtype=torch.float32,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
def t(f,k=5):
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(k): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/k*1e3
def seq():
    d1.copy_(h1,non_blocking=True); h2.copy_(d2,non_blocking=True)
def par():
    with torch.cuda.stream(s1): d1.copy_(h1,non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
def h2d(): d1.copy_(h1,non_blocking=True)
def d2h(): h2.copy_(d2,non_blocking=True)
for name,f in [("h2d",h2d),("d2h",d2h),("seq",seq),("par",par)]:
    f(); ms=t(f); print(name, round(ms,2),"ms")
