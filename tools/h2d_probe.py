import torch, numpy as np, time
n,H,W=60,4000,6000
host=torch.empty((n,H,W),dtype=torch.uint8,pin_memory=True); a=host.numpy()
dev=torch.empty((n,H*W),dtype=torch.uint8,device='cuda')
print("is_pinned view:", torch.from_numpy(a[3]).is_pinned())
for rep in range(3):
    torch.cuda.synchronize(); t0=time.perf_counter()
    for q in range(n):
        dev[q].view(H,W).copy_(torch.from_numpy(a[q]), non_blocking=True)
    t1=time.perf_counter(); torch.cuda.synchronize(); t2=time.perf_counter()
    print(f"enqueue {1e3*(t1-t0):.1f} ms, total {1e3*(t2-t0):.1f} ms, {n*H*W/(t2-t0)/1e9:.1f} GB/s")
b=np.empty((n,H,W),dtype=np.uint8)
torch.cuda.synchronize(); t0=time.perf_counter()
for q in range(n):
    dev[q].view(H,W).copy_(torch.from_numpy(b[q]), non_blocking=True)
torch.cuda.synchronize(); t2=time.perf_counter()
print(f"pageable total {1e3*(t2-t0):.1f} ms, {n*H*W/(t2-t0)/1e9:.1f} GB/s")
