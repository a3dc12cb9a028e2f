import torch, time
torch.backends.cuda.matmul.allow_tf32=False
for n,b in [(8192,1),(128,4096),(64,16384)]:
    a=torch.randn(b,n,n,dtype=torch.float64,device='cuda'); c=torch.randn(b,n,n,dtype=torch.float64,device='cuda')
    for _ in range(3): torch.bmm(a,c)
    torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    reps=5; e0.record()
    for _ in range(reps): torch.bmm(a,c)
    e1.record(); torch.cuda.synchronize(); ms=e0.elapsed_time(e1)/reps
    print(f"bmm f64 n={n} batch={b}: {ms:.3f} ms  {2*n**3*b/ms*1e-9:.2f} TFLOP/s")
