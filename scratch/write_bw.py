import torch, time
x = torch.empty(2 * 1024**3, dtype=torch.float64, device='cuda')   # 16 GB
for name, fn in [('zero_', lambda: x.zero_()), ('fill_', lambda: x.fill_(1.5)), ('copy half', lambda: x[:x.numel()//2].copy_(x[x.numel()//2:]))]:
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    b = x.numel() * 8 if 'copy' not in name else x.numel() * 8   # copy: read half + write half
    print(name, '%.2f ms  %.2f TB/s' % (ms, b / ms / 1e9))
