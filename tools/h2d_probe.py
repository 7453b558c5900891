import time, torch
dev = torch.device('cuda')
for nbytes in (64 << 10, 512 << 10, 2 << 20, 16 << 20):
    h = torch.empty(nbytes // 4, dtype=torch.int32).pin_memory()
    d = torch.empty(nbytes // 4, dtype=torch.int32, device=dev)
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    e0.record()
    for _ in range(20):
        h.copy_(d, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms2 = e0.elapsed_time(e1) / 20
    print("%8d KB  H2D %.3f ms %.1f GB/s   D2H %.3f ms %.1f GB/s" % (nbytes >> 10, ms, nbytes / ms / 1e6, ms2, nbytes / ms2 / 1e6))
