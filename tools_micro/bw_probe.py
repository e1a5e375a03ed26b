import torch, time
dev = torch.device("cuda:0")
n = 1 << 28  # 1 GiB fp32
a = torch.empty(n, device=dev); b = torch.empty(n, device=dev)
def t(f, nbytes, name):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(10):
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("%-12s %.3f ms  %.0f GB/s" % (name, best, nbytes / best / 1e6))
t(lambda: a.fill_(1.0), n * 4, "write fill")
t(lambda: b.copy_(a), n * 8, "copy")
t(lambda: a.sum(), n * 4, "read sum")
# small sizes matching our step (297 MB write / 148 MB read)
m = 297 * 1024 * 1024 // 4
t(lambda: a[:m].fill_(1.0), m * 4, "write 297MB")
m2 = 148 * 1024 * 1024 // 4
t(lambda: a[:m2].sum(), m2 * 4, "read 148MB")
