import torch, time
x = torch.empty(256*1024*1024, dtype=torch.float32).pin_memory()
d = torch.empty_like(x, device="cuda")
for n in (1, 2, 4):
    streams = [torch.cuda.Stream() for _ in range(n)]
    chunks = x.chunk(n); dch = d.chunk(n)
    torch.cuda.synchronize()
    for rep in range(3):
        t0 = time.perf_counter()
        for s, a, b in zip(streams, chunks, dch):
            with torch.cuda.stream(s):
                b.copy_(a, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(n, "streams:", x.numel()*4/dt/1e9, "GB/s")
