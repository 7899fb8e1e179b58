"""PCIe copy characteristics of the GPU box (pinned host memory, CUDA events)."""
import torch
for size in (64 << 10, 344 << 10, 1376 << 10, 5898 << 10, 64 << 20):
    h = torch.empty(size, dtype=torch.uint8).pin_memory(); d = torch.empty(size, dtype=torch.uint8, device="cuda")
    for name, dst, src in (("H2D", d, h), ("D2H", h, d)):
        for _ in range(3): dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        # latency of ONE copy (event around a single transfer) and throughput of 10 back-to-back
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dst.copy_(src, non_blocking=True); b.record(); torch.cuda.synchronize(); one = a.elapsed_time(b)
        a.record()
        for _ in range(10): dst.copy_(src, non_blocking=True)
        b.record(); torch.cuda.synchronize(); ten = a.elapsed_time(b) / 10
        print(f"{name} {size/1024:9.0f} KiB: single {one*1e3:7.1f} us, pipelined {ten*1e3:7.1f} us = {size/ten/1e6:6.1f} GB/s")
