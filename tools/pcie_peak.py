import torch, time
for mb in (16, 64, 256):
    h = torch.empty(mb << 20, dtype=torch.uint8).pin_memory(); d = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
    for direction in ("h2d", "d2h"):
        torch.cuda.synchronize(); best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4):
                (d.copy_(h, non_blocking=True) if direction == "h2d" else h.copy_(d, non_blocking=True))
            e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) / 4)
        print(mb, "MB", direction, round((mb << 20) / best / 1e6, 1), "GB/s")
