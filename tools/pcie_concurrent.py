"""Host-to-device copy rate with 1, 2, 4, 8 GPUs copying AT ONCE from pinned host memory (one process, one stream per
GPU): the host-side ceiling of the end-to-end path at N GPUs.  usage: python tools/pcie_concurrent.py"""
import json, time
import torch

n = torch.cuda.device_count()
MB, REPS = 256, 8
host = [torch.empty(MB << 20, dtype=torch.uint8).pin_memory() for _ in range(n)]
dev = [torch.empty(MB << 20, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]
print(json.dumps(dict(gpus_visible=n, cpu_affinity=len(__import__("os").sched_getaffinity(0)))), flush=True)
k = 1
while k <= n:
    for direction in ("h2d", "d2h"):
        for i in range(k):
            torch.cuda.synchronize(i)
        t0 = time.perf_counter()
        for _ in range(REPS):
            for i in range(k):
                with torch.cuda.stream(streams[i]):
                    if direction == "h2d":
                        dev[i].copy_(host[i], non_blocking=True)
                    elif direction == "d2h":
                        host[i].copy_(dev[i], non_blocking=True)
                    else:  # both directions queued on the one stream (they serialise per GPU, as in a simple host loop)
                        dev[i].copy_(host[i], non_blocking=True)
        for i in range(k):
            streams[i].synchronize()
        dt = time.perf_counter() - t0
        gbs = k * REPS * (MB << 20) / dt / 1e9
        print(json.dumps(dict(gpus=k, direction=direction, aggregate_gbs=round(gbs, 1), per_gpu_gbs=round(gbs / k, 1),
                              mpps_at_66B_per_packet=round(gbs * 1e9 / 66 / 1e6))), flush=True)
    k *= 2
