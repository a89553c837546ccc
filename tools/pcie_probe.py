#!/usr/bin/env python3
"""Host<->device copy bandwidth per rank, alone and with every rank copying at once (torchrun).
Tells a PCIe / host-memory limit of the box from a placement problem of the pinned buffers."""
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mode = os.environ.get("PROBE_AFFINITY", "nvml")
    if mode == "nvml":
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        ideal = {64 * i + b for i, w in enumerate(words) for b in range(64) if w >> b & 1}
        mine = sorted(ideal & os.sched_getaffinity(0))
        if mine:
            os.sched_setaffinity(0, mine)
    aff = sorted(os.sched_getaffinity(0))
    nbytes = 768 << 20
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    host.fill_(1)
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")

    def run(direction, ranks):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        if rank in ranks:
            for _ in range(8):
                if direction == "d2h":
                    host.copy_(dev, non_blocking=True)
                else:
                    dev.copy_(host, non_blocking=True)
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([8 * nbytes / dt / 1e9 if rank in ranks else 0.0], device="cuda")
        out = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [round(float(x), 1) for x in out]

    info = [None] * world
    dist.all_gather_object(info, (rank, len(aff), aff[0], aff[-1]))
    if rank == 0:
        print("cpus", os.cpu_count(), "affinity (rank, n, first, last):", info, flush=True)
        try:
            for n in sorted(os.listdir("/sys/devices/system/node")):
                if n.startswith("node"):
                    print(n, open(f"/sys/devices/system/node/{n}/cpulist").read().strip(), flush=True)
        except OSError as e:
            print("no numa info", e)
    for d in ("d2h", "h2d"):
        a = run(d, {0})
        b = run(d, set(range(world)))
        c = run(d, set(range(0, world, 2)))
        if rank == 0:
            print(d, "rank 0 alone", a[0], "GB/s | all ranks", b, "sum", round(sum(b), 1), "| even ranks", c, "sum", round(sum(c), 1), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
