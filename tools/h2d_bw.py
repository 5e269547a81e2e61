"""Host -> device bandwidth per rank with all ranks copying at once: what bounds the end-to-end number
(cnmf.nmf from host arrays) at N = 4 and 8 GPUs of one box.

    python tools/h2d_bw.py                                  (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/h2d_bw.py

Each rank copies a 2 GiB pinned buffer to its GPU ten times (cudaMemcpyAsync, CUDA events), then the same from a
pageable buffer through torch, and rank 0 prints per-rank GB/s, the aggregate, the NUMA node of each GPU and the
CPU affinity of each rank.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def numa_of_gpu(i):
    try:
        bus = torch.cuda.get_device_properties(i).pci_bus_id
    except Exception:
        try:
            import pynvml
            pynvml.nvmlInit()
            bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(i)).busId
            bus = bus.decode() if isinstance(bus, bytes) else bus
        except Exception:
            return None
    bus = str(bus).lower()
    if len(bus.split(":")[0]) == 8:
        bus = bus[4:]
    try:
        return int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
    except Exception:
        return None


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = 2 << 30
    src = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    src.fill_(1)
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    page = torch.empty(nbytes // 4, dtype=torch.uint8)
    page.fill_(2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    dst.copy_(src, non_blocking=True)
    sync()
    e0.record()
    for _ in range(10):
        dst.copy_(src, non_blocking=True)
    e1.record()
    sync()
    pinned = 10 * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
    dst[:page.numel()].copy_(page)
    sync()
    e0.record()
    for _ in range(4):
        dst[:page.numel()].copy_(page)
    e1.record()
    sync()
    pageable = 4 * page.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9
    mine = {"rank": rank, "gpu": local, "gpu_numa": numa_of_gpu(local), "cpus": len(os.sched_getaffinity(0)),
            "pinned_gbs": pinned, "pageable_torch_gbs": pageable}
    if world > 1:
        out = [None] * world
        dist.all_gather_object(out, mine)
    else:
        out = [mine]
    if rank == 0:
        print(json.dumps({"world": world, "aggregate_pinned_gbs": sum(o["pinned_gbs"] for o in out),
                          "aggregate_pageable_gbs": sum(o["pageable_torch_gbs"] for o in out), "host_cpus": os.cpu_count(),
                          "ranks": out}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
