#!/usr/bin/env python3
"""Host <-> GPU copy bandwidth of this box with k = 1, 2, 4, ... N GPUs copying at once: what the end-to-end pass of bench.py
can at best reach when every rank moves its frames up and its results down through the same host links.

  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py

Each active rank copies 64 MB chunks (H2D only, D2H only, both directions on two streams), walking through 512 MB of
pinned host memory per direction so that the host side of every copy is DRAM traffic as in the bench (a single 64 MB
buffer copied over and over is served from the CPU's last-level cache and overstates what eight ranks can get), timed
with CUDA events; rank 0 prints per-GPU and aggregate GB/s for every k. No kernels of this repository run here."""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 64 << 20
    n_chunks = 8
    h_up = torch.empty(nbytes * n_chunks, dtype=torch.uint8).pin_memory()
    h_dn = torch.empty(nbytes * n_chunks, dtype=torch.uint8).pin_memory()
    h_up.fill_(1)
    d_up = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_dn = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    reps = 24
    out = {}
    k = 1
    ks = []
    while k <= world:
        ks.append(k)
        k *= 2
    for k in ks:
        for mode in ("h2d", "d2h", "both"):
            active = rank < k
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if active:
                for i in range(reps):
                    c = (i % n_chunks) * nbytes
                    if mode in ("h2d", "both"):
                        with torch.cuda.stream(s_up):
                            d_up.copy_(h_up[c:c + nbytes], non_blocking=True)
                    if mode in ("d2h", "both"):
                        with torch.cuda.stream(s_dn):
                            h_dn[c:c + nbytes].copy_(d_dn, non_blocking=True)
            s_up.synchronize()
            s_dn.synchronize()
            torch.cuda.current_stream().wait_stream(s_up)
            torch.cuda.current_stream().wait_stream(s_dn)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            per_dir = reps * nbytes / (ms * 1e-3) / 1e9 if active else 0.0
            t = torch.tensor([per_dir], device="cuda", dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.SUM)
            agg = float(t.item())
            out["k=%d %s" % (k, mode)] = {"per_gpu_GBps_each_direction": round(agg / k, 2),
                                          "aggregate_GBps": round(agg * (2 if mode == "both" else 1), 1)}
    if rank == 0:
        print(json.dumps({"pcie_probe": out, "n_gpus": world, "chunk_MB": nbytes >> 20, "host_MB_per_direction": (nbytes * n_chunks) >> 20}, indent=1))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
