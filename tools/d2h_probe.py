"""What the box can ingest: every rank copies one (2N, T+1) trajectory-sized buffer (4.5 GB) from its GPU to
pinned host memory back to back, all ranks at once -- the ceiling of any API that returns the reference's
arrays to the host (bench.py's `e2e`).  One JSON line from rank 0.

    python tools/d2h_probe.py [--buffers 6]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \\
        --master-port 29544 tools/d2h_probe.py
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from finitedifference_b200.sweep import bind_host_to_gpu_node
    from tools.records import ClockSampler
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    numa = bind_host_to_gpu_node(dev)
    nbuf = int(sys.argv[sys.argv.index("--buffers") + 1]) if "--buffers" in sys.argv else 2   # distinct pinned destinations, round robin
    n2, t1 = 2 * 750 * 750, 501
    src = torch.ones((n2, t1), dtype=torch.float64, device=dev)
    dst = [torch.empty((n2, t1), dtype=torch.float64).pin_memory() for _ in range(nbuf)]
    reps = max(6, 2 * nbuf)
    for d in dst:
        d.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    with ClockSampler(local, 100) as clk:
        t0 = time.perf_counter()
        for i in range(reps):
            dst[i % nbuf].copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        el = time.perf_counter() - t0
    t = torch.tensor([el], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gb = reps * src.numel() * 8 / 1e9
    if rank == 0:
        print(json.dumps({"probe": "pinned D2H, all ranks at once", "n_gpus": world, "bytes_per_copy": src.numel() * 8,
                          "copies_per_rank": reps, "pinned_buffers": nbuf, "seconds_max_over_ranks": float(t.item()),
                          "GBps_per_gpu": gb / float(t.item()), "GBps_aggregate": world * gb / float(t.item()),
                          "host_numa_node_rank0": numa, "cpu_count": os.cpu_count(), "clocks": clk.summary()}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
