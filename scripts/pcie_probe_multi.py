"""Host<->device link ceiling with N ranks copying AT THE SAME TIME (one rank per GPU, torchrun): per-direction GB/s
per rank with H2D alone, D2H alone and both directions busy, pinned memory, 32 MiB copies. Answers whether the
host-buffer (e2e) number of bench.py stops scaling because of the host memory system rather than the codec:
the sum over ranks with both directions busy is the ceiling of sum(e2e bytes per direction) / time.
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P
scripts/pcie_probe_multi.py [--pin 1]   (N = 1 works without torchrun). --pin 1 binds each rank to the CPUs of its
GPU's NUMA node first (bench.py's pin_to_gpu_numa_node), so the two placements can be compared."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pin", type=int, default=1)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    placement = {"numa_node": None, "cpus": None, "source": "unpinned"}
    if a.pin:
        import bench
        placement = bench.pin_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, chunk = 512 << 20, 32 << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_in.zero_()
    h_out.zero_()
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        for _ in range(2):
            for off in range(0, n, chunk):
                if h2d:
                    with torch.cuda.stream(s1):
                        d_a[off:off + chunk].copy_(h_in[off:off + chunk], non_blocking=True)
                if d2h:
                    with torch.cuda.stream(s2):
                        h_out[off:off + chunk].copy_(d_b[off:off + chunk], non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        return 2 * n / e0.elapsed_time(e1) / 1e6  # GB/s per direction

    def gather(v):
        t = torch.tensor([float(v)], device="cuda", dtype=torch.float64)
        if world == 1:
            return [float(v)]
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        return [float(z.item()) for z in allv]

    res = {}
    for name, (x, y) in {"h2d": (1, 0), "d2h": (0, 1), "both": (1, 1)}.items():
        run(x, y)
        vals = gather(max(run(x, y) for _ in range(3)))
        res[name] = {"per_rank_GBps_each_direction": [round(z, 1) for z in vals], "sum": round(sum(vals), 1),
                     "min": round(min(vals), 1)}
    nodes = [int(z) for z in gather(-1 if placement["numa_node"] is None else placement["numa_node"])]
    if rank == 0:
        try:
            numa_nodes = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
        except Exception:
            numa_nodes = None
        print(json.dumps({"n_gpus": world, "pinned_to_gpu_node": bool(a.pin), "placement_source": placement["source"],
                          "host_numa_nodes": numa_nodes, "gpu_numa_node": nodes, "cpus": os.cpu_count(), **res}),
              flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
