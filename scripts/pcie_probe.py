"""Host<->device link probe: H2D alone, D2H alone and both at once (two streams), pinned memory.
Explains the ceiling of bench.py's e2e number (PCIe full duplex?)."""
import json
import torch

def main():
    n = 512 << 20
    chunk = 32 << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h, chunked):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        step = chunk if chunked else n
        for off in range(0, n, step):
            if h2d:
                with torch.cuda.stream(s1):
                    d_a[off:off + step].copy_(h_in[off:off + step], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out[off:off + step].copy_(d_b[off:off + step], non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b)

    out = {}
    for name, (x, y) in {"h2d": (1, 0), "d2h": (0, 1), "both": (1, 1)}.items():
        for chunked in (False, True):
            run(x, y, chunked)
            ms = min(run(x, y, chunked) for _ in range(3))
            out[f"{name}_{'32MiB' if chunked else 'whole'}"] = {"ms": round(ms, 3), "GBps_each_direction": round(n / ms / 1e6, 1)}
    print(json.dumps(out, indent=1))

main()
