"""Device-resident throughput of every BASELINE.json config, ours next to the reference's CUDA build (when
oracle/_ref/libdietgpu_ref.so exists) with the identical harness: same buffers, CUDA events on the current stream,
3 warm-ups, N timed calls, compressed outputs compared byte-count for byte-count. One JSON line per config.
Usage: python scripts/bench_configs.py [--steps 10] [--only ans,bf16,f32_1g,f32_64,f16,f64,sparse[,bf16_small,bf16_tiny,ans_big]]"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dietgpu_fork_b200 import _capi, codec  # noqa: E402

REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libdietgpu_ref.so")
REF = C.CDLL(REF_PATH) if os.path.exists(REF_PATH) else None
PEAK = 6545.9
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def pa(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


def u32(v):
    return (C.c_uint32 * len(v))(*v)


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


CPU_SAMPLE = 0  # set by --cpu-sample: batch elements of every config also timed on the host cores with the CPU oracle


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def cpu_baseline(kind, ft, ins, counts, sparse):
    """The CPU oracle (oracle/, a restatement of the same stream formats; the reference has no CPU codec) on a
    bounded sample of this config: the first CPU_SAMPLE elements, each cut to at most 4 Mi symbols / words."""
    import time

    import oracle
    k = min(len(ins), CPU_SAMPLE)
    cap = 4 << 20
    wdt = np.uint8 if kind == "ans" else oracle.WORD_DTYPE[ft]
    host = [ins[i].view(torch.uint8).reshape(-1)[: min(counts[i], cap) * (1 if kind == "ans" else wdt().itemsize)]
            .cpu().numpy().view(wdt).copy() for i in range(k)]
    nbytes = sum(h.nbytes for h in host)
    threads = oracle.num_threads()
    if not sparse:  # untimed warm-up: thread pool start, page faults of the library
        w = host[0][:65536]
        if kind == "ans":
            oracle.ans_encode_batch([w] * threads, [np.zeros(oracle.ans_max_compressed_size(w.size), np.uint8) for _ in range(threads)])
        else:
            oracle.float_compress_batch(ft, [w] * threads,
                                        [np.zeros(oracle.float_max_compressed_size(ft, w.size), np.uint8) for _ in range(threads)])
    t0 = time.perf_counter()
    if kind == "ans":
        comp = [np.zeros(oracle.ans_max_compressed_size(h.size), np.uint8) for h in host]
        outs = [np.zeros(h.size, np.uint8) for h in host]
        t0 = time.perf_counter()
        oracle.ans_encode_batch(host, comp)
        t1 = time.perf_counter()
        oracle.ans_decode_batch(comp, outs)
        t2 = time.perf_counter()
    elif sparse:  # no batched driver for the sparse path: one element after the other, one thread
        threads = 1
        t0 = time.perf_counter()
        comp = [oracle.sparse_float_compress(ft, h) for h in host]
        t1 = time.perf_counter()
        outs = [oracle.sparse_float_decompress(ft, c)[2] for c in comp]
        t2 = time.perf_counter()
    else:
        comp = [np.zeros(oracle.float_max_compressed_size(ft, h.size), np.uint8) for h in host]
        outs = [np.zeros(h.size, wdt) for h in host]
        t0 = time.perf_counter()
        oracle.float_compress_batch(ft, host, comp)
        t1 = time.perf_counter()
        oracle.float_decompress_batch(comp, outs)
        t2 = time.perf_counter()
    assert all(np.array_equal(a, b) for a, b in zip(host, outs)), "CPU oracle round trip mismatch"
    return {"kind": "port", "cores": threads, "cpu": cpu_model(),
            "sample": f"{k} element(s) of <= 4 Mi symbols each ({nbytes >> 20} MiB)",
            "compress_GBps": round(nbytes / (t1 - t0) / 1e9, 3), "decompress_GBps": round(nbytes / (t2 - t1) / 1e9, 3)}


def run_config(name, kind, ft, ins, counts, steps, sparse=False):
    """kind: 'ans' (bytes) or 'float'. ins: device tensors; counts: bytes (ans) or words (float)."""
    n = len(ins)
    wb = 1 if kind == "ans" else codec.WORD_BYTES[ft]
    U = sum(counts) * wb
    if kind == "ans":
        mx = max(codec.get_max_compressed_size(c) for c in counts)
    elif sparse:
        mx = max(codec.get_max_sparse_float_compressed_size(ft, c) for c in counts)
    else:
        mx = max(codec.get_max_float_compressed_size(ft, c) for c in counts)
    mx = (mx + 15) // 16 * 16
    comp = torch.zeros((n, mx), dtype=torch.uint8, device="cuda")
    rows = [comp[i] for i in range(n)]
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    outs = [torch.empty(c * wb, dtype=torch.uint8, device="cuda") for c in counts]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda")
    dsz = torch.zeros(n, dtype=torch.int32, device="cuda")
    temp = torch.empty(int(2.5 * U) + (64 << 20), dtype=torch.uint8, device="cuda")
    res = {"config": name, "batch": n, "uncompressed_bytes": U}

    def ours_c():
        if kind == "ans":
            codec.ans_encode_batch_pointer(ins, rows, sizes, 10, False, temp=temp, in_sizes=list(counts))
        elif sparse:
            codec.float_compress_sparse(ft, ins, rows, sizes, 10, temp=temp, in_sizes=list(counts))
        else:
            codec.float_compress(ft, ins, rows, sizes, 10, False, temp=temp, in_sizes=list(counts))

    def ours_d():
        if kind == "ans":
            codec.ans_decode_batch_pointer(rows, outs, ok, dsz, 10, False, temp=temp, out_capacity=list(counts))
        elif sparse:
            codec.float_decompress_sparse(ft, rows, outs, ok, dsz, 10, temp=temp, out_capacity=list(counts))
        else:
            codec.float_decompress(ft, rows, outs, ok, dsz, 10, False, temp=temp, out_capacity=list(counts))

    tc = timed(ours_c, steps)
    csz = sizes.cpu().numpy().astype(np.int64)
    td = timed(ours_d, steps)
    assert ok.cpu().numpy().all()
    for i in (0, n - 1):
        assert torch.equal(outs[i], ins[i].view(torch.uint8).reshape(-1)[: counts[i] * wb]), "round trip mismatch"
    Ctot = int(csz.sum())
    # per-kernel device time (CUDA events around every launch, csrc/prof.cu), 3 extra calls of each direction
    buf = C.create_string_buffer(1 << 16)
    _capi.lib.dgb_profile_report(buf, len(buf))  # clear
    _capi.lib.dgb_profile_enable(1)
    for _ in range(3):
        ours_c()
        ours_d()
    torch.cuda.synchronize()
    _capi.lib.dgb_profile_enable(0)
    _capi.lib.dgb_profile_report(buf, len(buf))
    kern = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.split()
        kern[name] = round(float(ms) / 3, 4)
    res["kernels_ms_per_call"] = kern
    res.update(ratio=round(Ctot / U, 4),
               ours={"compress_GBps": round(U / tc / 1e6, 1), "decompress_GBps": round(U / td / 1e6, 1),
                     "compress_ms": round(tc, 4), "decompress_ms": round(td, 4),
                     "compress_hbm_frac": round((U + Ctot) / tc / 1e6 / PEAK, 3),
                     "decompress_hbm_frac": round((U + Ctot) / td / 1e6 / PEAK, 3)})
    if REF is not None:
        rt = torch.empty(max(int(3 * U), 1 << 30), dtype=torch.uint8, device="cuda")
        rsizes = torch.zeros(n, dtype=torch.int32, device="cuda")
        rcomp = torch.zeros_like(comp)
        rrows = [rcomp[i] for i in range(n)]
        tp, tb = C.c_void_p(rt.data_ptr()), C.c_size_t(rt.numel())

        def ref_c():
            if kind == "ans":
                REF.ref_ans_encode_batch_pointer(tp, tb, 10, 0, n, pa(ins), u32(counts), None, pa(rrows),
                                                 C.c_void_p(rsizes.data_ptr()), stream())
            else:
                fn = REF.ref_float_compress_sparse if sparse else REF.ref_float_compress
                fn(tp, tb, ft, 10, 0, n, pa(ins), u32(counts), pa(rrows), C.c_void_p(rsizes.data_ptr()), stream())

        def ref_d():
            if kind == "ans":
                REF.ref_ans_decode_batch_pointer(tp, tb, 10, 0, n, pa(rrows), pa(outs), u32(counts),
                                                 C.c_void_p(ok.data_ptr()), C.c_void_p(dsz.data_ptr()), stream())
            else:
                fn = REF.ref_float_decompress_sparse if sparse else REF.ref_float_decompress
                fn(tp, tb, ft, 10, 0, n, pa(rrows), pa(outs), u32(counts), C.c_void_p(ok.data_ptr()),
                   C.c_void_p(dsz.data_ptr()), stream())

        rtc = timed(ref_c, steps)
        rtd = timed(ref_d, steps)
        same = bool((rsizes.cpu().numpy().astype(np.int64) == csz).all())
        res["reference"] = {"compress_GBps": round(U / rtc / 1e6, 1), "decompress_GBps": round(U / rtd / 1e6, 1),
                            "compress_ms": round(rtc, 4), "decompress_ms": round(rtd, 4), "same_sizes": same}
        res["speedup"] = {"compress": round(rtc / tc, 3), "decompress": round(rtd / td, 3),
                          "round_trip": round((rtc + rtd) / (tc + td), 3)}
        del rt
    if CPU_SAMPLE:
        res["cpu_baseline"] = cpu_baseline(kind, ft, ins, counts, sparse)
    print(json.dumps(res), flush=True)
    del temp, comp, outs
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--only", default="")
    ap.add_argument("--custom", default="", help="comma list of bf16:<batch>:<KiB per element> (e.g. bf16:32:2048)")
    ap.add_argument("--cpu-sample", type=int, default=0,
                    help="also time the CPU oracle on this many elements of every config (bounded: <= 4 Mi symbols each)")
    a = ap.parse_args()
    global CPU_SAMPLE
    CPU_SAMPLE = a.cpu_sample
    only = set(a.only.split(",")) if a.only else None
    torch.cuda.set_device(0)

    def want(k):
        return only is None or k in only

    if want("ans"):  # config 1: 16 x 1 MiB skewed bytes, lambda = 20
        ins = []
        for e in range(16):
            rng = np.random.default_rng(1000 + e)
            b = np.minimum(255, np.floor(256 * np.minimum(rng.exponential(1 / 20.0, 1 << 20), 1.0))).astype(np.uint8)
            ins.append(torch.from_numpy(b).cuda())
        run_config("ANS bytes 16 x 1 MiB lambda=20", "ans", 0, ins, [1 << 20] * 16, a.steps)
    g = torch.Generator(device="cuda")
    if want("bf16"):  # config 2
        g.manual_seed(2000)
        ins = [(torch.randn(1 << 21, device="cuda", generator=g) * 0.02).to(torch.bfloat16) for _ in range(128)]
        run_config("bf16 128 x 4 MiB N(0,0.02)", "float", codec.BFLOAT16, ins, [1 << 21] * 128, a.steps)
    if want("f16"):
        g.manual_seed(2001)
        ins = [(torch.randn(1 << 21, device="cuda", generator=g) * 0.02).to(torch.float16) for _ in range(128)]
        run_config("fp16 128 x 4 MiB N(0,0.02)", "float", codec.FLOAT16, ins, [1 << 21] * 128, a.steps)
    if want("f32_1g"):  # config 3 (i)
        g.manual_seed(3000)
        ins = [torch.randn(1 << 28, device="cuda", generator=g)]
        run_config("fp32 1 x 1 GiB N(0,1)", "float", codec.FLOAT32, ins, [1 << 28], a.steps)
    if want("f32_64"):  # config 3 (ii)
        g.manual_seed(3000)
        ins = [torch.randn(1 << 22, device="cuda", generator=g) for _ in range(64)]
        run_config("fp32 64 x 16 MiB N(0,1)", "float", codec.FLOAT32, ins, [1 << 22] * 64, a.steps)
    # opt-in extras (only with --only): small batches (launch/latency bound) and a large raw-ANS batch
    if only and "bf16_small" in only:
        g.manual_seed(2002)
        ins = [(torch.randn(1 << 19, device="cuda", generator=g) * 0.02).to(torch.bfloat16) for _ in range(16)]
        run_config("bf16 16 x 1 MiB N(0,0.02)", "float", codec.BFLOAT16, ins, [1 << 19] * 16, a.steps)
    if only and "bf16_tiny" in only:
        g.manual_seed(2003)
        ins = [(torch.randn(1 << 17, device="cuda", generator=g) * 0.02).to(torch.bfloat16) for _ in range(4)]
        run_config("bf16 4 x 256 KiB N(0,0.02)", "float", codec.BFLOAT16, ins, [1 << 17] * 4, a.steps)
    if only and "ans_big" in only:
        g.manual_seed(1001)
        ins = [torch.clamp(torch.floor(256 * torch.clamp(-torch.log(torch.rand(1 << 22, device="cuda", generator=g)) / 20.0,
                                                         max=1.0)), max=255).to(torch.uint8) for _ in range(128)]
        run_config("ANS bytes 128 x 4 MiB lambda=20", "ans", 0, ins, [1 << 22] * 128, a.steps)
    for spec in [c for c in a.custom.split(",") if c]:
        kind, b, kib = spec.split(":")
        b, words = int(b), int(kib) * 512
        g.manual_seed(2004)
        ins = [(torch.randn(words, device="cuda", generator=g) * 0.02).to(torch.bfloat16) for _ in range(b)]
        run_config(f"bf16 {b} x {kib} KiB N(0,0.02)", "float", codec.BFLOAT16, ins, [words] * b, a.steps)
    if a.custom and not a.only:
        return
    if want("f64"):  # config 4
        g.manual_seed(4000)
        ins = [torch.randn(1 << 20, device="cuda", generator=g, dtype=torch.float64) for _ in range(16)]
        run_config("fp64 16 x 8 MiB N(0,1)", "float", codec.FLOAT64, ins, [1 << 20] * 16, a.steps)
    if want("sparse"):  # config 4
        g.manual_seed(4001)
        ins = []
        for _ in range(16):
            t = torch.randn(1 << 22, device="cuda", generator=g)
            t[torch.rand(1 << 22, device="cuda", generator=g) < 0.9] = 0
            ins.append(t)
        run_config("sparse fp32 16 x 16 MiB, 90% zeros", "float", codec.FLOAT32, ins, [1 << 22] * 16, a.steps, sparse=True)


if __name__ == "__main__":
    main()
