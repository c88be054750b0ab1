#!/usr/bin/env python
"""bench.py — one JSON line per run for the DietGPU hot path on B200.

Workload (BASELINE.json configs[1]): bf16 float codec, batch 128 x 4 MiB Gaussian gradient-like
tensors (sigma 0.02), probBits 10, checksum off, 1 GPU; with --gpus N every rank gets its own
batch of the same shape (independent batch elements sharded across ranks, no data-path collective:
"scaling": "weak"). A step = floatCompress of the batch followed by floatDecompress of the result.

  value      (U_compress + U_decompress) / t  in GB/s of UNCOMPRESSED bytes, inputs resident in HBM,
             summed over ranks; t = device time (CUDA events), max over ranks.
  e2e        same metric with HOST buffers through the C ABI's host-buffer calls
             (dgb_float_compress_host / dgb_float_decompress_host): tensors and archives live in pinned
             host memory, every H2D / D2H copy is inside the timed region (see run_e2e).
  roofline   the dominant kernel (largest share of device time), its algorithmic bytes per launch
             over its event-timed duration, against MEASURED_PEAKS.json's hbm_gbs. The per-launch events
             run in a second pass of the same K steps right after the timed region (inside it they cost
             about 6 % of the step, which the reference arm does not pay).
  cpu_baseline  the CPU oracle (a port of the same stream format; the reference has NO CPU codec)
             timed on the host cores on a bounded sample, rank 0, N = 1 only.

--impl reference times the reference's own CUDA build (oracle/_ref/libdietgpu_ref.so, compiled
from /root/reference by oracle/build_ref.sh) with the identical harness; if that library is absent
it times the CPU oracle port instead and says so.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 128
ELEM_WORDS = 2 * 1024 * 1024  # 4 MiB of bf16
SIGMA = 0.02
BF16 = 2


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def _cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def pin_to_gpu_numa_node(index):
    """Runs this rank on the CPUs of its GPU's NUMA node, so that the pinned host buffers of the host-buffer leg are
    first touched, hence placed, next to the GPU. The node comes from the PCI device itself
    (/sys/bus/pci/devices/<bdf>/numa_node); NVML's "ideal CPU affinity" is only the fallback (on the round-1 pool it
    reported the same 0-31 mask for all eight GPUs). Both arms do this. Returns what was done, for the JSON line."""
    info = {"numa_node": None, "cpus": None, "source": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        cpus = set()
        try:
            bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
            bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
            if len(bdf.split(":")[0]) == 8:  # NVML prints an 8-digit PCI domain, sysfs a 4-digit one
                bdf = bdf[4:]
            node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
            if node >= 0:
                cpus = _cpulist(open(f"/sys/devices/system/node/node{node}/cpulist").read())
                info.update(numa_node=node, source="sysfs")
        except Exception:
            cpus = set()
        if not cpus:
            words = (os.cpu_count() + 63) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1}
            info["source"] = "nvml affinity"
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["cpus"] = len(cpus)
    except Exception:
        pass
    return info


def make_batch(rank):
    torch.manual_seed(2000 + rank)
    ts = []
    for _ in range(BATCH):
        ts.append((torch.randn(ELEM_WORDS, device="cuda") * SIGMA).to(torch.bfloat16))
    return ts


# ------------------------------------------------------------------------------ reference CUDA build
class RefLib:
    def __init__(self):
        path = os.path.join(ROOT, "oracle", "_ref", "libdietgpu_ref.so")
        self.lib = C.CDLL(path) if os.path.exists(path) else None
        self.temp = None

    def ok(self):
        return self.lib is not None

    def _temp(self):
        if self.temp is None:
            self.temp = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
        return self.temp

    @staticmethod
    def _pa(ts):
        return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])

    def compress(self, ins, rows, sizes):
        t = self._temp()
        n = len(ins)
        self.lib.ref_float_compress(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), BF16, 10, 0, n,
                                    self._pa(ins), (C.c_uint32 * n)(*[x.numel() for x in ins]), self._pa(rows),
                                    C.c_void_p(sizes.data_ptr()),
                                    C.c_void_p(torch.cuda.current_stream().cuda_stream))

    def decompress(self, rows, outs, ok, sizes):
        t = self._temp()
        n = len(rows)
        self.lib.ref_float_decompress(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), BF16, 10, 0, n,
                                      self._pa(rows), self._pa(outs),
                                      (C.c_uint32 * n)(*[x.numel() for x in outs]),
                                      C.c_void_p(ok.data_ptr()), C.c_void_p(sizes.data_ptr()),
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream))


# ------------------------------------------------------------------------------ CPU oracle baseline
def cpu_oracle_roundtrip(sample_elems, budget_s=15.0):
    """Times compress + decompress of `sample_elems` bf16 elements of the workload shape with the
    multithreaded CPU oracle. Returns (GB/s of uncompressed bytes, threads, description)."""
    import oracle
    rng = np.random.default_rng(2000)
    ins = []
    for _ in range(sample_elems):
        x = (rng.standard_normal(ELEM_WORDS) * SIGMA).astype(np.float32)
        ins.append((x.view(np.uint32) >> 16).astype(np.uint16))
    mx = oracle.float_max_compressed_size(BF16, ELEM_WORDS)
    comp = [np.zeros(mx, np.uint8) for _ in ins]
    outs = [np.zeros(ELEM_WORDS, np.uint16) for _ in ins]
    oracle.float_compress_batch(BF16, ins[:1], comp[:1])  # warm (page in, build)
    reps, t_total = 0, 0.0
    while True:
        t0 = time.perf_counter()
        oracle.float_compress_batch(BF16, ins, comp)
        oracle.float_decompress_batch(comp, outs)
        t_total += time.perf_counter() - t0
        reps += 1
        if t_total > budget_s or reps >= 3:
            break
    assert all(np.array_equal(a, b) for a, b in zip(ins, outs))
    nbytes = 2.0 * sample_elems * ELEM_WORDS * 2 * reps
    threads = oracle.num_threads()
    return nbytes / t_total / 1e9, threads, (f"{sample_elems} x 4 MiB bf16 elements of the workload, "
                                             f"compress+decompress, {reps} pass(es), {threads} OpenMP threads")


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip() + f" ({os.cpu_count()} logical CPUs)"
    except Exception:
        pass
    return "unknown"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (kernel sweeps only)")
    ap.add_argument("--no-collective", action="store_true", help="N > 1: skip the compressed-collective record")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local_rank = env_int("LOCAL_RANK", 0)

    ref = RefLib() if args.impl == "reference" else None
    if args.impl == "reference" and not ref.ok() and world > 1 and rank != 0:
        return 0  # CPU stand-in (reference build absent): rank 0 alone runs it

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local_rank)
    numa = pin_to_gpu_numa_node(local_rank)
    dist = None
    if world > 1 and (args.impl == "ours" or ref.ok()):
        # the reference's CUDA build shards the batch exactly like ours: one rank per GPU for both arms
        import torch.distributed as dist
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank),
                                timeout=datetime.timedelta(seconds=300))

    if args.impl == "reference" and not ref.ok():
        return reference_cpu_arm(args)

    # The reference arm must not load the repo's library: everything it needs comes from oracle/_ref.
    if ref is None:
        from dietgpu_fork_b200 import _capi, codec
        max_bytes = codec.get_max_float_compressed_size(BF16, ELEM_WORDS)
    else:
        _capi = codec = None
        ref.lib.ref_float_max_compressed_size.restype = C.c_uint32
        max_bytes = ref.lib.ref_float_max_compressed_size(BF16, C.c_uint32(ELEM_WORDS))

    ins = make_batch(rank)
    U = BATCH * ELEM_WORDS * 2
    comp = torch.empty((BATCH, max_bytes), dtype=torch.uint8, device="cuda")
    rows = [comp[i] for i in range(BATCH)]
    sizes = torch.zeros(BATCH, dtype=torch.int32, device="cuda")
    outs = [torch.empty(ELEM_WORDS, dtype=torch.bfloat16, device="cuda") for _ in range(BATCH)]
    ok = torch.zeros(BATCH, dtype=torch.uint8, device="cuda")
    dsizes = torch.zeros(BATCH, dtype=torch.int32, device="cuda")
    temp = None
    if ref is None:
        temp_bytes = max(_capi.lib.dgb_temp_bytes_required(2, BF16, BATCH, ELEM_WORDS),
                         _capi.lib.dgb_temp_bytes_required(3, BF16, BATCH, ELEM_WORDS))
        temp = torch.empty(temp_bytes, dtype=torch.uint8, device="cuda")

    if ref is not None:
        def do_compress():
            ref.compress(ins, rows, sizes)

        def do_decompress():
            ref.decompress(rows, outs, ok, dsizes)
    else:
        def do_compress():
            codec.float_compress(BF16, ins, rows, sizes, 10, False, temp=temp)

        def do_decompress():
            codec.float_decompress(BF16, rows, outs, ok, dsizes, 10, False, temp=temp)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        do_compress()
        do_decompress()
    torch.cuda.synchronize()
    # correctness inside the bench: the round trip must be bit exact
    csz = sizes.cpu().numpy().astype(np.int64)
    assert ok.cpu().numpy().all(), "decompress reported failure"
    for i in range(BATCH):  # every element, bit for bit
        assert torch.equal(outs[i].view(torch.int16), ins[i].view(torch.int16)), f"round trip mismatch in element {i}"
    assert dsizes.cpu().tolist() == [ELEM_WORDS] * BATCH, "decoded sizes"
    C_total = int(csz.sum())
    ratio = C_total / U

    sampler = ClockSampler(local_rank)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    launches0 = 0
    if ref is None:
        launches0 = _capi.lib.dgb_profile_launches()
    barrier()
    sampler.start()
    t_c = t_d = 0.0
    ev[0].record()
    pairs = []
    for _ in range(args.steps):
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a.record()
        do_compress()
        b.record()
        do_decompress()
        c.record()
        pairs.append((a, b, c))
    ev[1].record()
    barrier()
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[1])
    for a, b, c in pairs:
        t_c += a.elapsed_time(b)
        t_d += b.elapsed_time(c)
    launches = 0
    kern = {}
    if ref is None:
        launches = _capi.lib.dgb_profile_launches() - launches0  # launches inside the timed region
        # Per-kernel device times come from a SECOND pass of the same K steps with a CUDA event pair around every
        # launch (csrc/prof.cu). Inside the timed region those events cost about 6 % of the step (0.70 -> 0.75 ms),
        # which the reference arm does not pay, so the timed region above runs without them.
        _capi.lib.dgb_profile_enable(1)
        for _ in range(args.steps):
            do_compress()
            do_decompress()
        torch.cuda.synchronize()
        _capi.lib.dgb_profile_enable(0)
        buf = C.create_string_buffer(1 << 16)
        _capi.lib.dgb_profile_report(buf, len(buf))
        for line in buf.value.decode().splitlines():
            name, cnt, ms = line.split()
            kern[name] = (int(cnt), float(ms))

    if dist is not None:
        t = torch.tensor([total_ms, t_c, t_d], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, t_c, t_d = t.tolist()
    n_ranks = world if dist is not None else 1
    ms_per_step = total_ms / args.steps
    value = n_ranks * 2.0 * U / (ms_per_step * 1e-3) / 1e9
    comp_gbps = n_ranks * U / (t_c / args.steps * 1e-3) / 1e9
    dec_gbps = n_ranks * U / (t_d / args.steps * 1e-3) / 1e9

    # ---------------- roofline of the dominant kernel (this rank) ----------------
    peak, peak_kind = measured_peak()
    N = BATCH * ELEM_WORDS
    C_ans = C_total - BATCH * 32 - N  # bf16: raw section is N bytes
    alg = {  # algorithmic bytes per launch on this workload (DESIGN.md "Kernels")
        "k_floatSplit": U + N + N,      # read words; write raw section + exponent plane
        "k_encode": N + C_ans,          # read exponent plane; write ANS archives
        "k_compressFused": U + N + C_ans,  # read words; write raw section + ANS payload (plane traffic is extra)
        "k_coalesce": 2 * C_ans,        # read block slots; write ANS payload (extra pass, not algorithmic)
        "k_decode": C_total + U,        # read archives; write words
    }
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")  # dram bytes per launch from `ncu --set full`
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get("dram_bytes_per_launch", {})
    roofline = None
    if kern:
        top = max(kern, key=lambda k: kern[k][1])
        cnt, ms = kern[top]
        if top in alg and ms > 0:
            ach = alg[top] / (ms / cnt * 1e-3) / 1e9
            roofline = {"bound": "hbm", "kernel": top, "achieved": round(ach, 1), "peak": peak,
                        "peak_kind": peak_kind, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": traffic.get(top),
                        "algorithmic_bytes_per_launch": alg[top], "avg_launch_ms": round(ms / cnt, 4),
                        "timed": f"CUDA event pair around each of its {cnt} launches on the launching stream, in a "
                                 "second pass of the same steps right after the timed region"}
    phase = {
        "compress_GBps_uncompressed": round(comp_gbps, 1),
        "decompress_GBps_uncompressed": round(dec_gbps, 1),
        "compress_roofline_frac": round((U + C_total) / (t_c / args.steps * 1e-3) / 1e9 / peak, 4),
        "decompress_roofline_frac": round((U + C_total) / (t_d / args.steps * 1e-3) / 1e9 / peak, 4),
        "kernels_ms_per_step": {k: round(v[1] / args.steps, 4) for k, v in sorted(kern.items())},
    }

    # ---------------- e2e: host buffers through the public op API ----------------
    e2e = None if args.no_e2e else run_e2e(args, ins, ref, comp, rows, sizes, ok, dsizes, temp, dist, n_ranks)
    def build_line(collective):
        line = {
            "metric": "compress+decompress GB/s (uncompressed), bf16 float codec",
            "value": round(value, 2), "unit": "GB/s", "n_gpus": n_ranks, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8/u32 integer (bf16 words)", "data": "synthetic",
            "config": {"workload": "bf16 float codec, batch 128 x 4 MiB N(0, 0.02) tensors per GPU, probBits 10, "
                                   "checksum off (BASELINE.json configs[1])",
                       "batch": BATCH, "elem_bytes": ELEM_WORDS * 2, "l2": "inputs (512 MiB) larger than L2",
                       "parallelism": f"batch elements sharded over {n_ranks} GPU(s), no collective"},
            "compression_ratio": round(ratio, 4),
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "host_placement": numa,
            "roofline": roofline, "phases": phase,
        }
        if collective is not None:
            line["collective"] = collective
        return line

    collective = None
    if ref is None and dist is not None and not args.no_collective:
        # The sharded line must survive whatever the collective leg does: a watchdog thread (it runs even while the
        # main thread is blocked inside a CUDA / NCCL call) prints the line without the record and ends the process.
        def give_up():
            if rank == 0:
                sys.stdout.write(json.dumps(build_line({"error": "collective leg timed out after 180 s"})) + "\n")
                sys.stdout.flush()
            os._exit(0)

        dog = threading.Timer(180.0, give_up)
        dog.daemon = True
        dog.start()
        collective = run_collectives(dist, rank, world)
        dog.cancel()

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    line = build_line(collective)
    if args.impl == "reference":
        line["impl"] = "reference"
        line["reference_kind"] = "reference CUDA build (oracle/_ref/libdietgpu_ref.so) on the GPU"
        line["gpu_launches"] = None
        line["cpu_baseline"] = {"value": line["value"], "unit": "GB/s", "cores": 0, "kind": "reference",
                                "sample": "the reference has no CPU codec: this is its own CUDA build on the "
                                          "same GPU, full workload"}
    elif n_ranks == 1 and not args.no_cpu_baseline:
        v, threads, desc = cpu_oracle_roundtrip(min(BATCH, 16))
        line["cpu_baseline"] = {"value": round(v, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                                "sample": desc, "cpu": cpu_model()}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


def pcie_duplex_gbs(nbytes=256 << 20, chunk=32 << 20):
    """Host<->device link rate per direction with both directions busy (pinned memory, 32 MiB copies on two
    streams; scripts/pcie_probe.py is the long form). Measured in the same run, on this rank, so that the e2e
    number can be read as a fraction of what the link allows."""
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    best = None
    for _ in range(4):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s1.wait_stream(torch.cuda.current_stream())
        s2.wait_stream(torch.cuda.current_stream())
        for off in range(0, nbytes, chunk):
            with torch.cuda.stream(s1):
                d_a[off:off + chunk].copy_(h_in[off:off + chunk], non_blocking=True)
            with torch.cuda.stream(s2):
                h_out[off:off + chunk].copy_(d_b[off:off + chunk], non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1)
        torch.cuda.current_stream().wait_stream(s2)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        best = ms if best is None else min(best, ms)
    return nbytes / best / 1e6


def run_e2e(args, ins, ref, comp, rows, sizes, ok, dsizes, temp, dist, n_ranks):
    """The same step with HOST buffers: pinned host tensors in, archives in pinned host memory, decompressed
    tensors back in pinned host memory; every host<->device copy is inside the timed region.

    ours       codec.float_compress_host + codec.float_decompress_host: the C ABI's host-buffer entry points,
               which pipeline H2D / kernels / D2H over element groups inside the library (csrc/host_pipe.cu).
    reference  its API takes device pointers only, so the caller does what its users do: copy the batch in,
               floatCompress, read the sizes, copy the archives out (exact sizes); copy them back in,
               floatDecompress, copy the tensors out. (Nothing of this repo's library is loaded on that arm.)
    `unpipelined` is that same plain sequence around OUR device-pointer calls, to separate what the pipelining
    buys from what the kernels buy (PCIe dominates both). Every number is the median of 3 repeats of >= 20 steps;
    `pcie_bound` is the same metric if the step ran at the duplex link rate measured in this run."""
    codec = None
    if ref is None:
        from dietgpu_fork_b200 import codec
    U = BATCH * ELEM_WORDS * 2
    max_bytes = comp.shape[1]
    host_in = torch.empty((BATCH, ELEM_WORDS), dtype=torch.bfloat16).pin_memory()
    host_out = torch.empty((BATCH, ELEM_WORDS), dtype=torch.bfloat16).pin_memory()
    host_comp = torch.empty((BATCH, max_bytes), dtype=torch.uint8).pin_memory()
    for i in range(BATCH):
        host_in[i].copy_(ins[i])
    hin_rows = [host_in[i] for i in range(BATCH)]
    hout_rows = [host_out[i] for i in range(BATCH)]
    hcomp_rows = [host_comp[i] for i in range(BATCH)]
    dev_in = torch.empty((BATCH, ELEM_WORDS), dtype=torch.bfloat16, device="cuda")
    dev_out = torch.empty((BATCH, ELEM_WORDS), dtype=torch.bfloat16, device="cuda")
    in_rows = [dev_in[i] for i in range(BATCH)]
    out_rows = [dev_out[i] for i in range(BATCH)]
    comp2 = torch.empty_like(comp)  # the archives come back in from the host into a second device buffer
    rows2 = [comp2[i] for i in range(BATCH)]
    copied = {"h2d": 0, "d2h": 0}

    def plain_step(compress, decompress):
        dev_in.copy_(host_in, non_blocking=True)
        compress(in_rows, rows, sizes)
        csz = sizes.cpu().tolist()  # blocks: archive sizes exist only on the device
        for i, n in enumerate(csz):
            host_comp[i, :n].copy_(comp[i, :n], non_blocking=True)
        for i, n in enumerate(csz):
            comp2[i, :n].copy_(host_comp[i, :n], non_blocking=True)
        decompress(rows2, out_rows, ok, dsizes)
        host_out.copy_(dev_out, non_blocking=True)
        torch.cuda.synchronize()
        c = sum(csz)
        copied["h2d"], copied["d2h"] = U + c, c + U + BATCH * 4

    def pipelined_step():
        csz = codec.float_compress_host(BF16, hin_rows, hcomp_rows, 10, False)
        codec.float_decompress_host(BF16, hcomp_rows, csz, hout_rows, 10, False)
        c = sum(csz)
        copied["h2d"], copied["d2h"] = U + c, c + U + 2 * BATCH * 4 + BATCH

    steps = max(20, args.steps)
    repeats = 3

    def timed(step):
        step()
        vals = []
        for _ in range(repeats):
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            torch.cuda.synchronize()
            host_out.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(steps):
                step()
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            if dist is not None:
                t = torch.tensor([ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = t.item()
            assert torch.equal(host_out.view(torch.int16), host_in.view(torch.int16)), "e2e round trip mismatch"
            vals.append(n_ranks * 2.0 * U / (ms / steps * 1e-3) / 1e9)
        return float(np.median(vals)), vals

    if dist is not None:  # all ranks measure at the same time: they share the host's memory system
        torch.cuda.synchronize()
        dist.barrier()
    link = pcie_duplex_gbs()
    if dist is not None:  # the slowest rank's link bounds the job
        t = torch.tensor([link], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        link = t.item()

    def bound(h2d, d2h):  # GB/s of the metric if both copy directions ran at the duplex link rate, nothing else
        return n_ranks * 2.0 * U / (max(h2d, d2h) / (link * 1e9)) / 1e9

    if ref is not None:
        v, vals = timed(lambda: plain_step(ref.compress, ref.decompress))
        return {"value": round(v, 2), "unit": "GB/s", "h2d_bytes_per_step": copied["h2d"],
                "d2h_bytes_per_step": copied["d2h"], "steps": steps, "repeats": [round(x, 2) for x in vals],
                "pcie_duplex_GBps_per_direction": round(link, 1),
                "pcie_bound": round(bound(copied["h2d"], copied["d2h"]), 2),
                "path": "H2D, floatCompress, sizes D2H, archives D2H, archives H2D, floatDecompress, D2H "
                        "(the reference API takes device pointers only)"}
    plain, _ = timed(lambda: plain_step(
        lambda i, o, s: codec.float_compress(BF16, i, o, s, 10, False, temp=temp),
        lambda i, o, k, s: codec.float_decompress(BF16, i, o, k, s, 10, False, temp=temp)))
    v, vals = timed(pipelined_step)
    pb = bound(copied["h2d"], copied["d2h"])
    return {"value": round(v, 2), "unit": "GB/s", "h2d_bytes_per_step": copied["h2d"],
            "d2h_bytes_per_step": copied["d2h"], "steps": steps, "repeats": [round(x, 2) for x in vals],
            "pcie_duplex_GBps_per_direction": round(link, 1), "pcie_bound": round(pb, 2),
            "frac_of_pcie_bound": round(v / pb, 3),
            "path": "dgb_float_compress_host + dgb_float_decompress_host (copies and kernels pipelined over "
                    "element groups inside the call)",
            "unpipelined": round(plain, 2)}


def run_collectives(dist, rank, world):
    """BASELINE configs[4] next to the sharded number: compressed all-gather / all-to-all of 64 MiB of bf16 per rank
    against the uncompressed NCCL collectives (scripts/bench_collectives.py; bit equality asserted inside)."""
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import bench_collectives
    try:
        res = bench_collectives.measure(dist, rank, world, mib=64, steps=10, warmup=3, chunks=1)
    except Exception as e:  # the sharded line must not be lost to a collective failure; say what happened
        res = {"error": f"{type(e).__name__}: {e}"}
    try:  # BASELINE configs[2](ii): one fp32 64 x 16 MiB batch partitioned over the ranks
        res["sharded_fp32"] = bench_collectives.measure_sharded(dist, rank, world)
    except Exception as e:
        res["sharded_fp32"] = {"error": f"{type(e).__name__}: {e}"}
    return res


def reference_cpu_arm(args):
    """oracle/_ref is absent: time the CPU oracle port (the reference has no CPU codec of its own)."""
    v, threads, desc = cpu_oracle_roundtrip(min(BATCH, 16), budget_s=20.0)
    line = {"impl": "reference", "metric": "compress+decompress GB/s (uncompressed), bf16 float codec",
            "value": round(v, 3), "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "data": "synthetic",
            "reference_kind": "CPU oracle port (oracle/_ref/libdietgpu_ref.so not built)",
            "config": {"workload": "bf16 float codec, 4 MiB N(0, 0.02) tensors, probBits 10 (bounded sample)"},
            "cpu_baseline": {"value": round(v, 3), "unit": "GB/s", "cores": threads, "kind": "port",
                             "sample": desc},
            "e2e": {"value": round(v, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


if __name__ == "__main__":
    sys.exit(main())
