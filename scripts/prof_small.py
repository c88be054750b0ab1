"""Per-kernel times of one small call (ANS 16 x 1 MiB, fp64 16 x 8 MiB, fp32 1 GiB) through the library's launch profiler."""
import ctypes as C, sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import _capi, codec
def report(tag, steps):
    buf = C.create_string_buffer(1 << 16); _capi.lib.dgb_profile_report(buf, len(buf))
    print(tag, {l.split()[0]: round(float(l.split()[2]) / steps * 1000, 1) for l in buf.value.decode().splitlines()}, "us per call")
def run(tag, fn, steps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    print(tag, "wall per call", round(a.elapsed_time(b) / steps * 1000, 1), "us")
    _capi.lib.dgb_profile_enable(1)
    for _ in range(steps): fn()
    torch.cuda.synchronize(); _capi.lib.dgb_profile_enable(0)
    report(tag, steps)
torch.cuda.set_device(0)
ins = [torch.randint(0, 40, (1 << 20,), dtype=torch.uint8, device="cuda") for _ in range(16)]
mx = codec.get_max_compressed_size(1 << 20)
comp = torch.zeros((16, mx), dtype=torch.uint8, device="cuda"); rows = [comp[i] for i in range(16)]
sizes = torch.zeros(16, dtype=torch.int32, device="cuda"); temp = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
outs = [torch.empty_like(t) for t in ins]
run("ans enc", lambda: codec.ans_encode_batch_pointer(ins, rows, sizes, 10, False, temp=temp))
run("ans dec", lambda: codec.ans_decode_batch_pointer(rows, outs, None, None, 10, False, temp=temp))
ins = [torch.randn(1 << 20, device="cuda", dtype=torch.float64) for _ in range(16)]
mx = codec.get_max_float_compressed_size(codec.FLOAT64, 1 << 20)
comp = torch.zeros((16, mx), dtype=torch.uint8, device="cuda"); rows = [comp[i] for i in range(16)]
run("f64 enc", lambda: codec.float_compress(codec.FLOAT64, ins, rows, sizes, 10, False, temp=temp))
ins = [torch.randn(1 << 28, device="cuda")]
mx = codec.get_max_float_compressed_size(codec.FLOAT32, 1 << 28)
comp = torch.zeros((1, mx), dtype=torch.uint8, device="cuda"); rows = [comp[0]]
temp = torch.empty(3 << 30, dtype=torch.uint8, device="cuda")
run("f32 1g enc", lambda: codec.float_compress(codec.FLOAT32, ins, rows, sizes, 10, False, temp=temp), 5)
