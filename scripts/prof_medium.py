"""Per-kernel times (library launch profiler, CUDA events around every launch) of medium bf16 compress calls under
both schedules. Usage: python scripts/prof_medium.py"""
import ctypes as C, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import _capi, codec
def run(tag, fn, steps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    wall = a.elapsed_time(b) / steps * 1000
    _capi.lib.dgb_profile_enable(1)
    for _ in range(steps): fn()
    torch.cuda.synchronize(); _capi.lib.dgb_profile_enable(0)
    buf = C.create_string_buffer(1 << 16); _capi.lib.dgb_profile_report(buf, len(buf))
    print(tag, "wall", round(wall, 1), "us;", {l.split()[0]: round(float(l.split()[2]) / steps * 1000, 1) for l in buf.value.decode().splitlines()})
torch.cuda.set_device(0)
temp = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
for batch, words in ((1, 32 << 20), (8, 8 << 20), (1, 128 << 20)):
    ins = [(torch.randn(words, device="cuda") * 0.02).to(torch.bfloat16) for _ in range(batch)]
    mx = codec.get_max_float_compressed_size(codec.BFLOAT16, words)
    comp = torch.zeros((batch, mx), dtype=torch.uint8, device="cuda"); rows = [comp[i] for i in range(batch)]
    sizes = torch.zeros(batch, dtype=torch.int32, device="cuda")
    for sched, tiles in (("fused", 0), ("separate", 1 << 40)):
        _capi.lib.dgb_set_fused_min_tiles(tiles)
        run(f"bf16 {batch} x {words * 2 >> 20} MiB {sched}", lambda: codec.float_compress(codec.BFLOAT16, ins, rows, sizes, 10, False, temp=temp))
    _capi.lib.dgb_set_fused_min_tiles(-1)
    del ins, comp
