# Timeline of the fused kernel's last launch in bench.py (per CTA; per element when the trace build exists:
# make -C dietgpu_fork_b200/csrc trace). Run on the GPU box; swaps the trace build in for this process only.
if [ -f dietgpu_fork_b200/libdietgpu_b200_trace.so ]; then
  cp dietgpu_fork_b200/libdietgpu_b200.so /tmp/libdietgpu_b200.keep
  cp dietgpu_fork_b200/libdietgpu_b200_trace.so dietgpu_fork_b200/libdietgpu_b200.so
fi
DGB_FUSED_TRACE=/tmp/trace.txt python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > /dev/null 2>&1
[ -f /tmp/libdietgpu_b200.keep ] && cp /tmp/libdietgpu_b200.keep dietgpu_fork_b200/libdietgpu_b200.so
python scripts/trace_fused.py /tmp/trace.txt
