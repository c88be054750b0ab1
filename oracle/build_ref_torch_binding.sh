#!/bin/bash
# Drop-in proof for INTEGRATION.md section 2: compiles the reference's OWN torch binding translation unit
# (/root/reference/dietgpu/DietGpu.cpp, unmodified, where it lies) against include/compat + include/dietgpu_b200.hpp
# and links it with libdietgpu_b200.so instead of the reference's libraries. Output: oracle/_ref/libdietgpu_torch_b200.so
# (git-ignored, test infrastructure; tests/test_ref_binding.py loads it with torch.ops.load_library on the GPU).
set -e
HERE=$(cd "$(dirname "$0")" && pwd); ROOT=$(dirname "$HERE")
SRC=/root/reference/dietgpu/DietGpu.cpp
[ -f "$SRC" ] || { echo "no /root/reference: skipping"; exit 0; }
OUT=$HERE/_ref; mkdir -p "$OUT"
PY=${PYTHON:-python}
read -r TINC TLIB PYINC <<<"$($PY - <<'PYEOF'
import sysconfig, os, torch
t = os.path.dirname(torch.__file__)
print(t + "/include", t + "/lib", sysconfig.get_paths()["include"])
PYEOF
)"
g++ -std=c++17 -O1 -fPIC -shared -D_GLIBCXX_USE_CXX11_ABI=1 -DTORCH_API_INCLUDE_EXTENSION_H \
  -I"$ROOT/include/compat" -I"$ROOT/include" -I"$TINC" -I"$TINC/torch/csrc/api/include" -I"$PYINC" -I/usr/local/cuda/include \
  "$SRC" -o "$OUT/libdietgpu_torch_b200.so" \
  -L"$ROOT/dietgpu_fork_b200" -ldietgpu_b200 -L"$TLIB" -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -L/usr/local/cuda/lib64 -lcudart \
  -Wl,-rpath,'$ORIGIN/../../dietgpu_fork_b200' -Wl,-rpath,"$TLIB"
echo "built $OUT/libdietgpu_torch_b200.so"
