#!/usr/bin/env bash
# Builds the compiled PyTorch front (torch_binding.cpp) into ../lib/_b200_torch.so, linked against
# ../lib/libb200kv.so (rpath $ORIGIN, so both load from the same directory).  Host compiler only.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../lib"
mkdir -p "$OUT"
PY=${PYTHON:-python}
SRC="$HERE/torch_binding.cpp"
if [ -f "$OUT/_b200_torch.so" ] && [ "$OUT/_b200_torch.so" -nt "$SRC" ] && [ "$OUT/_b200_torch.so" -nt "$HERE/../../include/b200_kv.h" ]; then
  echo "[torch_binding] up to date"; exit 0
fi
TORCH_INC=$($PY - <<'PY'
from torch.utils.cpp_extension import include_paths
print(" ".join("-I"+p for p in include_paths(device_type="cuda")))
PY
)
TORCH_LIB=$($PY -c "import torch,os;print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
PY_INC=$($PY -c "import sysconfig;print('-I'+sysconfig.get_paths()['include'])")
ABI=$($PY -c "import torch;print(int(torch._C._GLIBCXX_USE_CXX11_ABI))")
g++ -O2 -std=c++17 -fPIC -shared -DTORCH_EXTENSION_NAME=_b200_torch -DTORCH_API_INCLUDE_EXTENSION_H \
    -D_GLIBCXX_USE_CXX11_ABI=$ABI $TORCH_INC $PY_INC "$SRC" -o "$OUT/_b200_torch.so" \
    -L"$OUT" -lb200kv -Wl,-rpath,'$ORIGIN' \
    -L"$TORCH_LIB" -Wl,-rpath,"$TORCH_LIB" -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -ltorch_python \
    -L/usr/local/cuda/lib64 -lcudart
echo "[torch_binding] built $OUT/_b200_torch.so"
