#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the UNMODIFIED reference CUDA kernels of the hot path for
# sm_100a into oracle/_ref/vllm_ref_C.so (git-ignored; travels to the GPU box with gpurun).
# Sources are compiled where they lie under /root/reference -- nothing is copied.  The only file
# of ours in the link is oracle/ref_build/ref_binding.cpp (a trimmed pybind front).
# Takes ~8 minutes on 8 cores (216 attention template instantiations).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
[ -d "$REF/csrc" ] || { echo "reference not present at $REF; keeping prebuilt $OUT"; exit 0; }
mkdir -p "$OUT"
PY=python
TORCH_INC=$($PY - <<'PY'
from torch.utils.cpp_extension import include_paths
print(" ".join("-I"+p for p in include_paths(device_type="cuda")))
PY
)
TORCH_LIB=$($PY -c "import torch,os;print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
PY_INC=$($PY -c "import sysconfig;print('-I'+sysconfig.get_paths()['include'])")
ABI=$($PY -c "import torch;print(int(torch._C._GLIBCXX_USE_CXX11_ABI))")
COMMON="-O3 -std=c++17 -DENABLE_FP8_E5M2 -DTORCH_EXTENSION_NAME=vllm_ref_C -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=$ABI $TORCH_INC $PY_INC"
NVCC="nvcc -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr -Xcompiler -fPIC -U__CUDA_NO_HALF_OPERATORS__ -U__CUDA_NO_HALF_CONVERSIONS__ -U__CUDA_NO_BFLOAT16_CONVERSIONS__ -U__CUDA_NO_HALF2_OPERATORS__ --threads 4"
build_cu() {  # $1 = source under $REF/csrc, $2 = object name
  if [ ! -f "$OUT/$2" ] || [ "$REF/csrc/$1" -nt "$OUT/$2" ]; then
    echo "[build_ref] nvcc $1"; $NVCC $COMMON -c "$REF/csrc/$1" -o "$OUT/$2"
  fi
}
build_cu attention/attention_kernels.cu attention_kernels.o &
P1=$!
build_cu cache_kernels.cu cache_kernels.o &
P2=$!
build_cu cuda_utils_kernels.cu cuda_utils_kernels.o
build_cu pos_encoding_kernels.cu pos_encoding_kernels.o
build_cu custom_all_reduce.cu custom_all_reduce.o
g++ -fPIC $COMMON -c "$HERE/ref_build/ref_binding.cpp" -o "$OUT/ref_binding.o"
wait $P1; wait $P2
g++ -shared -o "$OUT/vllm_ref_C.so" "$OUT/attention_kernels.o" "$OUT/cache_kernels.o" \
    "$OUT/cuda_utils_kernels.o" "$OUT/pos_encoding_kernels.o" "$OUT/custom_all_reduce.o" "$OUT/ref_binding.o" \
    -L"$TORCH_LIB" -Wl,-rpath,"$TORCH_LIB" -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -ltorch_python \
    -L/usr/local/cuda/lib64 -lcudart -L/usr/local/cuda/lib64/stubs -lcuda  # custom_all_reduce.cuh calls the driver API
echo "[build_ref] built $OUT/vllm_ref_C.so"
