// Tensor-core grouped paged-attention instantiations: query/out __nv_bfloat16, cache __nv_bfloat16.
#define INST_T __nv_bfloat16
#define INST_CT __nv_bfloat16
#include "attn_mma_inst.inc"
