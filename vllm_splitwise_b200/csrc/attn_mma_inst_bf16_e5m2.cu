// Tensor-core grouped paged-attention instantiations: query/out __nv_bfloat16, cache uint8_t.
#define INST_T __nv_bfloat16
#define INST_CT uint8_t
#include "attn_mma_inst.inc"
