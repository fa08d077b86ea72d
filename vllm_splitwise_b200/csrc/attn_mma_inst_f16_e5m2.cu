// Tensor-core grouped paged-attention instantiations: query/out __half, cache uint8_t.
#define INST_T __half
#define INST_CT uint8_t
#include "attn_mma_inst.inc"
