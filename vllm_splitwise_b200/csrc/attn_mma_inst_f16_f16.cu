// Tensor-core grouped paged-attention instantiations: query/out __half, cache __half.
#define INST_T __half
#define INST_CT __half
#include "attn_mma_inst.inc"
