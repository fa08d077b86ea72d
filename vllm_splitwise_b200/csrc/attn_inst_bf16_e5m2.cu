// SIMT paged-attention instantiations: query/out bf16, cache e5m2.
#define INST_T __nv_bfloat16
#define INST_CT uint8_t
#include "attn_inst.inc"
