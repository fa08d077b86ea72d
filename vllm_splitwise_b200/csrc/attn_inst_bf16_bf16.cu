// SIMT paged-attention instantiations: query/out bf16, cache bf16.
#define INST_T __nv_bfloat16
#define INST_CT __nv_bfloat16
#include "attn_inst.inc"
