// Prefix-prefill attention instantiations: __nv_bfloat16.
#define INST_MT __nv_bfloat16
#include "ctx_attn_inst.inc"
