// Prefix-prefill attention instantiations: __half.
#define INST_MT __half
#include "ctx_attn_inst.inc"
