// SIMT paged-attention instantiations: query/out f16, cache e5m2.
#define INST_T __half
#define INST_CT uint8_t
#include "attn_inst.inc"
