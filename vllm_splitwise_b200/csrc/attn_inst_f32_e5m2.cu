// SIMT paged-attention instantiations: query/out f32, cache e5m2.
#define INST_T float
#define INST_CT uint8_t
#include "attn_inst.inc"
