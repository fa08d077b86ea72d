// SIMT paged-attention instantiations: query/out f32, cache f32.
#define INST_T float
#define INST_CT float
#include "attn_inst.inc"
