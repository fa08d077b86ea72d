// SIMT paged-attention instantiations: query/out f16, cache f16.
#define INST_T __half
#define INST_CT __half
#include "attn_inst.inc"
