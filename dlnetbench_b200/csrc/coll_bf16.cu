// coll_bf16.cu — collective kernels instantiated for __nv_bfloat16 (see coll_inst.inc)
#define DLNB_T __nv_bfloat16
#define DLNB_SUFFIX bf16
#include "coll_inst.inc"
