// coll_f16.cu — collective kernels instantiated for __half (see coll_inst.inc)
#define DLNB_T __half
#define DLNB_SUFFIX f16
#include "coll_inst.inc"
