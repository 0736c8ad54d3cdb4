// coll_f32.cu — collective kernels instantiated for float (see coll_inst.inc)
#define DLNB_T float
#define DLNB_SUFFIX f32
#include "coll_inst.inc"
