// default-flags variant (-fmad=true, like the reference's own nvcc build)
#define RTB_NS fast
#include "rtb_render.cuh"
