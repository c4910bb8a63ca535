// -fmad=false variant: every + - * / sqrt rounds as written (bit-exact ids vs the host oracle)
#define RTB_NS strict
#include "rtb_render.cuh"
