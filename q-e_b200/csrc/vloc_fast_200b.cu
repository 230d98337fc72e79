// explicit instantiation of the fast V_loc kernels for FFT length 200 = 8 x 5 x 5 (low-register plan)
#define QEB_FAST_INSTANTIATE
#include "vloc_fast_inst.cuh"
namespace qeb {
template struct FastLaunch<SP200b>;
}
