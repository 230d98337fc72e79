// explicit instantiation of the fast V_loc kernels for FFT length 144
#define QEB_FAST_INSTANTIATE
#include "vloc_fast_inst.cuh"
namespace qeb {
template struct FastLaunch<SP144>;
}
