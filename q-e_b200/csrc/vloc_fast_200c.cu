// explicit instantiation of the fast V_loc kernels for FFT length 200 = 20 x 10 (tuning alternative, QEB_ZPLAN)
#define QEB_FAST_INSTANTIATE
#include "vloc_fast_inst.cuh"
namespace qeb {
template struct FastLaunch<SP200c>;
}
