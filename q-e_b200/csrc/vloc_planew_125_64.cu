// explicit instantiation of the warp-local plane kernel (vloc_planew.cuh) for a 125 x 64 (x,y) plane
#define QEB_PLANEW_INSTANTIATE
#include "vloc_planew_inst.cuh"
namespace qeb {
template struct PlaneWLaunch<SP125, SP64, 512>;
}
