// explicit instantiation of the warp-local plane kernel (vloc_planew.cuh) for a 25 x 25 (x,y) plane
#define QEB_PLANEW_INSTANTIATE
#include "vloc_planew_inst.cuh"
namespace qeb {
template struct PlaneWLaunch<SP25, SP25, 512>;
}
