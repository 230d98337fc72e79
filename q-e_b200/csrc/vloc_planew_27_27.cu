// explicit instantiation of the warp-local plane kernel (vloc_planew.cuh) for a 27 x 27 (x,y) plane
#define QEB_PLANEW_INSTANTIATE
#include "vloc_planew_inst.cuh"
namespace qeb {
template struct PlaneWLaunch<SP27, SP27, 512>;
}
