// explicit instantiation of the warp-local plane kernel (vloc_planew.cuh) for a 15 x 15 (x,y) plane
#define QEB_PLANEW_INSTANTIATE
#include "vloc_planew_inst.cuh"
namespace qeb {
template struct PlaneWLaunch<SP15, SP15, 512>;
}
