// explicit instantiation of the fused plane kernel for a 64 x 64 (x,y) plane
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP64, SP64, 512>;
}
