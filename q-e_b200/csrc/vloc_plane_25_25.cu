// explicit instantiation of the fused plane kernel for a 25 x 25 (x,y) plane
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP25, SP25, 512>;
}
