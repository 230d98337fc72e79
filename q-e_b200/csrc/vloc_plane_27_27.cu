// explicit instantiation of the fused plane kernel for a 27 x 27 (x,y) plane
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP27, SP27, 512>;
}
