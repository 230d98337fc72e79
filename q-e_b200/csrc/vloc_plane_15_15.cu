// explicit instantiation of the fused plane kernel for a 15 x 15 (x,y) plane
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP15, SP15, 512>;
}
