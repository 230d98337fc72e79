// explicit instantiation of the fused plane kernel for a 125 x 64 (x,y) plane
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP125, SP64, 512>;
}
