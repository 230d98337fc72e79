// explicit instantiation of the fused plane kernel: low-register plans SP125b x SP64, 1024 threads
#define QEB_PLANE_INSTANTIATE
#include "vloc_plane_inst.cuh"
namespace qeb {
template struct PlaneLaunch<SP125b, SP64, 1024>;
}
