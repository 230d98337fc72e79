// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 25 x 25 plane
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP25, SP25, 512>;
}
