// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 64 x 64 plane
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP64, SP64, 512>;
}
