// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 27 x 27 plane
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP27, SP27, 512>;
}
