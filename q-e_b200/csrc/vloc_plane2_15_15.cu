// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 15 x 15 plane
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP15, SP15, 512>;
}
