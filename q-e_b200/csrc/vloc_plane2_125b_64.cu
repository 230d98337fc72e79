// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 125 x 64 plane, x = 5*5*5 (low-register plan), 25 warps
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP125b, SP64, 800>;
}
