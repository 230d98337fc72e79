// explicit instantiation of the plane2 kernel (vloc_plane2.cuh): 125 x 64 plane, x = 25*5, 16 warps
#define QEB_PLANE2_INSTANTIATE
#include "vloc_plane2_inst.cuh"
namespace qeb {
template struct Plane2Launch<SP125, SP64, 512>;
}
