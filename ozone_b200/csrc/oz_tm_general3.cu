// Time-marching kernels (forward + reverse sweep, all stage counts, fp64 + fp32) for one
// registered right-hand side.  One translation unit per RHS so the build parallelises.
#include "oz_rhs.cuh"
#include "oz_tm_launch.cuh"
OZ_INSTANTIATE_TM(oz::RhsGeneral3, general3)
