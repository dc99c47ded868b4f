#include "oz_common.cuh"
#include "oz_tm_launch.cuh"
namespace oz {
int pde_dispatch(const oz_problem*, int, size_t*, const EvalArgs*) { set_error("grid RHS not built yet"); return -1; }
}
