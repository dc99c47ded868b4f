#include "oz_common.cuh"
namespace oz {
int colloc_dispatch(const oz_problem*, int, size_t*, const void*, const void*, const void*, const void*, void*, void*) { set_error("collocation not built yet"); return -1; }
}
