// Collocation routing: one dispatcher per registered right-hand side (instantiated in oz_tm_<rhs>.cu).
#include "oz_common.cuh"
namespace oz {
#define OZ_DECL(NAME)                                                                                          \
    int colloc_##NAME(const oz_problem*, int, size_t*, const void*, const void*, const void*, const void*, void*, \
                      void*);
OZ_DECL(lotka_volterra) OZ_DECL(van_der_pol) OZ_DECL(lunar_ascent) OZ_DECL(affine) OZ_DECL(quad_decay) OZ_DECL(general3)
#undef OZ_DECL

int colloc_dispatch(const oz_problem* pb, int mode, size_t* ws_bytes, const void* Y, const void* y, const void* mu1,
                    const void* mu2, void* o1, void* o2) {
    switch (pb->rhs_id) {
        case OZ_RHS_LOTKA_VOLTERRA: return colloc_lotka_volterra(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
        case OZ_RHS_VAN_DER_POL: return colloc_van_der_pol(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
        case OZ_RHS_LUNAR_ASCENT: return colloc_lunar_ascent(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
        case OZ_RHS_AFFINE: return colloc_affine(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
        case OZ_RHS_QUAD_DECAY: return colloc_quad_decay(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
        case OZ_RHS_GENERAL3: return colloc_general3(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);
    }
    set_error("Collocation is not available for rhs_id %d", pb->rhs_id);
    return -1;
}
}  // namespace oz
