// Picard routing: one dispatcher per registered right-hand side (instantiated in oz_tm_<rhs>.cu).
#include "oz_common.cuh"
namespace oz {
#define OZ_DECL(NAME) \
    int picard_##NAME(const oz_problem*, const void*, double, int, const int*, int);
OZ_DECL(lotka_volterra) OZ_DECL(van_der_pol) OZ_DECL(lunar_ascent) OZ_DECL(affine) OZ_DECL(quad_decay) OZ_DECL(general3)
#undef OZ_DECL

int picard_dispatch(const oz_problem* pb, int mode, size_t* ws_bytes, const void* guess, double tol, int max_sweeps) {
    if (mode == 1) {  // workspace query: the iterate lives in shared memory
        *ws_bytes = 0;
        return 0;
    }
    OZ_REQUIRE(pb->y0 && pb->h, "y0 and h are required");
    OZ_REQUIRE(tol > 0.0 && max_sweeps >= 0, "tolerance must be positive and max_sweeps non-negative");
    int blk[10] = {0};
    int nb = pb->num_state_blocks;
    int d = 0;
    oz_rhs_info(pb->rhs_id, &d, nullptr, nullptr, nullptr, nullptr, nullptr);
    if (nb <= 0) {
        nb = 1;
        blk[0] = 0;
        blk[1] = d;
    } else {
        OZ_REQUIRE(nb <= 8, "at most 8 state blocks");
        for (int b = 0; b <= nb; ++b) blk[b] = pb->state_block_lo[b];
    }
    switch (pb->rhs_id) {
        case OZ_RHS_LOTKA_VOLTERRA: return picard_lotka_volterra(pb, guess, tol, max_sweeps, blk, nb);
        case OZ_RHS_VAN_DER_POL: return picard_van_der_pol(pb, guess, tol, max_sweeps, blk, nb);
        case OZ_RHS_LUNAR_ASCENT: return picard_lunar_ascent(pb, guess, tol, max_sweeps, blk, nb);
        case OZ_RHS_AFFINE: return picard_affine(pb, guess, tol, max_sweeps, blk, nb);
        case OZ_RHS_QUAD_DECAY: return picard_quad_decay(pb, guess, tol, max_sweeps, blk, nb);
        case OZ_RHS_GENERAL3: return picard_general3(pb, guess, tol, max_sweeps, blk, nb);
    }
    set_error("PicardIteration is not available for rhs_id %d (grid right-hand sides use TimeMarching)", pb->rhs_id);
    return -1;
}
}  // namespace oz
