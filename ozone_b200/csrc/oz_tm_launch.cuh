// Host-side launch + (stage count, explicit/implicit, dtype) dispatch for one right-hand side.
// Each oz_tm_<rhs>.cu instantiates dispatch_tm<R> so the RHS files compile in parallel.
#pragma once
#include "oz_tm_kernels.cuh"
#include "oz_picard_kernels.cuh"
#include "oz_colloc_kernels.cuh"

namespace oz {

#ifndef OZ_ADJ_SPEC_STAGES   /* stage count whose explicit reverse sweep also gets the specialised kernels (RK4) */
#define OZ_ADJ_SPEC_STAGES 4
#endif

enum TmMode { TM_FORWARD = 0, TM_ADJOINT = 1, TM_ADJOINT_WS = 2, TM_RHS_EVAL = 3, TM_OUT_VJP = 4, TM_JAC_EVAL = 5, TM_JAC_PATTERN = 6 };

struct EvalArgs {
    long long nodes;
    const void *y, *p;
    void *dy, *prof, *field;
    // TM_OUT_VJP: cotangents of the outputs in, gradients per node out
    const void *vprof, *vfield;
    void *yb, *pb, *tb;
    long long blocks;     // TM_JAC_EVAL: number of [D][nodes] blocks (one per (step, stage))
    int compact;          // TM_JAC_EVAL flags: bit 0 write only the entries JP / JPP allow, bit 1 `p` is raw (interpolate to the stages)
};

template <class R, typename T, int S, bool IMPL>
int pick_block(size_t* smem) {
    if constexpr (!IMPL) {
        *smem = 0;
        return 128;
    } else if constexpr (use_reg_solve<R, S, T>()) {
        *smem = 0;      // Newton matrix lives in registers
        return OZ_IMPL_BLOCK;      // many small blocks: ~200 registers per thread, short tail on 148 SMs
    } else {
        const size_t per = SmemSystem<T, S * R::D>::bytes_per_thread();
        int block = 64;
        while (block > 32 && per * block > kSmemBudget) block >>= 1;
        *smem = per * block;
        return per * block <= 227 * 1024 ? block : -1;
    }
}

template <class R, typename T, int S, bool IMPL>
int run_tm(const oz_problem* pb, int mode, size_t* ws_bytes) {
    size_t smem = 0;
    const int block = pick_block<R, T, S, IMPL>(&smem);
    OZ_REQUIRE(block > 0, "implicit system of %d unknowns per step does not fit in shared memory", S * R::D);
    cudaStream_t stream = (cudaStream_t)pb->stream;
    const long long nblocks = (pb->N + block - 1) / block;
    OZ_REQUIRE(nblocks < (1ll << 31), "batch too large for one launch");
    TmArgs<T> a = make_tm_args<T>(pb);

    if (mode == TM_FORWARD) {
        auto kern = tm_forward_kernel<R, T, S, IMPL>;
        if (smem > 48 * 1024)
            OZ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblocks, block, smem, stream>>>(a);
        OZ_CUDA_CHECK(cudaGetLastError());
        return 0;
    }
    const AdjWs w = tm_adjoint_ws_layout<R>(pb, block);
    if (mode == TM_ADJOINT_WS) {
        *ws_bytes = (size_t)w.total * sizeof(T);
        return 0;
    }
    OZ_REQUIRE(pb->hist || (pb->ckpt && pb->num_inner > 0), "adjoint needs hist, or ckpt with num_inner > 0");
    OZ_REQUIRE((size_t)w.total * sizeof(T) <= pb->ws_bytes && (w.total == 0 || pb->ws),
               "adjoint workspace too small: need %zu bytes, got %zu", (size_t)w.total * sizeof(T), pb->ws_bytes);
    a.W = w.W;
    AdjWsPtr<T> wp;
    wp.h = a.ws + w.off_h;
    wp.p = a.ws + w.off_p;
    wp.th = a.ws + w.off_th;
    wp.seg = a.ws + w.off_seg;
    wp.W = w.W;
    // the specialised reverse sweeps (see tm_adjoint_kernel): final-state cotangent only, shared step vector with h_bar,
    // p_bar wanted, full history; explicit methods additionally need the RK4-shaped tableau the specialisation pins
    constexpr bool kSpecOk = R::P > 0 && ((!IMPL && S == OZ_ADJ_SPEC_STAGES) || (IMPL && use_reg_solve<R, S, T>()));
    int spec = 0;
    if constexpr (kSpecOk) {
        if ((IMPL || a.chain) && pb->hist && pb->final_bar && !pb->hist_bar && !pb->prof_bar && !pb->field_bar && pb->h_shared &&
            pb->h_bar && pb->p_bar)
            spec = pb->p_shared ? 1 : 2;
    }
    auto launch = [&](auto kern) -> int {
        if (smem > 48 * 1024)
            OZ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblocks, block, smem, stream>>>(a, wp);
        OZ_CUDA_CHECK(cudaGetLastError());
        return 0;
    };
    int rc = 0;
    if constexpr (kSpecOk) {
        if (spec == 1)
            rc = launch(tm_adjoint_kernel<R, T, S, IMPL, 1>);
        else if (spec == 2)
            rc = launch(tm_adjoint_kernel<R, T, S, IMPL, 2>);
        else
            rc = launch(tm_adjoint_kernel<R, T, S, IMPL, 0>);
    } else {
        rc = launch(tm_adjoint_kernel<R, T, S, IMPL, 0>);
    }
    if (rc) return rc;
    if (w.rows_h) reduce_rows_kernel<T><<<(unsigned)w.rows_h, 256, 0, stream>>>(wp.h, w.W, a.h_bar);
    if (w.rows_p) reduce_rows_kernel<T><<<(unsigned)w.rows_p, 256, 0, stream>>>(wp.p, w.W, a.p_bar);
    if (w.rows_th) reduce_rows_kernel<T><<<(unsigned)w.rows_th, 256, 0, stream>>>(wp.th, w.W, a.th_bar);
    OZ_CUDA_CHECK(cudaGetLastError());
    return 0;
}

template <class R, typename T>
int dispatch_stage(const oz_problem* pb, int mode, size_t* ws_bytes) {
    const int s = pb->num_stages;
#define OZ_TM_CASE(SS, IM)                                                             \
    if constexpr (OZ_WANT_S(SS) && OZ_WANT_IMPL(IM) && OZ_WANT_KIND(0))                \
        if (s == SS && (pb->is_explicit != 0) == !(IM)) return run_tm<R, T, SS, IM>(pb, mode, ws_bytes);
    OZ_TM_CASE(1, false)
    OZ_TM_CASE(2, false)
    OZ_TM_CASE(3, false)
    OZ_TM_CASE(4, false)
    OZ_TM_CASE(7, false)
    OZ_TM_CASE(1, true)
    OZ_TM_CASE(2, true)
    OZ_TM_CASE(3, true)
#undef OZ_TM_CASE
    set_error("no kernel for %s method with %d stages", pb->is_explicit ? "an explicit" : "an implicit", s);
    return -1;
}

template <class R>
int dispatch_tm(const oz_problem* pb, int mode, size_t* ws_bytes, const EvalArgs* ev) {
    if (mode == TM_JAC_PATTERN) {  // host only: which entries of df/dy and df/dp the functor declares possibly nonzero
        unsigned char* jp = (unsigned char*)ev->yb;
        unsigned char* jpp = (unsigned char*)ev->pb;
        if (jp)
            for (int e = 0; e < R::D * R::D; ++e) jp[e] = jac_nz<R>(e / R::D, e % R::D) ? 1 : 0;
        if constexpr (R::P > 0) {
            if (jpp)
                for (int e = 0; e < R::D * R::P; ++e) jpp[e] = jacp_nz<R>(e / R::P, e % R::P) ? 1 : 0;
        }
        return 0;
    }
    if (mode == TM_JAC_EVAL) {     // f -> ev->dy, Jy -> ev->yb, Jp -> ev->pb
        cudaStream_t stream = (cudaStream_t)pb->stream;
        const dim3 grid((unsigned)((ev->nodes + 127) / 128), (unsigned)ev->blocks);
        if (grid.x == 0 || grid.y == 0) return 0;
        OZ_REQUIRE(ev->blocks <= 65535, "at most 65535 blocks per Jacobian launch");
        auto go = [&](auto tag) -> int {
            using T = decltype(tag);
            JacLerp<T> lerp;
            lerp.S = (ev->compact & 2) ? pb->num_stages : 0;      // bit 1: p is the raw grid-point array, interpolate here
            lerp.p_shared = pb->p_shared;
            for (int j = 0; j < kMaxS; ++j) lerp.C[j] = (T)pb->C[j];
            if (lerp.S > 0) OZ_REQUIRE(ev->blocks % lerp.S == 0, "stage mode: the block count must be a multiple of the stage count");
            if (ev->compact & 1)
                rhs_jac_kernel<R, T, true><<<grid, 128, 0, stream>>>(ev->nodes, pb->th_shared, (const T*)ev->y, (const T*)ev->p,
                                                                    (const T*)pb->th, (T*)ev->dy, (T*)ev->yb, (T*)ev->pb, lerp);
            else
                rhs_jac_kernel<R, T, false><<<grid, 128, 0, stream>>>(ev->nodes, pb->th_shared, (const T*)ev->y, (const T*)ev->p,
                                                                     (const T*)pb->th, (T*)ev->dy, (T*)ev->yb, (T*)ev->pb, lerp);
            OZ_CUDA_CHECK(cudaGetLastError());
            return 0;
        };
        if constexpr (OZ_WANT_F64(true))
            if (pb->dtype == OZ_F64) return go(double{});
        if constexpr (OZ_WANT_F64(false))
            if (pb->dtype == OZ_F32) return go(float{});
        set_error("no Jacobian kernel for dtype %d in this build", pb->dtype);
        return -1;
    }
    if (mode == TM_OUT_VJP) {
        cudaStream_t stream = (cudaStream_t)pb->stream;
        const unsigned grid = (unsigned)((ev->nodes + 127) / 128);
        if (grid == 0) return 0;
        OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
        if constexpr (OZ_WANT_F64(true)) {
            if (pb->dtype == OZ_F64) {
                rhs_out_vjp_kernel<R, double><<<grid, 128, 0, stream>>>(
                    ev->nodes, pb->th_shared, (const double*)ev->y, (const double*)ev->p, (const double*)pb->th,
                    (const double*)ev->vprof, (const double*)ev->vfield, (double*)ev->yb, (double*)ev->pb, (double*)ev->tb);
                OZ_CUDA_CHECK(cudaGetLastError());
                return 0;
            }
        }
        if constexpr (OZ_WANT_F64(false)) {
            if (pb->dtype == OZ_F32) {
                rhs_out_vjp_kernel<R, float><<<grid, 128, 0, stream>>>(
                    ev->nodes, pb->th_shared, (const float*)ev->y, (const float*)ev->p, (const float*)pb->th,
                    (const float*)ev->vprof, (const float*)ev->vfield, (float*)ev->yb, (float*)ev->pb, (float*)ev->tb);
                OZ_CUDA_CHECK(cudaGetLastError());
                return 0;
            }
        }
        set_error("no output-VJP kernel for dtype %d in this build", pb->dtype);
        return -1;
    }
    if (mode == TM_RHS_EVAL) {
        cudaStream_t stream = (cudaStream_t)pb->stream;
        const unsigned grid = (unsigned)((ev->nodes + 127) / 128);
        if (grid == 0) return 0;
        OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
        if constexpr (OZ_WANT_F64(true)) {
            if (pb->dtype == OZ_F64)
                rhs_eval_kernel<R, double><<<grid, 128, 0, stream>>>(ev->nodes, pb->th_shared, (const double*)ev->y,
                                                                   (const double*)ev->p, (const double*)pb->th,
                                                                   (double*)ev->dy, (double*)ev->prof, (double*)ev->field);
        }
        if constexpr (OZ_WANT_F64(false)) {
            if (pb->dtype == OZ_F32)
                rhs_eval_kernel<R, float><<<grid, 128, 0, stream>>>(ev->nodes, pb->th_shared, (const float*)ev->y,
                                                                  (const float*)ev->p, (const float*)pb->th, (float*)ev->dy,
                                                                  (float*)ev->prof, (float*)ev->field);
        }
        OZ_CUDA_CHECK(cudaGetLastError());
        return 0;
    }
    if constexpr (OZ_WANT_F64(true))
        if (pb->dtype == OZ_F64) return dispatch_stage<R, double>(pb, mode, ws_bytes);
    if constexpr (OZ_WANT_F64(false))
        if (pb->dtype == OZ_F32) return dispatch_stage<R, float>(pb, mode, ws_bytes);
    set_error("no kernel for dtype %d in this build", pb->dtype);
    return -1;
}

}  // namespace oz

#define OZ_INSTANTIATE_TM(RHS, NAME)                                                                     \
    namespace oz {                                                                                       \
    int tm_##NAME(const oz_problem* pb, int mode, size_t* ws_bytes, const EvalArgs* ev) {                \
        return dispatch_tm<RHS>(pb, mode, ws_bytes, ev);                                                 \
    }                                                                                                    \
    int picard_##NAME(const oz_problem* pb, const void* guess, double tol, int max_sweeps, const int* blk_lo, \
                      int nblocks) {                                                                     \
        return dispatch_picard<RHS>(pb, guess, tol, max_sweeps, blk_lo, nblocks);                        \
    }                                                                                                    \
    int colloc_##NAME(const oz_problem* pb, int mode, size_t* ws_bytes, const void* Y, const void* y,    \
                      const void* mu1, const void* mu2, void* o1, void* o2) {                            \
        return dispatch_colloc<RHS>(pb, mode, ws_bytes, Y, y, mu1, mu2, o1, o2);                         \
    }                                                                                                    \
    }
