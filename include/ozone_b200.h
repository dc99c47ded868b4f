/*
 * ozone_b200 -- C ABI of the B200-native solve-and-differentiate engine.
 *
 * The reference (LSDOlab/ozone) is pure Python and has NO native interface: its integrators
 * hand every operation to csdl_alpha.  These entry points are therefore what a binding for the
 * hot path WOULD bind; each one names the reference code it replaces (paths relative to the
 * reference checkout).  INTEGRATION.md shows the ctypes stub a maintainer would add.
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer owned by the caller; the library never allocates
 *     or frees user-visible memory.  Scratch comes from a caller-provided workspace whose size
 *     is returned by the *_workspace_bytes query.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, no host sync
 *     unless stated.
 *   - return value: 0 ok; < 0 invalid argument / CUDA error (text via oz_last_error()).
 *     Numerical trouble (Newton / Picard not converged, non-finite) is reported per trajectory
 *     in the `status` array, never by the return code.
 *   - dtype: OZ_F64 or OZ_F32 selects the arithmetic AND storage type of every real array.
 *
 * Device layout (structure of arrays, trajectory index innermost):
 *     y0      [d][N]            initial conditions                 (collections.py:175-182)
 *     h       [n]  or [n][N]    step sizes        (h_shared = 1/0) (timespan.py:63-76)
 *     p       [nt][P] or [nt][P][N]  dynamic parameters at the nt = n+1 grid points
 *                               (p_shared = 1/0)                   (collections.py:239-249)
 *     th      [S]  or [S][N]    static parameters (th_shared = 1/0)
 *     hist    [nt][d][N]        state history incl. the IC row     (time_marching.py:120)
 *     ckpt    [nc+1][d][N]      states at the checkpoint boundaries (time_marching_uniform.py:43,128)
 *     prof    [nt][NPROF][N]    profile outputs                    (time_marching.py:97-122)
 *     field   [NFIELD][N]       field outputs (sum over time)      (time_marching.py:104-112)
 * For the grid right-hand side (OZ_RHS_DIFFUSION_REACTION) there is ONE trajectory whose
 * components are the grid cells: N = 1 and d = nx (set in `d_override`).
 */
#ifndef OZONE_B200_H
#define OZONE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define OZ_API __attribute__((visibility("default")))
#else
#define OZ_API
#endif

#define OZ_MAX_STAGES 7

enum { OZ_F64 = 0, OZ_F32 = 1 };

/* Registered device right-hand sides: the reference's Python callbacks restated as device
 * functors behind func_wrapper (ozone/func_wrapper.py:90-107). */
enum {
    OZ_RHS_LOTKA_VOLTERRA = 0,      /* examples/simple_examples/lotka_volterra.py:51-62 */
    OZ_RHS_VAN_DER_POL = 1,         /* ozone/paper_examples/van_der_pol/main_model.py:6-14 */
    OZ_RHS_LUNAR_ASCENT = 2,        /* ozone/paper_examples/ascent_system/main_model.py:6-21 */
    OZ_RHS_AFFINE = 3,              /* ozone/tests/test_correctness/test_1_simple/utils.py:6-7; tutorial.ipynb */
    OZ_RHS_QUAD_DECAY = 4,          /* ozone/tests/test_correctness/test_3_large/utils.py:23-30 */
    OZ_RHS_GENERAL3 = 5,            /* ozone/tests/test_correctness/test_2_general/utils.py:43-84 */
    OZ_RHS_DIFFUSION_REACTION = 6,  /* ozone/paper_examples/pde_control/pde_control.py:11-81 */
    OZ_RHS_COUNT = 7,
    OZ_RHS_PLUGIN_BASE = 1000       /* ids handed out by oz_register_rhs_plugin */
};

/* Slab mode runs a long solve as consecutive time segments (one call each).  The state at a segment
 * boundary belongs to both segments; these flags keep its output terms from being counted twice. */
enum {
    OZ_GRID_SKIP_IC_FIELD = 1,        /* forward: the k = 0 term of the field sum was counted by the previous segment */
    OZ_GRID_SKIP_TERMINAL_FIELD = 2,  /* adjoint: the k = n direct field cotangent is added by the next segment */
    OZ_GRID_NO_CONTROL_COST = 4       /* this rank does not own the 20 u^2 term of J_space (one rank does) */
};

enum { OZ_STATUS_OK = 0, OZ_STATUS_NOT_CONVERGED = 1, OZ_STATUS_SINGULAR = 2, OZ_STATUS_NONFINITE = 4 };

/* One descriptor for the forward solve and its adjoint.  Unused pointers are NULL. */
typedef struct oz_problem {
    /* ---- method: ozone/glms/GLMs.py:19-42 (A, B, C = A.1, explicit flag); U = 1, V = 1 ---- */
    int32_t rhs_id;
    int32_t dtype;
    int32_t num_stages;
    int32_t is_explicit;
    double A[OZ_MAX_STAGES * OZ_MAX_STAGES];   /* row-major, leading dimension OZ_MAX_STAGES */
    double B[OZ_MAX_STAGES];
    double C[OZ_MAX_STAGES];
    /* ---- sizes ---- */
    int64_t N;            /* trajectories */
    int32_t n;            /* steps; nt = n + 1 */
    int32_t num_inner;    /* checkpointing: inner steps per segment (time_marching_uniform.py:27); 0 = none */
    int32_t h_shared, p_shared, th_shared;
    int32_t d_override;   /* grid RHS only: number of cells nx */
    /* ---- implicit stage solver: time_marching.py:217-221 (Newton, tolerance 1e-7) ---- */
    double newton_tol;
    int32_t newton_max_iter;
    int32_t reserved0;
    double rhs_cfg[8];    /* RHS constants that are not differentiated (grid RHS: dx, beta_t, beta_u, gamma, H_lo, H_hi) */
    /* Picard only: the packed-component ranges of the states in insertion order, relaxed one after
     * another (block Gauss-Seidel, picard_iteration.py:35).  0 blocks = one block = Jacobi. */
    int32_t num_state_blocks;
    int32_t state_block_lo[9];
    /* Grid RHS, slab mode (grid_nx_global > 0): this call integrates cells [grid_x_off, grid_x_off + d_override)
     * of a grid of grid_nx_global cells; only local cells [grid_own_lo, grid_own_hi) count in the grid-wide
     * sums (J_space, u_bar, h_bar).  Used by the spatial shard; 0 = the whole grid is local. */
    int64_t grid_x_off, grid_nx_global, grid_own_lo, grid_own_hi;
    int64_t grid_flags;   /* OZ_GRID_* bits: how a time segment of a sharded solve joins its neighbours */
    /* ---- inputs ---- */
    const void* y0;
    const void* h;
    const void* p;
    const void* th;
    /* ---- forward outputs ---- */
    void* hist;
    void* ckpt;
    void* yfinal;
    void* prof;
    void* field;
    int32_t* status;      /* [N] OR-ed OZ_STATUS_* flags, or NULL */
    int32_t* iters;       /* [N] total Newton iterations (implicit) / sweeps (Picard), or NULL */
    /* ---- adjoint inputs: cotangents of the forward outputs ---- */
    const void* hist_bar;
    const void* final_bar;
    const void* prof_bar;
    const void* field_bar;
    /* ---- adjoint outputs (gradients, same layout as the inputs; shared inputs get the sum) ---- */
    void* y0_bar;
    void* h_bar;
    void* p_bar;
    void* th_bar;
    /* ---- scratch + stream ---- */
    void* ws;
    size_t ws_bytes;
    void* stream;
    /* ---- optional stage cache (implicit methods): [n][s][d][N].  oz_tm_forward stores the converged
     * stage values Y here; oz_tm_adjoint, given the same buffer, reads them instead of repeating the
     * Newton solve of every step (it then only factors the transposed system once per step). ---- */
    void* stages;
} oz_problem;

/* Library / build identification. */
OZ_API int oz_version(void);
OZ_API const char* oz_last_error(void);

/* Sizes of a registered right-hand side: what FuncWrapper discovers on its first call
 * (ozone/func_wrapper.py:100-104).  kind: 0 = per-trajectory (batched), 1 = grid. */
OZ_API int oz_rhs_info(int rhs_id, int* d, int* P, int* S, int* nprof, int* nfield, int* kind);

/* Evaluate f (and the outputs) at `nodes` points: FuncWrapper.evaluate (ozone/func_wrapper.py:90-107).
 * y [d][nodes], p [P][nodes], th [S] or [S][nodes]; dy [d][nodes], prof [NPROF][nodes], field [NFIELD][nodes]. */
OZ_API int oz_rhs_eval(const oz_problem* pb, int64_t nodes, const void* y, const void* p, void* dy, void* prof, void* field);

/* Transposed Jacobians of the profile / field outputs at `nodes` points: the reverse pass of the outputs that
 * VectorBased.integrate evaluates on a whole history (ozone/integrators/vector_based.py:33-49; csdl differentiates
 * them as graph nodes).  vprof [NPROF][nodes], vfield [NFIELD][nodes] (NULL = zero) -> y_bar [d][nodes],
 * p_bar [P][nodes], th_bar [S][nodes]: one gradient per node, the caller sums what is shared.  NULL outputs are skipped. */
OZ_API int oz_rhs_out_vjp(const oz_problem* pb, int64_t nodes, const void* y, const void* p, const void* vprof,
                          const void* vfield, void* y_bar, void* p_bar, void* th_bar);

/* f and its Jacobians at blocks x nodes points, trajectory index innermost: y [blocks][d][nodes], p [blocks][P][nodes] ->
 * f [blocks][d][nodes], Jy [blocks][d][d][nodes] (d f_r / d y_c at [r][c]), Jp [blocks][d][P][nodes]; NULL = skip.
 * With blocks = n s and nodes = N the inputs are the Collocation stage variables as they lie in memory, and the blocks
 * of the constraint Jacobian are products of these with tableau entries and step sizes
 * (ozone/integrators/collocation.py:48-57: c1 = A hF + U y - Y, c2 = B hF + V y - y; the reference leaves the
 * differentiation to csdl, an optimiser consumes it as `dc`, ozone/paper_examples/utils/utils.py:237,273-286). */
enum { OZ_JAC_COMPACT = 1,        /* write only the entries oz_rhs_jac_pattern marks (see below) */
       OZ_JAC_STAGE_PARAMS = 2 }; /* blocks = steps x stages and `p` is the RAW array at the grid points, [steps+1][P][nodes] or
                                     [steps+1][P] (p_shared): interpolated to the stage abscissae with pb->C inside the kernel */
OZ_API int oz_rhs_jac_eval(const oz_problem* pb, int64_t nodes, int64_t blocks, int32_t flags, const void* y, const void* p,
                           void* f, void* Jy, void* Jp);
/* Which entries of df/dy (jp [d*d], row-major) and df/dp (jpp [d*P]) the right-hand side declares possibly nonzero
 * (1) or identically zero (0); NULL = skip.  With OZ_JAC_COMPACT, oz_rhs_jac_eval writes only the possibly-nonzero entries,
 * in row-major order: Jy [blocks][nnz_y][nodes], Jp [blocks][nnz_p][nodes] (lunar ascent: 9 of 49 and 3 of 21). */
OZ_API int oz_rhs_jac_pattern(int rhs_id, uint8_t* jp, uint8_t* jpp);

/* TimeMarching.integrate: ozone/integrators/time_marching.py:6-59 (loop), :143-245 (stages,
 * explicit and Newton), :247-260 (update), :97-141 (outputs).  With num_inner > 0 it is
 * TimeMarchingUniform.integrate (ozone/integrators/time_marching_uniform.py:7-150): only `ckpt`
 * needs to be retained for the reverse sweep. */
OZ_API size_t oz_tm_forward_workspace_bytes(const oz_problem* pb);   /* 0 for per-trajectory RHS; grid RHS: ping-pong + partial sums */
OZ_API int oz_tm_forward(const oz_problem* pb);

/* Reverse sweep of the same loop: what csdl derives from the stacked loop
 * (time_marching.py:51 stack_all=True; call sites examples/simple_examples/lotka_volterra.py:76-79,
 * ozone/tests/run_ode.py:66-69).  Reads `hist` (or `ckpt` + recompute when num_inner > 0). */
OZ_API size_t oz_tm_adjoint_workspace_bytes(const oz_problem* pb);
OZ_API int oz_tm_adjoint(const oz_problem* pb);

/* PicardIteration.solve: ozone/integrators/picard_iteration.py:9-65 (sweep), vector_based.py:33-53
 * (outputs).  `guess` [n][d][N] or NULL (collections.py:125-133).  Writes hist/prof/field/yfinal,
 * per-trajectory sweep counts in `iters`, OZ_STATUS_NOT_CONVERGED in `status`. */
OZ_API size_t oz_picard_workspace_bytes(const oz_problem* pb);
OZ_API int oz_picard_forward(const oz_problem* pb, const void* guess, double tol, int max_sweeps);

/* Collocation.solve: ozone/integrators/collocation.py:33-50.  Y [n][s][d][N], y [n][d][N];
 * c1 [n][s][d][N], c2 [n][d][N].  The VJP returns (dc)^T mu for an optimiser
 * (ozone/tests/run_ode.py:49-58): Y_bar, y_bar plus pb->y0_bar/h_bar/p_bar/th_bar. */
OZ_API int oz_colloc_residual(const oz_problem* pb, const void* Y, const void* y, void* c1, void* c2);
OZ_API size_t oz_colloc_vjp_workspace_bytes(const oz_problem* pb);
OZ_API int oz_colloc_vjp(const oz_problem* pb, const void* Y, const void* y, const void* mu1, const void* mu2,
                  void* Y_bar, void* y_bar);

/* Measurement aid (no counterpart in the reference): enqueue a register-resident FMA loop and return its
 * flop count, so the caller can time the FP64 (or FP32) vector peak of the box with events.  bench.py reports it
 * beside the HBM roofline because the implicit path is FP64-bound.  `out`: one device scalar. */
OZ_API double oz_fma_probe(int dtype, int iters, void* out, void* stream);

/* Register a right-hand side compiled separately (ozone_b200/trace.py turns a Python callback into CUDA
 * source and builds it against the same kernel templates).  `so_path` is a shared object exporting
 * `oz_plugin_describe`; returns the rhs_id to put in oz_problem.rhs_id (>= OZ_RHS_PLUGIN_BASE) or < 0.
 * This is what stands in for "any Python f" of ozone/func_wrapper.py:90-107.  Two kinds (oz_rhs_info reports which): one ODE per
 * trajectory, or a grid model (ozone_b200/trace_grid.py, from a whole-grid callback like pde_control.py:11-81) that runs through
 * the fused stencil kernels of the built-in diffusion-reaction system; a grid model's per-cell arrays are passed as `th`. */
OZ_API int oz_register_rhs_plugin(const char* so_path);

/* No counterpart in the reference (it has no distributed code).  Plumbing of the two multi-GPU paths over peer-mapped
 * memory (NVLink / NVSwitch; each rank holds device pointers into every other rank's symmetric-memory arena): plain
 * stores into the peer's arena + a release store of an epoch number, acquire polling on the receiver's own memory.
 * Epochs are *epoch_base + epoch so a captured CUDA graph can be replayed (oz_peer_epoch_advance moves the base).  A wait
 * that sees no signal for ~2 s sets *err to 1 and returns without copying.  NULL pointers = no neighbour on that side.
 *   oz_peer_halo_push / oz_peer_halo_wait (/ oz_peer_halo_exchange = both in one launch): neighbour exchange of the spatial shard (ozone_b200/parallel.py), `count`
 *       values per side; push writes into the neighbours' slots, wait copies this rank's slots into its halo zones.
 *   oz_peer_sum: all-reduce (sum, fixed rank order) of `rows` values: slots[r] -> rank r's [2][world][rows] slot array,
 *       flags[r] -> rank r's [world] flags (both device arrays of `world` pointers). */
OZ_API int oz_peer_halo_push(int dtype, const void* left_src, const void* right_src, int64_t count, void* left_dst,
                             void* right_dst, uint64_t* left_flag, uint64_t* right_flag, const uint64_t* epoch_base,
                             uint64_t epoch, void* stream);
OZ_API int oz_peer_halo_wait(int dtype, const void* from_left_slot, const void* from_right_slot, int64_t count,
                             void* left_halo, void* right_halo, const uint64_t* flag_left, const uint64_t* flag_right,
                             const uint64_t* epoch_base, uint64_t epoch, int32_t* err, void* stream);
/* push + wait in one launch: only when every rank's kernel can be resident at once (one process per GPU) */
OZ_API int oz_peer_halo_exchange(int dtype, const void* left_src, const void* right_src, int64_t count, void* left_dst,
                                 void* right_dst, uint64_t* left_flag, uint64_t* right_flag, const void* from_left_slot,
                                 const void* from_right_slot, void* left_halo, void* right_halo, const uint64_t* flag_left,
                                 const uint64_t* flag_right, const uint64_t* epoch_base, uint64_t epoch, int32_t* err,
                                 void* stream);
OZ_API int oz_peer_epoch_advance(uint64_t* epoch_base, uint64_t by, void* stream);
OZ_API int oz_peer_sum(int dtype, const void* local, int32_t rows, int32_t rank, int32_t world, void* const* slots,
                       uint64_t* const* flags, const uint64_t* epoch_base, uint64_t epoch, void* out, int32_t* err,
                       void* stream);

/* No counterpart in the reference.  Same reverse sweep as oz_tm_adjoint, for callers that cannot
 * allocate device memory (plug-ins inside a host application that owns the GPU, real-time loops,
 * FFI stubs without an allocator): pb->ws / pb->ws_bytes are ignored and the scratch is a 4 MiB
 * array built into the library (oz_static_scratch_bytes()).  Problems needing more are refused
 * (rc < 0, see oz_last_error).  The first call on a device resolves the array's address: make it
 * outside stream capture.  All calls on a device share the array, so calls that may overlap on the
 * GPU race; nothing in it needs clearing between calls.  Allocates nothing. */
OZ_API size_t oz_static_scratch_bytes(void);
OZ_API int oz_tm_adjoint_static(const oz_problem* pb);

#ifdef __cplusplus
}
#endif
#endif /* OZONE_B200_H */
