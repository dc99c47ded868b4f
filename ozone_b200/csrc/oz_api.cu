// extern "C" boundary of liboz_b200.so (declared in include/ozone_b200.h).
#include <cstdarg>
#include <dlfcn.h>
#include <atomic>
#include <mutex>
#include "oz_plugin.h"
#include "oz_common.cuh"
#include "oz_rhs.cuh"
#include "oz_tm_launch.cuh"

namespace oz {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// one dispatcher per batched RHS (defined in oz_tm_<rhs>.cu)
int tm_lotka_volterra(const oz_problem*, int, size_t*, const EvalArgs*);
int tm_van_der_pol(const oz_problem*, int, size_t*, const EvalArgs*);
int tm_lunar_ascent(const oz_problem*, int, size_t*, const EvalArgs*);
int tm_affine(const oz_problem*, int, size_t*, const EvalArgs*);
int tm_quad_decay(const oz_problem*, int, size_t*, const EvalArgs*);
int tm_general3(const oz_problem*, int, size_t*, const EvalArgs*);
// grid RHS (oz_pde.cu)
int pde_dispatch(const oz_problem*, int, size_t*, const EvalArgs*);
// parallel-in-time paths (oz_picard.cu, oz_colloc.cu)
int picard_dispatch(const oz_problem*, int mode, size_t* ws_bytes, const void* guess, double tol, int max_sweeps);
int colloc_dispatch(const oz_problem*, int mode, size_t* ws_bytes, const void* Y, const void* y, const void* mu1,
                    const void* mu2, void* o1, void* o2);

struct RhsInfo {
    int d, P, S, nprof, nfield, kind;
};
static const RhsInfo kInfo[OZ_RHS_COUNT] = {
    {RhsLotkaVolterra::D, RhsLotkaVolterra::P, RhsLotkaVolterra::S, RhsLotkaVolterra::NPROF, RhsLotkaVolterra::NFIELD, 0},
    {RhsVanDerPol::D, RhsVanDerPol::P, RhsVanDerPol::S, RhsVanDerPol::NPROF, RhsVanDerPol::NFIELD, 0},
    {RhsLunarAscent::D, RhsLunarAscent::P, RhsLunarAscent::S, RhsLunarAscent::NPROF, RhsLunarAscent::NFIELD, 0},
    {RhsAffine::D, RhsAffine::P, RhsAffine::S, RhsAffine::NPROF, RhsAffine::NFIELD, 0},
    {RhsQuadDecay::D, RhsQuadDecay::P, RhsQuadDecay::S, RhsQuadDecay::NPROF, RhsQuadDecay::NFIELD, 0},
    {RhsGeneral3::D, RhsGeneral3::P, RhsGeneral3::S, RhsGeneral3::NPROF, RhsGeneral3::NFIELD, 0},
    {-1, 1, 0, 0, 1, 1},  // diffusion-reaction: d = nx comes from d_override
};

// ---- right-hand sides registered at run time (plugins built by ozone_b200/trace.py) -------------------
constexpr int kMaxPlugins = 256;
static oz_plugin_info g_plugins[kMaxPlugins];
static std::atomic<int> g_num_plugins{0};   // published with release order after the slot is filled: readers take no lock
static std::mutex g_plugin_mu;

static const oz_plugin_info* plugin_of(int rhs_id) {
    const int slot = rhs_id - OZ_RHS_PLUGIN_BASE;
    if (slot < 0 || slot >= g_num_plugins.load(std::memory_order_acquire)) return nullptr;
    return &g_plugins[slot];
}
static bool known_rhs(int rhs_id) { return (rhs_id >= 0 && rhs_id < OZ_RHS_COUNT) || plugin_of(rhs_id) != nullptr; }
// 0: one ODE per trajectory, 1: grid (only valid for a known rhs_id)
static int rhs_kind(int rhs_id) {
    if (const oz_plugin_info* pl = plugin_of(rhs_id)) return pl->kind;
    return kInfo[rhs_id].kind;
}
static int plugin_rc(const oz_plugin_info* pl, int rc) {
    if (rc < 0 && pl->last_error) set_error("%s", pl->last_error());
    return rc;
}

static int check_problem(const oz_problem* pb) {
    OZ_REQUIRE(pb != nullptr, "null problem descriptor");
    OZ_REQUIRE(known_rhs(pb->rhs_id), "unknown rhs_id %d", pb->rhs_id);
    OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
    OZ_REQUIRE(pb->num_stages >= 1 && pb->num_stages <= OZ_MAX_STAGES, "num_stages %d out of range", pb->num_stages);
    OZ_REQUIRE(pb->N >= 1, "N must be >= 1 (got %lld)", (long long)pb->N);
    OZ_REQUIRE(pb->n >= 1, "at least one step is required (got %d)", pb->n);
    OZ_REQUIRE(pb->num_inner >= 0 && (pb->num_inner == 0 || pb->n % pb->num_inner == 0),
               "number of steps (%d) must be divisible by the inner segment length (%d)", pb->n, pb->num_inner);
    return 0;
}

static int route(const oz_problem* pb, int mode, size_t* ws_bytes, const EvalArgs* ev) {
    if (const oz_plugin_info* pl = plugin_of(pb->rhs_id)) return plugin_rc(pl, pl->tm(pb, mode, ws_bytes, ev));
    switch (pb->rhs_id) {
        case OZ_RHS_LOTKA_VOLTERRA: return tm_lotka_volterra(pb, mode, ws_bytes, ev);
        case OZ_RHS_VAN_DER_POL: return tm_van_der_pol(pb, mode, ws_bytes, ev);
        case OZ_RHS_LUNAR_ASCENT: return tm_lunar_ascent(pb, mode, ws_bytes, ev);
        case OZ_RHS_AFFINE: return tm_affine(pb, mode, ws_bytes, ev);
        case OZ_RHS_QUAD_DECAY: return tm_quad_decay(pb, mode, ws_bytes, ev);
        case OZ_RHS_GENERAL3: return tm_general3(pb, mode, ws_bytes, ev);
        case OZ_RHS_DIFFUSION_REACTION: return pde_dispatch(pb, mode, ws_bytes, ev);
    }
    set_error("unknown rhs_id %d", pb->rhs_id);
    return -1;
}

}  // namespace oz

using namespace oz;

extern "C" {

int oz_version(void) { return 100; }

const char* oz_last_error(void) { return g_err; }

int oz_register_rhs_plugin(const char* so_path) {
    OZ_REQUIRE(so_path != nullptr, "null plugin path");
    std::lock_guard<std::mutex> lock(g_plugin_mu);
    const int slot = g_num_plugins.load(std::memory_order_relaxed);
    OZ_REQUIRE(slot < kMaxPlugins, "too many right-hand side plugins (%d)", kMaxPlugins);
    void* handle = dlopen(so_path, RTLD_NOW | RTLD_LOCAL);
    OZ_REQUIRE(handle != nullptr, "cannot load plugin %s: %s", so_path, dlerror());
    auto describe = (void (*)(oz_plugin_info*))dlsym(handle, "oz_plugin_describe");
    OZ_REQUIRE(describe != nullptr, "%s does not export oz_plugin_describe", so_path);
    oz_plugin_info info;
    memset(&info, 0, sizeof(info));
    describe(&info);
    OZ_REQUIRE(info.abi == OZ_PLUGIN_ABI, "plugin %s was built for plugin ABI %d, this library speaks %d", so_path, info.abi,
               OZ_PLUGIN_ABI);
    OZ_REQUIRE(info.tm && info.picard && info.colloc && info.d >= 1 && (info.kind == 0 || info.kind == 1), "plugin %s is incomplete",
               so_path);
    g_plugins[slot] = info;
    g_num_plugins.store(slot + 1, std::memory_order_release);
    return OZ_RHS_PLUGIN_BASE + slot;
}

int oz_rhs_info(int rhs_id, int* d, int* P, int* S, int* nprof, int* nfield, int* kind) {
    OZ_REQUIRE(known_rhs(rhs_id), "unknown rhs_id %d", rhs_id);
    RhsInfo plug;
    if (const oz_plugin_info* pl = plugin_of(rhs_id)) plug = RhsInfo{pl->d, pl->P, pl->S, pl->nprof, pl->nfield, pl->kind};
    const RhsInfo& r = plugin_of(rhs_id) ? plug : kInfo[rhs_id];
    if (d) *d = r.d;
    if (P) *P = r.P;
    if (S) *S = r.S;
    if (nprof) *nprof = r.nprof;
    if (nfield) *nfield = r.nfield;
    if (kind) *kind = r.kind;
    return 0;
}

int oz_rhs_eval(const oz_problem* pb, int64_t nodes, const void* y, const void* p, void* dy, void* prof, void* field) {
    OZ_REQUIRE(pb != nullptr, "null problem descriptor");
    OZ_REQUIRE(known_rhs(pb->rhs_id), "unknown rhs_id %d", pb->rhs_id);
    OZ_REQUIRE(nodes >= 0, "negative node count");
    OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
    EvalArgs ev{nodes, y, p, dy, prof, field, nullptr, nullptr, nullptr, nullptr, nullptr, 1, 0};
    return route(pb, TM_RHS_EVAL, nullptr, &ev);
}

int oz_rhs_out_vjp(const oz_problem* pb, int64_t nodes, const void* y, const void* p, const void* vprof, const void* vfield,
                   void* y_bar, void* p_bar, void* th_bar) {
    OZ_REQUIRE(pb != nullptr, "null problem descriptor");
    OZ_REQUIRE(known_rhs(pb->rhs_id), "unknown rhs_id %d", pb->rhs_id);
    OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
    OZ_REQUIRE(nodes >= 0, "negative node count");
    OZ_REQUIRE(rhs_kind(pb->rhs_id) == 0, "output VJPs are defined for per-trajectory right-hand sides");
    EvalArgs ev{nodes, y, p, nullptr, nullptr, nullptr, vprof, vfield, y_bar, p_bar, th_bar, 1, 0};
    return route(pb, TM_OUT_VJP, nullptr, &ev);
}

int oz_rhs_jac_pattern(int rhs_id, uint8_t* jp, uint8_t* jpp) {
    OZ_REQUIRE(known_rhs(rhs_id), "unknown rhs_id %d", rhs_id);
    OZ_REQUIRE(rhs_kind(rhs_id) == 0, "Jacobian patterns are defined for per-trajectory right-hand sides");
    oz_problem pb;
    memset(&pb, 0, sizeof(pb));
    pb.rhs_id = rhs_id;
    EvalArgs ev{0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, jp, jpp, nullptr, 0, 0};
    return route(&pb, TM_JAC_PATTERN, nullptr, &ev);
}

int oz_rhs_jac_eval(const oz_problem* pb, int64_t nodes, int64_t blocks, int32_t flags, const void* y, const void* p, void* f,
                    void* Jy, void* Jp) {
    OZ_REQUIRE(pb != nullptr, "null problem descriptor");
    OZ_REQUIRE(known_rhs(pb->rhs_id), "unknown rhs_id %d", pb->rhs_id);
    OZ_REQUIRE(pb->dtype == OZ_F64 || pb->dtype == OZ_F32, "unknown dtype %d", pb->dtype);
    OZ_REQUIRE(nodes >= 0 && blocks >= 0, "negative node / block count");
    OZ_REQUIRE(rhs_kind(pb->rhs_id) == 0, "node Jacobians are defined for per-trajectory right-hand sides");
    OZ_REQUIRE(!(flags & OZ_JAC_STAGE_PARAMS) || (pb->num_stages >= 1 && pb->num_stages <= OZ_MAX_STAGES), "stage mode needs the tableau");
    EvalArgs ev{nodes, y, p, f, nullptr, nullptr, nullptr, nullptr, Jy, Jp, nullptr, blocks, flags};
    return route(pb, TM_JAC_EVAL, nullptr, &ev);
}

size_t oz_tm_forward_workspace_bytes(const oz_problem* pb) {
    if (check_problem(pb)) return (size_t)-1;
    if (rhs_kind(pb->rhs_id) == 0) return 0;
    size_t bytes = 0;
    if (route(pb, 100, &bytes, nullptr)) return (size_t)-1;      // a grid right-hand side (built in or plugin): mode 100 = forward scratch
    return bytes;
}

int oz_tm_forward(const oz_problem* pb) {
    if (int rc = check_problem(pb)) return rc;
    OZ_REQUIRE(pb->y0 && pb->h, "y0 and h are required");
    return route(pb, TM_FORWARD, nullptr, nullptr);
}

size_t oz_tm_adjoint_workspace_bytes(const oz_problem* pb) {
    if (check_problem(pb)) return (size_t)-1;
    size_t bytes = 0;
    if (route(pb, TM_ADJOINT_WS, &bytes, nullptr)) return (size_t)-1;
    return bytes;
}

int oz_tm_adjoint(const oz_problem* pb) {
    if (int rc = check_problem(pb)) return rc;
    OZ_REQUIRE(pb->y0 && pb->h, "y0 and h are required");
    return route(pb, TM_ADJOINT, nullptr, nullptr);
}

size_t oz_picard_workspace_bytes(const oz_problem* pb) {
    if (check_problem(pb)) return (size_t)-1;
    size_t bytes = 0;
    if (plugin_of(pb->rhs_id)) return 0;
    if (picard_dispatch(pb, 1, &bytes, nullptr, 0.0, 0)) return (size_t)-1;
    return bytes;
}

int oz_picard_forward(const oz_problem* pb, const void* guess, double tol, int max_sweeps) {
    if (int rc = check_problem(pb)) return rc;
    if (const oz_plugin_info* pl = plugin_of(pb->rhs_id)) {
        OZ_REQUIRE(pb->y0 && pb->h, "y0 and h are required");
        OZ_REQUIRE(tol > 0.0 && max_sweeps >= 0, "tolerance must be positive and max_sweeps non-negative");
        int blk[10] = {0, pl->d};
        int nb = pb->num_state_blocks;
        if (nb <= 0) nb = 1;
        else {
            OZ_REQUIRE(nb <= 8, "at most 8 state blocks");
            for (int b = 0; b <= nb; ++b) blk[b] = pb->state_block_lo[b];
        }
        return plugin_rc(pl, pl->picard(pb, guess, tol, max_sweeps, blk, nb));
    }
    return picard_dispatch(pb, 0, nullptr, guess, tol, max_sweeps);
}

int oz_colloc_residual(const oz_problem* pb, const void* Y, const void* y, void* c1, void* c2) {
    if (int rc = check_problem(pb)) return rc;
    if (const oz_plugin_info* pl = plugin_of(pb->rhs_id)) return plugin_rc(pl, pl->colloc(pb, 0, nullptr, Y, y, nullptr, nullptr, c1, c2));
    return colloc_dispatch(pb, 0, nullptr, Y, y, nullptr, nullptr, c1, c2);
}

size_t oz_colloc_vjp_workspace_bytes(const oz_problem* pb) {
    if (check_problem(pb)) return (size_t)-1;
    size_t bytes = 0;
    if (const oz_plugin_info* pl = plugin_of(pb->rhs_id)) {
        if (pl->colloc(pb, 2, &bytes, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr)) return (size_t)-1;
        return bytes;
    }
    if (colloc_dispatch(pb, 2, &bytes, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr)) return (size_t)-1;
    return bytes;
}

int oz_colloc_vjp(const oz_problem* pb, const void* Y, const void* y, const void* mu1, const void* mu2, void* Y_bar,
                  void* y_bar) {
    if (int rc = check_problem(pb)) return rc;
    if (const oz_plugin_info* pl = plugin_of(pb->rhs_id)) return plugin_rc(pl, pl->colloc(pb, 1, nullptr, Y, y, mu1, mu2, Y_bar, y_bar));
    return colloc_dispatch(pb, 1, nullptr, Y, y, mu1, mu2, Y_bar, y_bar);
}

}  // extern "C"
