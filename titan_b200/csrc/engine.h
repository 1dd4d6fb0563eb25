// engine.h -- internal state of one tb_engine (one CUDA device, one stream).
//
// Data layout in HBM ("canonical" state, API mass/spring order), all fp64 structure-of-arrays:
//   pos[2], vel[2]   : ping-pong buffers, each 3 planes of nm doubles (x[], y[], z[])
//   acc              : 3 planes, written by the last step of every run
//   ext              : 3 planes, allocated only once a non-zero external force is uploaded
//   m, flags, drag   : per-mass scalars (flags: bit0 fixed, bit1 valid, bit2 has local constraints)
//   left,right,k,rest,damping,type,omega,l_max,l_min,rate : per-spring arrays
//   row_ptr, adj     : CSR incidence; adj = 2*spring + (1 if this mass is the spring's right endpoint),
//                      rows in ascending spring order, right entry before left entry for self-loops
// This replaces the reference's one-cudaMalloc-per-object AoS + pointer tables
// (include/Titan/mass.h:89-126, spring.h:77-97, src/sim.cu:940-1025).
#pragma once

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <unordered_map>
#include <vector>

#include <cuda_runtime.h>

#include "physics.cuh"

namespace tb {

enum : uint8_t { F_FIXED = 1, F_VALID = 2, F_LOCAL = 4, F_GHOST = 8 };   // ghost: state supplied by a peer engine (halo exchange)
// what a setter invalidated in the tiled layout: structure (tiles/ELL), spring parameters, mass state, mass flags
enum : uint32_t { TS_STRUCT = 1, TS_PARAMS = 2, TS_STATE = 4, TS_FLAGS = 8, TS_PATCH = 16 };   // PATCH: masses / springs were appended

struct DtSegment { int64_t first_step; double dt; };

struct TiledState;  // tiled.cu

struct Engine {
    int device = 0;
    uint32_t flags = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // tb_wait_steps: one event per tb_run_async in a small ring, tagged with the step count reached when it completes
    static constexpr int RUN_RING = 8;
    cudaEvent_t run_ev[RUN_RING] = {};
    int64_t run_ev_steps[RUN_RING] = {};
    int64_t run_count = 0;
    std::string err;
    int sm_count = 0;

    int64_t nm = 0, ns = 0;
    // device state
    double *pos[2] = {nullptr, nullptr}, *vel[2] = {nullptr, nullptr};
    double *acc = nullptr, *ext = nullptr, *m = nullptr;
    uint8_t *mflags = nullptr;
    int32_t *drag = nullptr;
    int32_t *left = nullptr, *right = nullptr, *type = nullptr;
    double *k = nullptr, *rest = nullptr, *damping = nullptr, *omega = nullptr, *l_max = nullptr,
           *l_min = nullptr, *rate = nullptr;
    int32_t *row_ptr = nullptr, *adj = nullptr;
    // ACTUATED_EXPAND / ACTUATED_CONTRACT springs (RobotLink), ascending spring index: the per-step _rest mutation
    // (src/sim.cu:1174-1181) runs over this list only
    int32_t *act_idx = nullptr; int64_t n_act = 0; bool act_dirty = true;
    std::vector<int32_t> h_act;
    int32_t *lc_ptr = nullptr;
    tb_local *lc = nullptr;
    tb_plane *more_planes = nullptr;
    tb_ball *more_balls = nullptr;
    double *rk2_bvel = nullptr;    // [3][nm] RK2 + local constraints: velocity backups of stage 1 (MassExtras::rk2_bvel)
    double *t_dev = nullptr;       // [2] simulation time, ping-pong with the step parity
    long long *step_dev = nullptr; // [2] steps completed, ping-pong (halo exchange sequence number)
    float *export_f32 = nullptr;
    uint32_t *export_idx = nullptr;   // [2 * ns] line list of the last tb_export_indices_dev
    // magnet / occupancy-grid pass (magnet.cu), allocated with the first magnetic mass: per-mass rad, shell
    // stiffness, max_mag_force, mag_scale_factor (include/Titan/mass.h:97-100), the pass's output (uploaded one-shot
    // external force + pair forces, [3][nm]) and the cell lists
    double *mag_rad = nullptr, *mag_stiff = nullptr, *mag_max = nullptr, *mag_scale = nullptr, *mag_force = nullptr;
    bool stiff_seen = false, scale_seen = false;         // shell stiffness / mag_scale uploaded so far are uniform:
    double stiff_uniform = 0.0, scale_uniform = 0.0;     //   remembered here instead of stored per mass
    int32_t *og_cell = nullptr, *og_items = nullptr, *og_sorted = nullptr, *og_counter = nullptr;
    double *og_pos = nullptr;        // positions in og_sorted order, [3][nm]
    float4 *og_posf = nullptr;       // (float x, y, z, mass index) in og_sorted order
    int cur = 0;                   // which pos/vel buffer holds the current state

    // host-side knowledge
    Globals G{};
    std::vector<tb_plane> planes;
    std::vector<tb_ball> balls;
    std::vector<uint8_t> h_flags;  // mirror of mflags
    std::vector<uint8_t> h_type;   // mirror of the spring types
    std::vector<uint8_t> h_removed; // tombstones (tb_remove_springs); empty until the first removal
    uint8_t *removed_dev = nullptr;
    int64_t n_removed = 0;
    std::vector<double> h_omega;   // mirror of the actuation frequencies (distinct-omega sine table of the tiled kernel)
    std::vector<int32_t> h_left, h_right, h_row_ptr, h_adj;  // host copy of the topology + CSR (tile builder)
    // the CSR covers masses [0, csr_nm) and springs [0, csr_ns); incidences of springs added since (tb_add_springs) wait in
    // extra_adj (mass -> 2*spring+side, ascending) until ensure_csr folds them in
    int64_t csr_nm = 0, csr_ns = 0, row_ptr_cap = 0, adj_cap = 0;
    // elements allocated per per-spring array / per plane of a per-mass array (0 = exactly ns / nm): tb_add_springs and
    // tb_add_masses leave room, so that the next edit re-allocates nothing.  [3][nm] arrays keep their plane stride nm inside
    // a buffer of 3 * cap_m elements; a changed nm re-strides them through m_scratch (buffers swap roles, nothing is allocated)
    int64_t cap_s = 0, cap_m = 0;
    double *m_scratch = nullptr;
    int64_t s_alloc() const { return cap_s ? cap_s : ns; }
    int64_t m_alloc() const { return cap_m ? cap_m : nm; }
    std::vector<int32_t> patch_masses, patch_springs;        // added / removed since the tiled layout was last brought up to date (TS_PATCH)
    bool patch_needs_layout = false;      // ... and some of them are additions: if the tiles cannot be patched, everything is rebuilt
    bool compact_requested = false;       // tb_compact: the next full build leaves the incidence words of removed springs out
    bool removed_unpatched = false;       // tombstones whose incidence words are still in their tiles (they read k = 0 from their record)
    std::unordered_map<int32_t, std::vector<int32_t>> extra_adj;
    std::vector<double> h_T0;      // per-mass T at its last upload
    std::vector<int64_t> h_n0;     // steps_done at that upload
    std::vector<DtSegment> dt_log;
    int64_t steps_done = 0;
    bool ext_pending = false;      // a non-zero extern force waits for the next step
    bool any_damping = false, any_active = false, any_actuated = false, any_local = false, any_magnet = false;
    bool any_fixed_or_invalid = false, any_ghost = false;
    double last_run_ms = 0.0;
    int64_t launches = 0;
    int64_t device_bytes = 0;

    // CUDA-graph cache for multi-step batches (one per parity)
    cudaGraphExec_t graph[2] = {nullptr, nullptr};
    int graph_steps = 0;
    bool graph_dirty = true;

    // staging
    void *stage_dev = nullptr; size_t stage_dev_bytes = 0;
    void *stage_host = nullptr; size_t stage_host_bytes = 0;   // pinned

    std::vector<void *> pinned;    // tb_alloc_host allocations

    TiledState *tiled = nullptr;
    uint32_t tiled_stale = 0xf;    // TS_* bits: which parts of the tiled layout must be refreshed
    bool canon_stale = false;      // canonical pos/vel must be refreshed from the tiled layout
    std::vector<int32_t> tile_hint; // optional caller-supplied tile id per mass (tb_set_tile_hint)
};

// error plumbing: store message, return status
int fail(Engine *e, int status, const char *fmt, ...);
#define TB_CUDA(e, call)                                                                         \
    do {                                                                                         \
        cudaError_t _err = (call);                                                               \
        if (_err != cudaSuccess)                                                                 \
            return tb::fail((e), TB_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_err), \
                            __FILE__, __LINE__);                                                 \
    } while (0)

// TB_TIMING=1: wall time of the host-side phases (edits, layout builds), one line each on stderr
struct TimeScope {
    const char *what; std::chrono::steady_clock::time_point t0; bool on;
    explicit TimeScope(const char *w) : what(w), on(timing_on()) { if (on) t0 = std::chrono::steady_clock::now(); }
    ~TimeScope() { if (on) fprintf(stderr, "[tb timing] %s: %.2f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count()); }
    static bool timing_on() { static const bool v = getenv("TB_TIMING") != nullptr; return v; }
};

#ifndef TB_TRY
#define TB_TRY(x) do { int _rc = (x); if (_rc != TB_OK) return _rc; } while (0)
#endif

template <class T> int dev_alloc(Engine *e, T **p, size_t count);
void dev_free(Engine *e, void *p, size_t bytes);

// engine.cu: host + device CSR brought up to date with the current topology (after tb_add_masses / tb_add_springs)
int ensure_csr(Engine *e);
// engine.cu: (re)builds e->act_idx / e->h_act, the list of ACTUATED_* springs; allocates, so never under stream capture
int prepare_actuated(Engine *e);
// csr_step.cu
int launch_csr_step(Engine *e, bool use_ext, bool write_acc, int rk2_stage = 0);
// magnet.cu
int magnet_alloc(Engine *e);
void magnet_free(Engine *e);
int launch_magnet_pass(Engine *e, bool use_ext, bool tiled);   // occupancy grid + pair forces -> e->mag_force
// Where the start-of-step positions live: canonical planes [3][nm] (generic kernel), or the tile-blocked state of the
// tiled kernel (blocks of 6 x 128 doubles: x y z vx vy vz of the tile's 128 slots) addressed through api_to_slot.
struct PosView {
    const double *base = nullptr; const int32_t *api_to_slot = nullptr; int64_t nm = 0;
#ifdef __CUDACC__
    __device__ __forceinline__ double ld(int64_t i, int c) const {
        if (!api_to_slot) return base[c * nm + i];
        const int32_t s = api_to_slot[i];
        return base[(int64_t)(s >> 7) * 768 + c * 128 + (s & 127)];
    }
#endif
};
// tiled.cu
int tiled_prepare(Engine *e);                    // (re)build tiles if stale
int launch_tiled_step(Engine *e, bool use_ext, bool write_acc, int rk2_stage = 0);
int tiled_to_canonical(Engine *e);               // refresh canonical pos/vel from tiles
void tiled_destroy(Engine *e);
bool tiled_supported(const Engine *e);
int tiled_info(const Engine *e, tb_info *info);
bool tiled_halo_connected(const Engine *e);
bool tiled_resident_ok(const Engine *e);         // <= 8 tiles, no peers: all steps of a run in one cluster launch
int launch_tiled_resident(Engine *e, int nsteps, bool write_acc);
void tiled_pos_view(const Engine *e, PosView *v); // the current state buffer of the tiled kernel
int tiled_halo_check(Engine *e);                 // after a sync: did a halo wait time out?

}  // namespace tb

struct tb_engine : tb::Engine {};
