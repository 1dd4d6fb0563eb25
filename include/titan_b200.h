/*
 * titan_b200.h -- C ABI of the B200-native mass-spring step engine (libtitan_b200.so).
 *
 * The reference (GitWyd/Titan) has no FFI layer: its boundary is the C++ class titan::Simulation
 * (include/Titan/sim.h:38-129).  This ABI is the "thin C-ABI layer" the host C++ facade
 * (the headers in include/Titan + titan_b200/host/) calls; each entry point names the reference code it replaces
 * (paths relative to the reference root).  Plain pointers and sizes only; no C++ or torch types.
 *
 * Conventions
 *  - every function returns 0 on success, else a tb_status / cudaError_t-derived code; the message is in
 *    tb_last_error(e).  Nothing throws across this boundary.
 *  - pointers are HOST pointers to caller-owned plain arrays unless the name ends in _dev.
 *  - xyz arrays are [count][3] row-major doubles (the layout of titan::Vec, include/Titan/vec.h:165).
 *  - optional array arguments may be NULL (= leave that field unchanged; at first upload = default).
 *  - one engine owns one CUDA device + one stream; calls on one engine are not re-entrant.  The facade's
 *    GPU thread calls tb_run_async/tb_sync, the user thread calls upload/download only while paused,
 *    exactly the discipline of the reference (docs/README.md:440-443).
 *  - there is NO CPU fallback: without a CUDA device tb_create fails with TB_ERR_NO_DEVICE.
 */
#ifndef TITAN_B200_H
#define TITAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB_ABI_VERSION 1

#if defined(__GNUC__)
#define TB_API __attribute__((visibility("default")))
#else
#define TB_API
#endif

typedef struct tb_engine tb_engine;

typedef enum {
    TB_OK = 0,
    TB_ERR_INVALID = 1,     /* bad argument / call order */
    TB_ERR_NO_DEVICE = 2,   /* no CUDA device: there is no CPU path */
    TB_ERR_CUDA = 3,        /* a CUDA call failed; see tb_last_error */
    TB_ERR_UNSUPPORTED = 4, /* feature outside the built scope (e.g. ghost masses on the generic kernel) */
    TB_ERR_OOM = 5
} tb_status;

/* tb_create flags */
#define TB_MATH_FAST 0u      /* default: fp64 with FMA contraction and one reciprocal per spring;  */
                             /* <= 1e-12 relative per step against the oracle (tests state it)      */
#define TB_MATH_STRICT 1u    /* unfused IEEE ops in the reference's operation order: bit-identical  */
                             /* to oracle/titan_oracle.c (passive springs; sin() may differ by 1ulp) */
#define TB_NO_GRAPH 2u       /* launch every step individually instead of CUDA-Graph batches         */
#define TB_KERNEL_CSR 4u     /* force the generic CSR-gather step kernel (disable the tiled one)     */
#define TB_NO_RESIDENT 16u   /* never use the resident small-system kernel (<= 8 tiles: one cluster launch runs all steps   */
                             /* of a tb_run_async with the state in shared memory); stream every step instead           */
#define TB_INTEGRATOR_VERLET 8u /* the reference's -DVERLET build (CMakeLists.txt:13,90-93; src/sim.cu:1350-1354):    */
                             /* vel += 0.5*(acc + f/m)*dt; acc = f/m; pos += vel*dt + 0.5*acc*dt^2, no speed clamp; */
                             /* acc is persistent state (zero at upload, like CUDA_MASS(Mass&)); both step kernels    */

#define TB_INTEGRATOR_RK2 32u /* the reference's -DRK2 build (CMakeLists.txt:14,95-100; src/sim.cu:1336-1349,1778-1799) with its two call sites  */
                             /* repaired to pass the arguments the kernel declares (as shipped they do not compile): midpoint rule,      */
                             /* two force evaluations (two kernel launches) per step, T += 0.5 dt per stage, no speed clamp; ACTUATED_*  */
                             /* springs advance their rest length in both evaluations, as computeSpringForces does.  Streaming and      */
                             /* generic kernels; not on a partition, not with magnetic masses (fails loudly)                             */

/* SpringType, include/Titan/spring.h:17-18 */
enum { TB_PASSIVE_SOFT = 0, TB_PASSIVE_STIFF = 1, TB_ACTIVE_CONTRACT_THEN_EXPAND = 2,
       TB_ACTIVE_EXPAND_THEN_CONTRACT = 3, TB_ACTUATED_EXPAND = 4, TB_ACTUATED_CONTRACT = 5 };
/* CONSTRAINT_TYPE, include/Titan/object.h:225-227 */
enum { TB_CONSTRAINT_PLANE = 0, TB_CONTACT_PLANE = 1, TB_BALL = 2, TB_DIRECTION = 3 };

/* CudaContactPlane (include/Titan/object.h:138-149); normal already normalised by the caller, as the
 * ContactPlane constructor does (object.h:101-109). */
typedef struct { double normal[3]; double offset; double friction_k; double friction_s; } tb_plane;
/* CudaBall (include/Titan/object.h:90-99) */
typedef struct { double center[3]; double radius; } tb_ball;
/* one record of a mass's LOCAL_CONSTRAINTS (include/Titan/object.h:181-220; Mass::addConstraint,
 * src/mass.cu:104-122): vec = normalised normal / tangent, or ball center; num = friction / offset / radius */
typedef struct { int32_t kind; int32_t pad; double vec[3]; double num; } tb_local;

/* ---- lifetime ------------------------------------------------------------------------------------------- */
TB_API int tb_abi_version(void);
TB_API int tb_device_count(void);
/* replaces the implicit device-0 context of the reference (no cudaSetDevice anywhere in src/sim.cu) */
TB_API int tb_create(int device, uint32_t flags, tb_engine **out);
TB_API void tb_destroy(tb_engine *e);                 /* Simulation::freeGPU, src/sim.cu:131-182 */
TB_API const char *tb_last_error(const tb_engine *e); /* replaces gpuErrchk's fprintf, src/sim.cu:60-75 */

/* ---- topology: replaces massToArray/springToArray pointer tables, src/sim.cu:940-1025 ------------------ */
/* Declares nm masses and ns springs; left/right are mass indices, -1 = nullptr endpoint (spring skipped,
 * src/sim.cu:1163).  The engine derives the CSR incidence (ascending spring index per mass) on upload.
 * State and parameters are reset to defaults (m=0.1, k=10000, valid, not fixed, PASSIVE_SOFT, ...). */
TB_API int tb_set_topology(tb_engine *e, int64_t nm, int64_t ns, const int32_t *left, const int32_t *right);
/* Optional: tile id per mass (nm entries; NULL clears).  Masses sharing an id are kept together in one tile
 * of the step kernel (split when more than 128 share one); ids order the tiles.  Builders that know the geometry
 * (Lattice::Lattice, src/object.cu:235-296 -> bricks of lattice cells) pass it so that congruent bricks share
 * their incidence tables; without a hint the engine bisects the current positions.  Results do not depend on it
 * beyond the summation order of TB_MATH_FAST (TB_MATH_STRICT is bit-identical for any tiling). */
TB_API int tb_set_tile_hint(tb_engine *e, const int32_t *tile_of_mass);

/* ---- mass state: CUDA_MASS(Mass&) src/mass.cu:74-100, Simulation::set/setAll src/sim.cu:603-765 -------- */
/* idx == NULL: masses [first, first+count); otherwise the `count` masses listed in idx (container upload). */
TB_API int tb_set_mass_state(tb_engine *e, int64_t first, int64_t count, const int32_t *idx,
                      const double *pos, const double *vel, const double *m, const double *T,
                      const double *extern_force, const uint8_t *fixed, const int32_t *drag,
                      const uint8_t *valid);
/* rad, stiffness, max_mag_force, mag_scale_factor (include/Titan/mass.h:97-100).  Once any mass has a non-zero rad
 * or max_mag_force every step runs the occupancy-grid + magnet pass (recomputeOG src/sim.cu:893-906,
 * computeExternalMagnetForcesOG :1250-1281) in front of the step kernel (tiled or generic); until then the pass is elided
 * (it would add exactly +0).  Not available together with ghost masses.  Pair forces are summed in (cell, ascending
 * mass index) order from start-of-step positions. */
TB_API int tb_set_mass_magnet(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, const double *rad,
                       const double *stiffness, const double *max_mag_force, const double *mag_scale);
/* Mass::operator=(CUDA_MASS&) src/mass.cu:21-49, Simulation::get/getAll src/sim.cu:589-654,1051-1116 */
TB_API int tb_get_mass_state(tb_engine *e, int64_t first, int64_t count, const int32_t *idx,
                      double *pos, double *vel, double *acc, double *T, double *extern_force);

/* ---- spring parameters: CUDA_SPRING(const Spring&) src/spring.cu:37-63 --------------------------------- */
TB_API int tb_set_spring_params(tb_engine *e, int64_t first, int64_t count, const int32_t *idx,
                         const double *k, const double *rest, const double *damping, const int32_t *type,
                         const double *omega, const double *l_max, const double *l_min, const double *rate);
/* Simulation::deleteSpring after start() (src/sim.cu:382-412) without a topology rebuild: the listed springs become
 * tombstones -- they keep their index but exert no force from now on (k and damping forced to zero, also after later
 * tb_set_spring_params calls), which is bit-identical to their absence.  Compaction: tb_compact.
 * Used by the topology-optimisation workload (BASELINE config C5: remove the weakest springs every 100 steps). */
TB_API int tb_remove_springs(tb_engine *e, int64_t n, const int32_t *idx);
/* Periodic compaction (thrust::remove in the reference's delete paths, src/sim.cu:371,401,519-552): rebuilds the device layout
 * at the next run without the incidence words of removed springs.  Mass and spring indices do not change.  Removals of up to a
 * few thousand springs are compacted on the spot, tile by tile; large scattered removals (topology optimisation) are better
 * left in place: their dead words cost ~0.1 % of a step, while dropping them makes every tile's incidence block unique and
 * defeats the content dedupe that keeps a lattice's blocks L2-resident (measured on C5: 168 -> 151 G spring-steps/s). */
TB_API int tb_compact(tb_engine *e);
/* Simulation::createMass / createSpring after start() (src/sim.cu:231-253,300-322), without re-uploading the system: the arrays
 * grow by n entries with the defaults of tb_set_topology (the new masses are [old nm, old nm + n), the new springs
 * [old ns, old ns + n): fill them with tb_set_mass_state / tb_set_spring_params), and at the next run only the tiles the new
 * springs touch are rebuilt; new masses get tiles of their own.  Results equal a fresh upload of the enlarged system (the
 * per-mass summation order is by spring index whatever the tiling).  TB_ERR_UNSUPPORTED where only a re-upload will do
 * (engines of a partition; tb_add_masses with magnet fields or per-mass constraints on the device).
 * Simulation::deleteMass after start (:343-380) is tb_set_mass_state(valid = 0): the mass freezes and its springs become
 * tombstones, again without a rebuild. */
TB_API int tb_add_masses(tb_engine *e, int64_t n);
TB_API int tb_add_springs(tb_engine *e, int64_t n, const int32_t *left, const int32_t *right);
/* Spring::update copies back only _rest (src/spring.cu:12-14) */
TB_API int tb_get_spring_rest(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, double *rest);

/* ---- constraints ---------------------------------------------------------------------------------------- */
/* createPlane/createBall, src/sim.cu:2251-2288; applied in array order, planes before balls (:1303-1309) */
TB_API int tb_set_planes(tb_engine *e, int32_t n, const tb_plane *planes);
TB_API int tb_set_balls(tb_engine *e, int32_t n, const tb_ball *balls);
/* per-mass local constraints as CSR over all nm masses (row_ptr[nm+1]); NULL/empty clears them.
 * Evaluation order per mass is the reference's: contact planes, balls, constraint planes, directions,
 * each kind in record order (src/sim.cu:1312-1326). */
TB_API int tb_set_local_constraints(tb_engine *e, const int64_t *row_ptr, const tb_local *records);

/* setGlobalAcceleration src/sim.cu:2334-2340, setTimeStep :798-808 */
TB_API int tb_set_globals(tb_engine *e, const double g[3], double dt);

/* ---- stepping: Simulation::execute, src/sim.cu:1704-1841 (Euler branch :1800-1811) -------------------- */
/* Enqueue nsteps steps on the engine stream and return.  t0 is the host Simulation::T at the first of
 * them; step j uses t = t0 (+ dt) j times, accumulated in fp64 exactly like `T += dt` (:1811).
 * nsteps == 0 validates the scene and brings the device layout up to date (tiles, records) without launching. */
TB_API int tb_run_async(tb_engine *e, int64_t nsteps, double t0);
TB_API int tb_sync(tb_engine *e);                                  /* cudaDeviceSynchronize at a breakpoint, :1708 */
TB_API int tb_steps_done(tb_engine *e, int64_t *steps);            /* total steps enqueued so far */
/* Blocks until the device has completed the first `steps` steps ever enqueued (<= tb_steps_done), later ones may still
 * be running: lets the run loop keep the next batch queued behind the one it waits for, where the reference drains the
 * device after every single step (cudaDeviceSynchronize, src/sim.cu:1771-1776). */
TB_API int tb_wait_steps(tb_engine *e, int64_t steps);
/* device time in ms of the most recent tb_run_async (CUDA events on the engine stream); call after tb_sync */
TB_API int tb_last_run_ms(tb_engine *e, double *ms);

/* ---- partitioned simulation: boundary-mass halo exchange over NVLink peer memory ---------------------------
 * One large system is split by mass over several engines (one per GPU / process).  Each engine holds
 * its OWN masses plus GHOST copies of the foreign endpoints of its cut springs; every cut spring is stored (and
 * evaluated) by each engine owning one of its endpoints, so no forces are exchanged and summation order is kept.
 * The reference has no such layer (single GPU, src/sim.cu:1704-1841); this replaces its single-device loop for
 * systems beyond one GPU.
 *
 * The exchange is fused into the step kernel: tiles whose springs reach ghosts are processed first; the thread that
 * updates an exported mass also stores its new state straight into the owning slot of the peer's state buffer
 * (P2P store over NVLink), and when the last such tile of a step is done a step counter is released into each
 * peer's flag word.  A peer's next step acquires that counter before it reads its ghost slots -- no host
 * involvement, no separate pack/copy/unpack kernels, capturable in the CUDA graphs of tb_run_async.
 * All engines of a partition must be advanced by the same tb_run_async calls (same step counts).
 *
 * Call order: tb_set_topology / state / parameters (local numbering) -> tb_halo_set_ghosts -> tb_halo_prepare ->
 * [exchange handles and ghost offsets between engines] -> tb_halo_connect -> tb_halo_set_exports -> tb_run_async. */
#define TB_MAX_PEERS 8
typedef struct { unsigned char bytes[320]; } tb_halo_handle;   /* opaque; valid across processes on one node */
/* Marks `n` local masses as ghosts: never updated or written by this engine; a peer supplies their state. */
TB_API int tb_halo_set_ghosts(tb_engine *e, int64_t n, const int32_t *mass_idx);
/* Builds the device layout now (it is otherwise built lazily by the first tb_run_async) and returns the handle a
 * peer needs to reach this engine's state buffers and flag words. */
TB_API int tb_halo_prepare(tb_engine *e, tb_halo_handle *out);
/* For local masses (normally ghosts): the offset, in this engine's state buffers, a peer must store their state at
 * (position x at offset, the other five components at a fixed stride). */
TB_API int tb_halo_state_offsets(tb_engine *e, int64_t n, const int32_t *mass_idx, int64_t *offsets);
/* Opens the peers' buffers.  flag_slot_at_peer[j] = index of THIS engine in peer j's peer list (which of peer j's
 * flag words this engine signals).  Handles from the same process are used directly, others through CUDA IPC. */
TB_API int tb_halo_connect(tb_engine *e, int32_t n_peers, const tb_halo_handle *peers, const int32_t *flag_slot_at_peer);
/* After every step the new state of local mass mass_idx[j] is stored at remote_offset[j] of peer peer_idx[j]. */
TB_API int tb_halo_set_exports(tb_engine *e, int64_t n, const int32_t *mass_idx, const int32_t *peer_idx,
                               const int64_t *remote_offset);
/* Pushes the CURRENT state of every exported mass to the peers (after an upload while paused) and returns when the
 * stores are visible; the caller then synchronises the engines (barrier) before the next tb_run_async.  Once
 * connected, a ghost's state is whatever its owner last pushed: uploads to ghost masses are ignored. */
TB_API int tb_halo_push_now(tb_engine *e);

/* ---- several GPUs behind one handle (tb_create_partitioned, tb_batch_create) -------------------------
 * The reference runs one Simulation on one device (no cudaSetDevice anywhere in src/sim.cu) and cannot run two at once
 * (static flags, include/Titan/sim.h:150-154).  A tb_group owns one engine per listed device IN ONE PROCESS and presents
 * them as one system in GLOBAL mass / spring numbering:
 *   tb_create_partitioned  one large system cut by an owner map (engine index per mass; NULL = equal contiguous index
 *                          ranges, i.e. i-slabs of a reference-ordered lattice).  Each engine holds its own masses, ghost
 *                          copies of the foreign endpoints of its cut springs and every spring with an own endpoint (global
 *                          order, so sums are ordered as on one engine: bit-identical under TB_MATH_STRICT); the boundary
 *                          state is exchanged inside the step kernel over peer memory (tb_halo_*, wired by the group).
 *   tb_batch_create        independent simulations (the RL batch, BASELINE config C3) laid end to end in the global
 *                          numbering; whole simulations go to engines in contiguous blocks balanced by mass count: no
 *                          spring is cut, nothing is exchanged.
 * The tb_group_* calls mirror their tb_* namesakes with global indices: mass uploads go to the owner (ghost copies follow
 * by a push before the next run), spring parameters to every engine holding the spring, downloads come from the owner.
 * One host thread drives the group: launches are asynchronous and the engines meet inside their kernels. */
typedef struct tb_group tb_group;
TB_API int tb_create_partitioned(int32_t ndev, const int32_t *devices, uint32_t flags, tb_group **out);
TB_API int tb_batch_create(int32_t ndev, const int32_t *devices, uint32_t flags, int64_t nsims, const int64_t *masses_per_sim, tb_group **out);
TB_API void tb_group_destroy(tb_group *g);
TB_API const char *tb_group_last_error(const tb_group *g);
TB_API int32_t tb_group_size(const tb_group *g);
TB_API tb_engine *tb_group_engine(tb_group *g, int32_t r);            /* for tb_get_info and the like; owned by the group */
TB_API int tb_group_set_topology(tb_group *g, int64_t nm, int64_t ns, const int32_t *left, const int32_t *right, const int32_t *owner);
TB_API int tb_group_set_tile_hint(tb_group *g, const int32_t *tile_of_mass);
TB_API int tb_group_set_mass_state(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *pos, const double *vel,
                                   const double *m, const double *T, const double *extern_force, const uint8_t *fixed,
                                   const int32_t *drag, const uint8_t *valid);
TB_API int tb_group_set_mass_magnet(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *rad,
                                    const double *stiffness, const double *max_mag_force, const double *mag_scale);
TB_API int tb_group_get_mass_state(tb_group *g, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel,
                                   double *acc, double *T, double *extern_force);
TB_API int tb_group_set_spring_params(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *k,
                                      const double *rest, const double *damping, const int32_t *type, const double *omega,
                                      const double *l_max, const double *l_min, const double *rate);
TB_API int tb_group_get_spring_rest(tb_group *g, int64_t first, int64_t count, const int32_t *idx, double *rest);
TB_API int tb_group_remove_springs(tb_group *g, int64_t n, const int32_t *idx);
TB_API int tb_group_set_planes(tb_group *g, int32_t n, const tb_plane *planes);
TB_API int tb_group_set_balls(tb_group *g, int32_t n, const tb_ball *balls);
TB_API int tb_group_set_globals(tb_group *g, const double gravity[3], double dt);
TB_API int tb_group_run_async(tb_group *g, int64_t nsteps, double t0);   /* every engine, same step count */
TB_API int tb_group_sync(tb_group *g);
TB_API int tb_group_wait_steps(tb_group *g, int64_t steps);
TB_API int tb_group_last_run_ms(tb_group *g, double *ms);                /* max over the engines, after tb_group_sync */
TB_API int tb_group_steps_done(tb_group *g, int64_t *steps);

/* ---- introspection used by tests / bench ---------------------------------------------------------------- */
typedef struct {
    int64_t nm, ns;
    int32_t kernel;              /* 0 = CSR gather, 1 = tiled */
    int32_t n_tiles;
    int64_t launches;            /* kernel launches (graph nodes included) since creation */
    int64_t device_bytes;        /* bytes of device memory held */
    double algorithmic_bytes_per_step; /* nm*104 + ns*B_s */
    int32_t sm_count;
    int32_t graph_steps;         /* steps per CUDA-Graph launch (0 = no graph) */
    int32_t threads_per_mass;    /* tiled kernel: threads sharing one mass's incident springs */
    int32_t grid;                /* tiled kernel: persistent CTAs of the last launch */
    int32_t smem_bytes;          /* tiled kernel: dynamic shared memory per CTA */
    int32_t ell_unique_tiles;    /* tiled kernel: distinct incidence (ELL) blocks after content dedupe */
    int64_t tiled_bytes_per_step;/* tiled kernel: bytes one step streams from its tables (blob+state+halo+ELL+meta) */
    int64_t ns_removed;          /* springs turned into tombstones by tb_remove_springs */
    int32_t dict_records;        /* tiled kernel, dictionary mode: most distinct (k, rest, ...) records in a tile; 0 = one
                                    record per spring (parameters too diverse, strict arithmetic or TB_DICT=0) */
    int32_t dict_split_k;        /* 1 = split dictionary: (rest, damping, omega) records shared, stiffness one value per
                                    record slot (systems whose k is edited spring by spring, BASELINE config C5) */
    int32_t layout_uploads;      /* tiled kernel: times the tile tables were allocated and uploaded whole (first build, record-mode
                                    change, an edit that outgrew the room left in them) */
    int32_t layout_patches;      /* tiled kernel: topology edits written into the room left in the tables, nothing re-allocated */
} tb_info;
TB_API int tb_get_info(tb_engine *e, tb_info *info);
/* headless position export, fp64 -> packed fp32 xyz (updateVertices, src/sim.cu:1944-1952); out is a
 * DEVICE pointer to nm*3 floats owned by the engine, valid until the next call. */
TB_API int tb_export_positions_f32_dev(tb_engine *e, const float **out_dev);
/* the matching line list (updateIndices, src/sim.cu:1957-1980): 2*ns uint32 (left, right) mass indices in spring order;
 * springs the step skips (null or invalid endpoint, tombstones) are the degenerate line (0, 0).  DEVICE pointer owned by
 * the engine, valid until the next call. */
TB_API int tb_export_indices_dev(tb_engine *e, const uint32_t **out_dev);
/* copy `bytes` from a device pointer handed out by this engine (the two exports above) into host memory: what a headless
 * consumer uses in place of the reference's OpenGL buffer mapping (src/sim.cu:1982-2043) */
TB_API int tb_read_device(tb_engine *e, void *dst_host, const void *src_dev, size_t bytes);

/* ---- download that allocates nothing ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as tb_get_mass_state for
 * pos and/or vel (NULL = skip), bit for bit, but the gather kernel's scratch is a file-scope `static __device__` array
 * inside the library instead of the engine's growable staging buffer: the call makes no cudaMalloc / cudaFree.  At most
 * tb_static_scratch_max_masses() (65,536) masses per call; more -> TB_ERR_INVALID with a message saying so (download in
 * windows, or use tb_get_mass_state).  The array costs 1.75 MiB on every GPU that loads the library, until the process
 * exits.  All calls on one device share it and it keeps the previous call's contents (nothing needs clearing: a call
 * reads back only what it wrote); calls that overlap on the GPU -- engines of the same device driven from different
 * threads -- race, so serialise them. */
TB_API int64_t tb_static_scratch_max_masses(void);
TB_API int tb_get_mass_state_static(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel);

/* page-locked host memory for upload/download buffers (the reference stages through pageable `new[]`
 * arrays, src/sim.cu:948-962,1099-1101); owned by the engine, released by tb_destroy */
TB_API int tb_alloc_host(tb_engine *e, size_t bytes, void **out);

#ifdef __cplusplus
}
#endif
#endif /* TITAN_B200_H */
