// engine.cu -- C-ABI entry points of libtitan_b200.so (include/titan_b200.h): device memory, SoA
// pack/unpack at the host<->device boundary, constraint tables, and the run loop that batches steps
// into CUDA graphs.  The step kernels themselves live in csr_step.cu and tiled.cu.
#include "engine.h"

#include <mutex>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>

namespace tb {

int fail(Engine *e, int status, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (e) e->err = buf;
    return status;
}

template <class T> int dev_alloc(Engine *e, T **p, size_t count) {
    *p = nullptr;
    if (count == 0) count = 1;
    cudaError_t err = cudaMalloc((void **)p, count * sizeof(T));
    if (err != cudaSuccess) {
        cudaGetLastError();
        return fail(e, TB_ERR_OOM, "cudaMalloc of %zu bytes failed: %s", count * sizeof(T), cudaGetErrorString(err));
    }
    e->device_bytes += (int64_t)(count * sizeof(T));
    return TB_OK;
}
template int dev_alloc<double>(Engine *, double **, size_t);
template int dev_alloc<int32_t>(Engine *, int32_t **, size_t);
template int dev_alloc<uint8_t>(Engine *, uint8_t **, size_t);
template int dev_alloc<float>(Engine *, float **, size_t);
template int dev_alloc<uint32_t>(Engine *, uint32_t **, size_t);
template int dev_alloc<uint64_t>(Engine *, uint64_t **, size_t);
template int dev_alloc<uint16_t>(Engine *, uint16_t **, size_t);
template int dev_alloc<tb_local>(Engine *, tb_local **, size_t);
template int dev_alloc<tb_plane>(Engine *, tb_plane **, size_t);
template int dev_alloc<tb_ball>(Engine *, tb_ball **, size_t);
template int dev_alloc<float4>(Engine *, float4 **, size_t);

void dev_free(Engine *e, void *p, size_t bytes) {
    if (!p) return;
    cudaFree(p);
    e->device_bytes -= (int64_t)bytes;
}


// ---- staging buffers (pinned host + device), grown on demand ------------------------------------------------
static int stage(Engine *e, size_t bytes) {
    if (bytes > e->stage_dev_bytes) {
        if (e->stage_dev) { cudaFree(e->stage_dev); e->device_bytes -= (int64_t)e->stage_dev_bytes; }
        e->stage_dev = nullptr; e->stage_dev_bytes = 0;
        size_t want = std::max(bytes, (size_t)1 << 20);
        cudaError_t err = cudaMalloc(&e->stage_dev, want);
        if (err != cudaSuccess) { cudaGetLastError(); return fail(e, TB_ERR_OOM, "staging cudaMalloc(%zu) failed", want); }
        e->stage_dev_bytes = want; e->device_bytes += (int64_t)want;
    }
    return TB_OK;
}

// ---- SoA pack / unpack kernels (replace createMassPointers / fromMassPointers, src/sim.cu:933-939,1043-1049) --
// src: [count][3] AoS (titan::Vec order); dst: 3 planes of stride nm.  idx == nullptr -> contiguous from first.
__global__ void scatter_vec3(double *dst, int64_t nm, const double *src, const int32_t *idx, int64_t first, int64_t count) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    int64_t i = idx ? idx[j] : first + j;
    dst[i] = src[3 * j]; dst[nm + i] = src[3 * j + 1]; dst[2 * nm + i] = src[3 * j + 2];
}
__global__ void gather_vec3(double *dst, const double *src, int64_t nm, const int32_t *idx, int64_t first, int64_t count) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    int64_t i = idx ? idx[j] : first + j;
    dst[3 * j] = src[i]; dst[3 * j + 1] = src[nm + i]; dst[3 * j + 2] = src[2 * nm + i];
}
template <class T>
__global__ void scatter_scalar(T *dst, const T *src, const int32_t *idx, int64_t first, int64_t count) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    dst[idx ? idx[j] : first + j] = src[j];
}
template <class T>
__global__ void gather_scalar(T *dst, const T *src, const int32_t *idx, int64_t first, int64_t count) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    dst[j] = src[idx ? idx[j] : first + j];
}
// tombstones stay dead whatever the caller uploads later (tb_remove_springs)
__global__ void apply_tombstones_kernel(double *k, double *damping, const uint8_t *removed, int64_t ns) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ns || !removed[s]) return;
    k[s] = 0.0;
    if (damping) damping[s] = 0.0;
}
template <class T> __global__ void fill_kernel(T *dst, T v, int64_t n) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) dst[j] = v;
}
// after a step that consumed the one-shot external force: zero it for every mass that was updated
// (src/sim.cu:1365); fixed masses return before the reset (:1294) and keep theirs.
__global__ void clear_ext_kernel(double *ext, const uint8_t *flags, int64_t nm) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    uint8_t f = flags[i];
    if ((f & F_VALID) && !(f & F_FIXED)) { ext[i] = 0.0; ext[nm + i] = 0.0; ext[2 * nm + i] = 0.0; }
}
// set(Mass*) overwrites the whole CUDA_MASS with a freshly constructed one, whose acc is Vec() = 0 (src/sim.cu:603-616,
// src/mass.cu:74-100); only the Verlet integrator reads acc back, so only there it matters
__global__ void zero_vec3_kernel(double *dst, int64_t nm, const int32_t *idx, int64_t first, int64_t count) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    int64_t i = idx ? idx[j] : first + j;
    dst[i] = 0.0; dst[nm + i] = 0.0; dst[2 * nm + i] = 0.0;
}
__global__ void export_f32_kernel(float *dst, const double *pos, int64_t nm) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    dst[3 * i] = (float)pos[i]; dst[3 * i + 1] = (float)pos[nm + i]; dst[3 * i + 2] = (float)pos[2 * nm + i];
}

// updateIndices, src/sim.cu:1957-1980: the line list (left, right) per spring; skipped springs (null / invalid endpoint,
// here also tombstones) become the degenerate line (0, 0).  The reference searches the mass table per spring
// (O(Ns * Nm)); here the endpoints ARE indices.
__global__ void export_indices_kernel(uint32_t *dst, const int32_t *left, const int32_t *right, const uint8_t *flags,
                                      const uint8_t *removed, int64_t ns) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ns) return;
    int l = left[s], r = right[s];
    bool skip = l < 0 || r < 0 || !(flags[l] & F_VALID) || !(flags[r] & F_VALID) || (removed && removed[s]);
    dst[2 * s] = skip ? 0u : (uint32_t)l;
    dst[2 * s + 1] = skip ? 0u : (uint32_t)r;
}

static inline unsigned blocks_for(int64_t n, int t = 256) { return (unsigned)std::max<int64_t>(1, (n + t - 1) / t); }

template <class T> static int fill(Engine *e, T *dst, T v, int64_t n) {
    if (n <= 0) return TB_OK;
    fill_kernel<T><<<blocks_for(n), 256, 0, e->stream>>>(dst, v, n);
    TB_CUDA(e, cudaGetLastError());
    return TB_OK;
}

// Upload `count` elements of host array `src` (elem bytes each) to the device staging area at `offset`.
static int stage_in(Engine *e, const void *src, size_t bytes, size_t offset) {
    TB_CUDA(e, cudaMemcpyAsync((char *)e->stage_dev + offset, src, bytes, cudaMemcpyHostToDevice, e->stream));
    return TB_OK;
}

static int check_range(Engine *e, int64_t n, int64_t first, int64_t count, const int32_t *idx, const char *what) {
    if (count < 0) return fail(e, TB_ERR_INVALID, "%s: negative count", what);
    if (!idx && (first < 0 || first + count > n)) return fail(e, TB_ERR_INVALID, "%s: range [%lld,+%lld) outside [0,%lld)", what, (long long)first, (long long)count, (long long)n);
    if (idx) for (int64_t j = 0; j < count; j++) if (idx[j] < 0 || idx[j] >= n) return fail(e, TB_ERR_INVALID, "%s: index %d out of range", what, idx[j]);
    return TB_OK;
}

static const size_t ALIGN = 256;
static size_t align_up(size_t x) { return (x + ALIGN - 1) / ALIGN * ALIGN; }

// stage idx (if any) + one payload, returning device pointers
struct Staged { const int32_t *idx; const void *payload; };
static int stage_idx_and(Engine *e, const int32_t *idx, int64_t count, const void *src, size_t elem, Staged *out) {
    size_t ib = idx ? align_up((size_t)count * 4) : 0, pb = (size_t)count * elem;
    TB_TRY(stage(e, ib + pb));
    if (idx) TB_TRY(stage_in(e, idx, (size_t)count * 4, 0));
    TB_TRY(stage_in(e, src, pb, ib));
    out->idx = idx ? (const int32_t *)e->stage_dev : nullptr;
    out->payload = (const char *)e->stage_dev + ib;
    return TB_OK;
}

static int up_vec3(Engine *e, double *dst, const double *src, const int32_t *idx, int64_t first, int64_t count) {
    Staged s;
    TB_TRY(stage_idx_and(e, idx, count, src, 24, &s));
    scatter_vec3<<<blocks_for(count), 256, 0, e->stream>>>(dst, e->nm, (const double *)s.payload, s.idx, first, count);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));  // staging area is reused by the next field
    return TB_OK;
}
template <class T> static int up_scalar(Engine *e, T *dst, const T *src, const int32_t *idx, int64_t first, int64_t count) {
    Staged s;
    TB_TRY(stage_idx_and(e, idx, count, src, sizeof(T), &s));
    scatter_scalar<T><<<blocks_for(count), 256, 0, e->stream>>>(dst, (const T *)s.payload, s.idx, first, count);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return TB_OK;
}
static int down_vec3(Engine *e, double *dst_host, const double *src, const int32_t *idx, int64_t first, int64_t count) {
    size_t ib = idx ? align_up((size_t)count * 4) : 0;
    TB_TRY(stage(e, ib + (size_t)count * 24));
    if (idx) TB_TRY(stage_in(e, idx, (size_t)count * 4, 0));
    double *tmp = (double *)((char *)e->stage_dev + ib);
    gather_vec3<<<blocks_for(count), 256, 0, e->stream>>>(tmp, src, e->nm, idx ? (const int32_t *)e->stage_dev : nullptr, first, count);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaMemcpyAsync(dst_host, tmp, (size_t)count * 24, cudaMemcpyDeviceToHost, e->stream));
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return TB_OK;
}
template <class T> static int down_scalar(Engine *e, T *dst_host, const T *src, const int32_t *idx, int64_t first, int64_t count) {
    size_t ib = idx ? align_up((size_t)count * 4) : 0;
    TB_TRY(stage(e, ib + (size_t)count * sizeof(T)));
    if (idx) TB_TRY(stage_in(e, idx, (size_t)count * 4, 0));
    T *tmp = (T *)((char *)e->stage_dev + ib);
    gather_scalar<T><<<blocks_for(count), 256, 0, e->stream>>>(tmp, src, idx ? (const int32_t *)e->stage_dev : nullptr, first, count);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaMemcpyAsync(dst_host, tmp, (size_t)count * sizeof(T), cudaMemcpyDeviceToHost, e->stream));
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return TB_OK;
}

static void free_all(Engine *e) {
    auto F = [&](auto *&p, size_t bytes) { dev_free(e, (void *)p, bytes); p = nullptr; };
    const size_t ma = (size_t)e->m_alloc(), sa = (size_t)e->s_alloc();
    size_t m3 = ma * 24, m1 = ma * 8, s1 = sa * 8, s4 = sa * 4;
    F(e->pos[0], m3); F(e->pos[1], m3); F(e->vel[0], m3); F(e->vel[1], m3); F(e->acc, m3); F(e->ext, m3); F(e->rk2_bvel, m3);
    F(e->m, m1); F(e->mflags, ma); F(e->drag, ma * 4); F(e->m_scratch, m3);
    F(e->left, s4); F(e->right, s4); F(e->type, s4);
    F(e->k, s1); F(e->rest, s1); F(e->damping, s1); F(e->omega, s1); F(e->l_max, s1); F(e->l_min, s1); F(e->rate, s1);
    F(e->row_ptr, (size_t)e->row_ptr_cap * 4); F(e->adj, (size_t)e->adj_cap * 4); e->row_ptr_cap = e->adj_cap = 0;
    e->csr_nm = e->csr_ns = 0; e->extra_adj.clear(); e->patch_masses.clear(); e->patch_springs.clear();
    e->patch_needs_layout = e->removed_unpatched = e->compact_requested = false;
    F(e->export_f32, (size_t)e->nm * 12);
    F(e->export_idx, (size_t)e->ns * 8);
    F(e->removed_dev, sa);
    e->cap_s = e->cap_m = 0;
    F(e->act_idx, (size_t)e->n_act * 4); e->n_act = 0; e->act_dirty = true; e->h_act.clear();
    magnet_free(e);
    e->h_removed.clear(); e->n_removed = 0;
    if (e->lc_ptr) { cudaFree(e->lc_ptr); e->lc_ptr = nullptr; }
    if (e->lc) { cudaFree(e->lc); e->lc = nullptr; }
    for (int p = 0; p < 2; p++) if (e->graph[p]) { cudaGraphExecDestroy(e->graph[p]); e->graph[p] = nullptr; }
    tiled_destroy(e);
    e->nm = e->ns = 0;
}

template <class T> static int ensure_spring_array(Engine *e, T **p, T init) {
    if (*p) return TB_OK;
    TB_TRY(dev_alloc(e, p, (size_t)e->s_alloc()));
    return fill<T>(e, *p, init, e->ns);
}

// Per-mass local time (CUDA_MASS::T, advanced by `T += dt` once per step for every non-fixed valid mass,
// src/sim.cu:1362) is not stored per step on the device: it is replayed on the host from the value at the
// mass's last upload with the very same sequence of fp64 additions.
static double replay_T(Engine *e, int64_t i, uint8_t flags) {
    double T = e->h_T0[(size_t)i];
    if ((flags & F_FIXED) || !(flags & F_VALID)) return T;
    int64_t from = e->h_n0[(size_t)i], to = e->steps_done;
    for (size_t s = 0; s < e->dt_log.size() && from < to; s++) {
        int64_t seg_end = (s + 1 < e->dt_log.size()) ? e->dt_log[s + 1].first_step : to;
        if (seg_end <= from) continue;
        int64_t n = std::min(seg_end, to) - from;
        double dt = e->dt_log[s].dt;
        if (e->flags & TB_INTEGRATOR_RK2) { const double h = 0.5 * dt; for (int64_t q = 0; q < n; q++) { T += h; T += h; } }   // :1342,1348
        else for (int64_t q = 0; q < n; q++) T += dt;
        from += n;
    }
    return T;
}

}  // namespace tb

using namespace tb;

#define E_CHECK(e) do { if (!(e)) return TB_ERR_INVALID; cudaSetDevice((e)->device); } while (0)

extern "C" {

int tb_abi_version(void) { return TB_ABI_VERSION; }

int tb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

static std::string g_create_error;

int tb_create(int device, uint32_t flags, tb_engine **out) {
    if (!out) return TB_ERR_INVALID;
    *out = nullptr;
    int n = tb_device_count();
    if (n <= 0) { g_create_error = "no CUDA device: titan_b200 has no CPU path"; return TB_ERR_NO_DEVICE; }
    if (device < 0 || device >= n) { g_create_error = "device index out of range"; return TB_ERR_INVALID; }
    tb_engine *e = new (std::nothrow) tb_engine();
    if (!e) return TB_ERR_OOM;
    e->device = device;
    e->flags = flags;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&e->ev0) != cudaSuccess || cudaEventCreate(&e->ev1) != cudaSuccess) {
        g_create_error = std::string("CUDA initialisation failed: ") + cudaGetErrorString(cudaGetLastError());
        delete e;
        return TB_ERR_CUDA;
    }
    for (cudaEvent_t &ev : e->run_ev)
        if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { g_create_error = "cudaEventCreate failed"; delete e; return TB_ERR_CUDA; }
    cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (cudaMalloc((void **)&e->t_dev, 2 * sizeof(double)) != cudaSuccess) { delete e; return TB_ERR_OOM; }
    if (cudaMalloc((void **)&e->step_dev, 2 * sizeof(long long)) != cudaSuccess) { delete e; return TB_ERR_OOM; }
    e->G.g[0] = 0; e->G.g[1] = 0; e->G.g[2] = -9.81;  // Simulation::Simulation, src/sim.cu:86
    e->G.dt = 0.0001;                                   // src/sim.cu:78
    e->dt_log.push_back({0, e->G.dt});
    *out = e;
    return TB_OK;
}

void tb_destroy(tb_engine *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    free_all(e);
    if (e->stage_dev) cudaFree(e->stage_dev);
    if (e->stage_host) cudaFreeHost(e->stage_host);
    for (void *p : e->pinned) cudaFreeHost(p);
    if (e->more_planes) cudaFree(e->more_planes);
    if (e->more_balls) cudaFree(e->more_balls);
    cudaFree(e->t_dev);
    cudaFree(e->step_dev);
    cudaEventDestroy(e->ev0); cudaEventDestroy(e->ev1);
    for (cudaEvent_t ev : e->run_ev) if (ev) cudaEventDestroy(ev);
    cudaStreamDestroy(e->stream);
    delete e;
}

const char *tb_last_error(const tb_engine *e) { return e ? e->err.c_str() : g_create_error.c_str(); }

int tb_set_topology(tb_engine *e, int64_t nm, int64_t ns, const int32_t *left, const int32_t *right) {
    E_CHECK(e);
    if (nm <= 0 || ns < 0 || (ns > 0 && (!left || !right))) return fail(e, TB_ERR_INVALID, "tb_set_topology: bad sizes/pointers");
    if (nm >= (int64_t)1 << 31 || ns >= (int64_t)1 << 30) return fail(e, TB_ERR_INVALID, "tb_set_topology: index range exceeds int32 encoding");
    for (int64_t s = 0; s < ns; s++)
        if (left[s] < -1 || left[s] >= nm || right[s] < -1 || right[s] >= nm)
            return fail(e, TB_ERR_INVALID, "tb_set_topology: spring %lld has an endpoint outside [-1,%lld)", (long long)s, (long long)nm);
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    free_all(e);
    e->nm = nm; e->ns = ns;
    size_t n3 = (size_t)nm * 3;
    for (int p = 0; p < 2; p++) { TB_TRY(dev_alloc(e, &e->pos[p], n3)); TB_TRY(dev_alloc(e, &e->vel[p], n3)); }
    TB_TRY(dev_alloc(e, &e->acc, n3));
    TB_TRY(dev_alloc(e, &e->m, (size_t)nm));
    TB_TRY(dev_alloc(e, &e->mflags, (size_t)nm));
    TB_TRY(dev_alloc(e, &e->left, (size_t)ns)); TB_TRY(dev_alloc(e, &e->right, (size_t)ns));
    TB_TRY(dev_alloc(e, &e->k, (size_t)ns)); TB_TRY(dev_alloc(e, &e->rest, (size_t)ns));
    for (int p = 0; p < 2; p++) {
        TB_CUDA(e, cudaMemsetAsync(e->pos[p], 0, n3 * 8, e->stream));
        TB_CUDA(e, cudaMemsetAsync(e->vel[p], 0, n3 * 8, e->stream));
    }
    TB_CUDA(e, cudaMemsetAsync(e->acc, 0, n3 * 8, e->stream));
    TB_TRY(fill<double>(e, e->m, 0.1, nm));                 // Mass default, include/Titan/mass.h:18
    TB_TRY(fill<uint8_t>(e, e->mflags, F_VALID, nm));
    TB_TRY(fill<double>(e, e->k, 10000.0, ns));             // Spring default, include/Titan/spring.h:25
    TB_TRY(fill<double>(e, e->rest, 1.0, ns));
    if (ns) {
        TB_CUDA(e, cudaMemcpyAsync(e->left, left, (size_t)ns * 4, cudaMemcpyHostToDevice, e->stream));
        TB_CUDA(e, cudaMemcpyAsync(e->right, right, (size_t)ns * 4, cudaMemcpyHostToDevice, e->stream));
    }
    e->h_left.reserve((size_t)(ns + ns / 8 + 4096)); e->h_right.reserve((size_t)(ns + ns / 8 + 4096));   // tb_add_springs appends without a copy
    e->h_left.assign(left, left + ns); e->h_right.assign(right, right + ns);
    e->csr_nm = e->csr_ns = 0; e->extra_adj.clear();
    TB_TRY(ensure_csr(e));
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    e->h_flags.assign((size_t)nm, F_VALID);
    e->h_T0.assign((size_t)nm, 0.0);
    e->h_n0.assign((size_t)nm, e->steps_done);
    e->cur = 0;
    e->ext_pending = false;
    e->any_damping = e->any_active = e->any_actuated = e->any_local = e->any_magnet = e->any_fixed_or_invalid = e->any_ghost = false;
    e->graph_dirty = true; e->tiled_stale = 0xf; e->canon_stale = false; e->tile_hint.clear();
    e->h_type.assign((size_t)ns, 0);
    e->h_omega.clear();
    e->stiff_seen = e->scale_seen = false;
    return TB_OK;
}

int tb_set_tile_hint(tb_engine *e, const int32_t *tile_of_mass) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_set_tile_hint before tb_set_topology");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (tile_of_mass) e->tile_hint.assign(tile_of_mass, tile_of_mass + e->nm);
    else e->tile_hint.clear();
    e->tiled_stale |= TS_STRUCT;
    e->graph_dirty = true;
    return TB_OK;
}

static bool any_nonzero(const double *a, int64_t n) { for (int64_t i = 0; i < n; i++) if (a[i] != 0.0) return true; return false; }

int tb_set_mass_state(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, const double *pos,
                      const double *vel, const double *m, const double *T, const double *extern_force,
                      const uint8_t *fixed, const int32_t *drag, const uint8_t *valid) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_set_mass_state before tb_set_topology");
    TB_TRY(check_range(e, e->nm, first, count, idx, "tb_set_mass_state"));
    if (count == 0) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if ((pos || vel) && e->canon_stale) TB_TRY(tiled_to_canonical(e));
    if (pos) TB_TRY(up_vec3(e, e->pos[e->cur], pos, idx, first, count));
    if (vel) TB_TRY(up_vec3(e, e->vel[e->cur], vel, idx, first, count));
    if ((pos || vel) && (e->flags & TB_INTEGRATOR_VERLET)) {
        int32_t *didx = nullptr;
        if (idx) { TB_TRY(stage(e, (size_t)count * 4)); TB_TRY(stage_in(e, idx, (size_t)count * 4, 0)); didx = (int32_t *)e->stage_dev; }
        zero_vec3_kernel<<<blocks_for(count), 256, 0, e->stream>>>(e->acc, e->nm, didx, first, count);
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    if (m) TB_TRY(up_scalar<double>(e, e->m, m, idx, first, count));
    if (extern_force) {
        bool nz = any_nonzero(extern_force, count * 3);
        if (nz && !e->ext) {
            TB_TRY(dev_alloc(e, &e->ext, (size_t)e->m_alloc() * 3));
            TB_CUDA(e, cudaMemsetAsync(e->ext, 0, (size_t)e->nm * 24, e->stream));
        }
        if (e->ext) TB_TRY(up_vec3(e, e->ext, extern_force, idx, first, count));
        if (nz) e->ext_pending = true;
    }
    if (drag) {
        bool nz = false;
        for (int64_t j = 0; j < count; j++) nz |= drag[j] != 0;
        if (nz && !e->drag) {
            TB_TRY(dev_alloc(e, &e->drag, (size_t)e->m_alloc()));
            TB_CUDA(e, cudaMemsetAsync(e->drag, 0, (size_t)e->nm * 4, e->stream));
        }
        if (e->drag) TB_TRY(up_scalar<int32_t>(e, e->drag, drag, idx, first, count));
    }
    bool flags_changed = false, valid_changed = false;
    std::vector<int32_t> orphaned;        // springs of masses that just became invalid
    for (int64_t j = 0; j < count; j++) {
        int64_t i = idx ? idx[j] : first + j;
        uint8_t f = e->h_flags[(size_t)i], nf = f;
        if (fixed) nf = (uint8_t)((nf & ~F_FIXED) | (fixed[j] ? F_FIXED : 0));
        if (valid) nf = (uint8_t)((nf & ~F_VALID) | (valid[j] ? F_VALID : 0));
        if (nf != f) {
            e->h_flags[(size_t)i] = nf; flags_changed = true;
            if ((f & F_VALID) && !(nf & F_VALID) && !(e->flags & TB_KERNEL_CSR)) {
                // Simulation::deleteMass after start (src/sim.cu:343-380): the mass is frozen and every spring attached to it
                // is skipped from now on (:1163).  No structure rebuild: the tile's flag byte freezes the mass, and its springs
                // become tombstones (tb_remove_springs), which is bit-identical to skipping them.
                if (i < e->csr_nm) for (int32_t q = e->h_row_ptr[(size_t)i]; q < e->h_row_ptr[(size_t)i + 1]; q++) orphaned.push_back(e->h_adj[(size_t)q] >> 1);
                auto it = e->extra_adj.find((int32_t)i);
                if (it != e->extra_adj.end()) for (int32_t a : it->second) orphaned.push_back(a >> 1);
            } else valid_changed |= ((nf ^ f) & F_VALID) != 0;
        }
        if (T) { e->h_T0[(size_t)i] = T[j]; e->h_n0[(size_t)i] = e->steps_done; }
        else if (nf != f) {
            // fixed/valid status changes: freeze the T accumulated so far (under the OLD flags) and restart
            e->h_T0[(size_t)i] = replay_T(e, i, f);
            e->h_n0[(size_t)i] = e->steps_done;
        }
    }
    if (flags_changed) {
        TB_CUDA(e, cudaMemcpyAsync(e->mflags, e->h_flags.data(), (size_t)e->nm, cudaMemcpyHostToDevice, e->stream));
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
        e->any_fixed_or_invalid = false;
        for (uint8_t f : e->h_flags) if ((f & F_FIXED) || !(f & F_VALID)) { e->any_fixed_or_invalid = true; break; }
    }
    if (pos || vel || m) e->tiled_stale |= TS_STATE;
    if (flags_changed) e->tiled_stale |= TS_FLAGS | (valid_changed ? TS_STRUCT : 0u);
    if (flags_changed || drag) e->graph_dirty = true;
    if (!orphaned.empty()) {
        std::sort(orphaned.begin(), orphaned.end());
        orphaned.erase(std::unique(orphaned.begin(), orphaned.end()), orphaned.end());
        TB_TRY(tb_remove_springs(e, (int64_t)orphaned.size(), orphaned.data()));
    }
    return TB_OK;
}

}  // extern "C"

extern "C" {

int tb_set_mass_magnet(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, const double *rad,
                       const double *stiffness, const double *max_mag_force, const double *mag_scale) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_set_mass_magnet before tb_set_topology");
    TB_TRY(check_range(e, e->nm, first, count, idx, "tb_set_mass_magnet"));
    if (count == 0) return TB_OK;
    // The occupancy-grid magnet pass (src/sim.cu:865-906,1223-1281) contributes exactly +0 while every mass has
    // rad == 0 and max_mag_force == 0 (all lattice / cube / beam masses): until the first magnetic mass shows up the
    // fields are not even stored and the pass is elided.  From then on the engine runs it every step (magnet.cu).
    bool magnetic = (rad && any_nonzero(rad, count)) || (max_mag_force && any_nonzero(max_mag_force, count));
    if (!e->mag_rad) {
        // Not stored yet.  A non-magnetic mass still takes part once another mass is magnetic -- its shell stiffness
        // scales the intersection force it feels (:1230-1232), its mag_scale_factor the pull it exerts (:1235-1237) --
        // so remember those two while they are uniform (1000 and 0 for every Mass(pos), include/Titan/mass.h:18-19)
        // and store the arrays as soon as they are not.
        auto uniform = [&](const double *src, bool &seen, double &value) {
            if (!src) return true;
            for (int64_t j = 0; j < count; j++) {
                if (!seen) { seen = true; value = src[j]; }
                else if (memcmp(&value, &src[j], 8) != 0) return false;
            }
            return true;
        };
        bool keep_lazy = uniform(stiffness, e->stiff_seen, e->stiff_uniform) & uniform(mag_scale, e->scale_seen, e->scale_uniform);
        if (!magnetic && keep_lazy) return TB_OK;
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
        TB_TRY(magnet_alloc(e));
        if (e->stiff_seen) TB_TRY(fill<double>(e, e->mag_stiff, e->stiff_uniform, e->nm));
        if (e->scale_seen) TB_TRY(fill<double>(e, e->mag_scale, e->scale_uniform, e->nm));
    }
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (rad) TB_TRY(up_scalar<double>(e, e->mag_rad, rad, idx, first, count));
    if (stiffness) TB_TRY(up_scalar<double>(e, e->mag_stiff, stiffness, idx, first, count));
    if (max_mag_force) TB_TRY(up_scalar<double>(e, e->mag_max, max_mag_force, idx, first, count));
    if (mag_scale) TB_TRY(up_scalar<double>(e, e->mag_scale, mag_scale, idx, first, count));
    if (magnetic && !e->any_magnet) { e->any_magnet = true; e->graph_dirty = true; }
    return TB_OK;
}

int tb_get_mass_state(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel,
                      double *acc, double *T, double *extern_force) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_get_mass_state before tb_set_topology");
    TB_TRY(check_range(e, e->nm, first, count, idx, "tb_get_mass_state"));
    if (count == 0) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if ((pos || vel) && e->canon_stale) TB_TRY(tiled_to_canonical(e));
    if (pos) TB_TRY(down_vec3(e, pos, e->pos[e->cur], idx, first, count));
    if (vel) TB_TRY(down_vec3(e, vel, e->vel[e->cur], idx, first, count));
    if (acc) TB_TRY(down_vec3(e, acc, e->acc, idx, first, count));
    if (extern_force) {
        if (e->ext) TB_TRY(down_vec3(e, extern_force, e->ext, idx, first, count));
        else memset(extern_force, 0, (size_t)count * 24);
    }
    if (T) {
        // masses uploaded together share (T0, n0): memoise the last replay
        double key_T0 = 0, val = 0; int64_t key_n0 = -1;
        for (int64_t j = 0; j < count; j++) {
            int64_t i = idx ? idx[j] : first + j;
            uint8_t f = (uint8_t)(e->h_flags[(size_t)i] & (F_FIXED | F_VALID));
            if ((f & F_FIXED) || !(f & F_VALID)) { T[j] = e->h_T0[(size_t)i]; continue; }   // frozen: no replay, memo untouched
            if (!(key_n0 == e->h_n0[(size_t)i] && memcmp(&key_T0, &e->h_T0[(size_t)i], 8) == 0)) {
                key_T0 = e->h_T0[(size_t)i]; key_n0 = e->h_n0[(size_t)i];
                val = replay_T(e, i, f);
            }
            T[j] = val;
        }
    }
    return TB_OK;
}

int tb_set_spring_params(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, const double *k,
                         const double *rest, const double *damping, const int32_t *type, const double *omega,
                         const double *l_max, const double *l_min, const double *rate) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_set_spring_params before tb_set_topology");
    TB_TRY(check_range(e, e->ns, first, count, idx, "tb_set_spring_params"));
    if (count == 0) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (k) TB_TRY(up_scalar<double>(e, e->k, k, idx, first, count));
    if (rest) TB_TRY(up_scalar<double>(e, e->rest, rest, idx, first, count));
    if (damping && (e->damping || any_nonzero(damping, count))) {
        TB_TRY(ensure_spring_array<double>(e, &e->damping, 0.0));
        TB_TRY(up_scalar<double>(e, e->damping, damping, idx, first, count));
        e->any_damping = true;
    }
    bool type_changed = false;    // the tile structure encodes the actuation class per incidence word: rebuilt only on a real change
    if (type) {
        bool nonpassive = false;
        for (int64_t j = 0; j < count; j++) {
            if (type[j] < 0 || type[j] > 5) return fail(e, TB_ERR_INVALID, "tb_set_spring_params: bad SpringType %d", type[j]);
            nonpassive |= type[j] >= 2;
        }
        if (e->h_type.size() != (size_t)e->ns) e->h_type.assign((size_t)e->ns, 0);
        for (int64_t j = 0; j < count; j++) {
            uint8_t &held = e->h_type[(size_t)(idx ? idx[j] : first + j)];
            type_changed |= (held < 2 ? 0 : held) != (type[j] < 2 ? 0 : type[j]);   // PASSIVE_SOFT and PASSIVE_STIFF are one class
            held = (uint8_t)type[j];
        }
        e->any_active = e->any_actuated = false;
        for (uint8_t ty : e->h_type) { if (ty == 2 || ty == 3) e->any_active = true; if (ty >= 4) e->any_actuated = true; }
        if (type_changed) e->act_dirty = true;
        if (e->type || nonpassive) {
            TB_TRY(ensure_spring_array<int32_t>(e, &e->type, 0));
            TB_TRY(up_scalar<int32_t>(e, e->type, type, idx, first, count));
        }
    }
    if (omega) {
        if (e->h_omega.size() != (size_t)e->ns) e->h_omega.assign((size_t)e->ns, 0.0);
        for (int64_t j = 0; j < count; j++) e->h_omega[(size_t)(idx ? idx[j] : first + j)] = omega[j];
    }
    struct { const double *src; double **dst; } opt[4] = {{omega, &e->omega}, {l_max, &e->l_max}, {l_min, &e->l_min}, {rate, &e->rate}};
    for (auto &o : opt) {
        if (!o.src) continue;
        if (!*o.dst && !any_nonzero(o.src, count)) continue;
        TB_TRY(ensure_spring_array<double>(e, o.dst, 0.0));
        TB_TRY(up_scalar<double>(e, *o.dst, o.src, idx, first, count));
    }
    if (e->n_removed > 0 && (k || damping)) {
        apply_tombstones_kernel<<<blocks_for(e->ns), 256, 0, e->stream>>>(e->k, e->damping, e->removed_dev, e->ns);
        TB_CUDA(e, cudaGetLastError());
    }
    e->tiled_stale |= TS_PARAMS | (type_changed ? TS_STRUCT : 0u);
    e->graph_dirty = true;
    return TB_OK;
}

int tb_remove_springs(tb_engine *e, int64_t n, const int32_t *idx) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_remove_springs before tb_set_topology");
    if (n < 0 || (n > 0 && !idx)) return fail(e, TB_ERR_INVALID, "tb_remove_springs: bad arguments");
    TB_TRY(check_range(e, e->ns, 0, n, idx, "tb_remove_springs"));
    if (n == 0) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->h_removed.empty()) {
        e->h_removed.assign((size_t)e->ns, 0);
        TB_TRY(dev_alloc(e, &e->removed_dev, (size_t)e->s_alloc()));
        TB_CUDA(e, cudaMemsetAsync(e->removed_dev, 0, (size_t)e->ns, e->stream));
    }
    std::vector<uint8_t> ones((size_t)n, 1);
    for (int64_t j = 0; j < n; j++)
        if (!e->h_removed[(size_t)idx[j]]) { e->h_removed[(size_t)idx[j]] = 1; e->n_removed++; }
    TB_TRY(up_scalar<uint8_t>(e, e->removed_dev, ones.data(), idx, 0, n));
    apply_tombstones_kernel<<<blocks_for(e->ns), 256, 0, e->stream>>>(e->k, e->damping, e->removed_dev, e->ns);
    TB_CUDA(e, cudaGetLastError());
    // A handful of springs: the tiles holding them are rebuilt without them (no dead incidence words, parameter classes
    // untouched).  Many (topology optimisation removes 0.1 % of all springs at a time, everywhere): they stay in place with
    // k = 0 until the next rebuild of their tile or tb_compact.
    if (n <= 4096 && e->patch_springs.size() + (size_t)n <= 8192 && !e->any_ghost && !tiled_halo_connected(e)) {
        for (int64_t j = 0; j < n; j++) e->patch_springs.push_back(idx[j]);
        e->tiled_stale |= TS_PATCH;
    } else e->removed_unpatched = true;
    e->tiled_stale |= TS_PARAMS;
    e->graph_dirty = true;
    return TB_OK;
}

int tb_compact(tb_engine *e) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_compact before tb_set_topology");
    if (tiled_halo_connected(e)) return fail(e, TB_ERR_UNSUPPORTED, "tb_compact: not on a connected engine of a partition (its peers hold offsets into the layout)");
    e->compact_requested = true;
    e->tiled_stale |= TS_STRUCT;          // this rebuild leaves the incidence words of removed springs out of the rows
    e->graph_dirty = true;
    return TB_OK;
}

}  // extern "C"

namespace tb {
// n_old -> n_new elements (x3 planes when `planes` == 3: [3][n] arrays change their plane stride), new entries = value
template <class T> __global__ void regrow_kernel(T *dst, const T *src, int64_t n_old, int64_t n_new, int planes, T value) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_new * planes) return;
    const int64_t c = q / n_new, i = q % n_new;
    dst[q] = i < n_old ? src[c * n_old + i] : value;
}
// a_old / a_new: elements allocated per plane before / after.  Enough room: planes == 1 -> the new entries are filled in place;
// planes == 3 -> the planes move apart inside a buffer of the same size, through *scratch, which takes over as the array (the old
// buffer becomes the scratch of the next array).  Not enough room: a new buffer of a_new elements per plane.
template <class T> static int regrow(Engine *e, T **p, int64_t n_old, int64_t n_new, int planes, T value, int64_t a_old, int64_t a_new, T **scratch = nullptr) {
    if (!*p) return TB_OK;             // optional array that was never allocated: stays that way
    if (a_new == a_old && planes == 1) {
        fill_kernel<T><<<blocks_for(n_new - n_old), 256, 0, e->stream>>>(*p + n_old, value, n_new - n_old);
        TB_CUDA(e, cudaGetLastError());
        return TB_OK;
    }
    if (a_new == a_old && scratch && *scratch) {
        regrow_kernel<T><<<blocks_for(n_new * planes), 256, 0, e->stream>>>(*scratch, *p, n_old, n_new, planes, value);
        TB_CUDA(e, cudaGetLastError());
        std::swap(*p, *scratch);
        return TB_OK;
    }
    T *bigger = nullptr;
    TB_TRY(dev_alloc(e, &bigger, (size_t)(a_new * planes)));
    regrow_kernel<T><<<blocks_for(n_new * planes), 256, 0, e->stream>>>(bigger, *p, n_old, n_new, planes, value);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    dev_free(e, (void *)*p, (size_t)(a_old * planes) * sizeof(T));
    *p = bigger;
    return TB_OK;
}
static int64_t with_room(int64_t n) { return n + n / 8 + 4096; }
}  // namespace tb

extern "C" {

int tb_add_masses(tb_engine *e, int64_t n) {
    E_CHECK(e);
    TimeScope ts("tb_add_masses");
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_add_masses before tb_set_topology");
    if (n < 0 || e->nm + n >= (int64_t)1 << 31) return fail(e, TB_ERR_INVALID, "tb_add_masses: bad count");
    if (n == 0) return TB_OK;
    if (e->any_ghost || tiled_halo_connected(e)) return fail(e, TB_ERR_UNSUPPORTED, "tb_add_masses: not on an engine of a partition (re-upload the topology)");
    if (e->mag_rad || e->lc_ptr) return fail(e, TB_ERR_UNSUPPORTED, "tb_add_masses: not with magnet fields or per-mass constraints on the device (re-upload the topology)");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->canon_stale) TB_TRY(tiled_to_canonical(e));     // the canonical arrays are about to move
    const int64_t n0 = e->nm, n1 = e->nm + n;
    const int64_t a0 = e->m_alloc(), a1 = n1 <= a0 ? a0 : with_room(n1);
    if (a1 != a0) { dev_free(e, (void *)e->m_scratch, (size_t)a0 * 24); e->m_scratch = nullptr; }
    else if (!e->m_scratch) TB_TRY(dev_alloc(e, &e->m_scratch, (size_t)a0 * 3));
    for (int p = 0; p < 2; p++) {
        TB_TRY(regrow<double>(e, &e->pos[p], n0, n1, 3, 0.0, a0, a1, &e->m_scratch));
        TB_TRY(regrow<double>(e, &e->vel[p], n0, n1, 3, 0.0, a0, a1, &e->m_scratch));
    }
    TB_TRY(regrow<double>(e, &e->acc, n0, n1, 3, 0.0, a0, a1, &e->m_scratch));
    TB_TRY(regrow<double>(e, &e->ext, n0, n1, 3, 0.0, a0, a1, &e->m_scratch));
    if (e->rk2_bvel) TB_TRY(regrow<double>(e, &e->rk2_bvel, n0, n1, 3, 0.0, a0, a1, &e->m_scratch));
    TB_TRY(regrow<double>(e, &e->m, n0, n1, 1, 0.1, a0, a1));                    // Mass default, include/Titan/mass.h:18
    TB_TRY(regrow<uint8_t>(e, &e->mflags, n0, n1, 1, (uint8_t)F_VALID, a0, a1));
    TB_TRY(regrow<int32_t>(e, &e->drag, n0, n1, 1, 0, a0, a1));
    e->cap_m = a1;
    dev_free(e, (void *)e->export_f32, (size_t)n0 * 12); e->export_f32 = nullptr;
    e->h_flags.resize((size_t)n1, F_VALID);
    e->h_T0.resize((size_t)n1, 0.0);
    e->h_n0.resize((size_t)n1, e->steps_done);
    if (!e->tile_hint.empty()) {       // the new masses stay together in tiles of their own at the next full build
        int32_t fresh = 0;
        for (int32_t h : e->tile_hint) fresh = std::max(fresh, h + 1);
        e->tile_hint.resize((size_t)n1, fresh);
    }
    for (int64_t i = n0; i < n1; i++) e->patch_masses.push_back((int32_t)i);
    e->patch_needs_layout = true;
    e->nm = n1;
    e->tiled_stale |= TS_PATCH | TS_STATE | TS_FLAGS;
    e->graph_dirty = true;
    return TB_OK;
}

int tb_add_springs(tb_engine *e, int64_t n, const int32_t *left, const int32_t *right) {
    E_CHECK(e);
    TimeScope ts("tb_add_springs");
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_add_springs before tb_set_topology");
    if (n < 0 || (n > 0 && (!left || !right)) || e->ns + n >= (int64_t)1 << 30) return fail(e, TB_ERR_INVALID, "tb_add_springs: bad arguments");
    if (n == 0) return TB_OK;
    for (int64_t j = 0; j < n; j++)
        if (left[j] < -1 || left[j] >= e->nm || right[j] < -1 || right[j] >= e->nm) return fail(e, TB_ERR_INVALID, "tb_add_springs: spring %lld has an endpoint outside [-1,%lld)", (long long)j, (long long)e->nm);
    if (tiled_halo_connected(e)) return fail(e, TB_ERR_UNSUPPORTED, "tb_add_springs: not on a connected engine of a partition (re-upload the topology)");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    const int64_t n0 = e->ns, n1 = e->ns + n;
    const int64_t a0 = e->s_alloc(), a1 = (n0 > 0 && n1 <= a0) ? a0 : with_room(n1);
    if (n0 == 0) {       // a system that had no spring at all: the arrays were allocated with one placeholder element
        dev_free(e, (void *)e->left, 4); dev_free(e, (void *)e->right, 4); dev_free(e, (void *)e->k, 8); dev_free(e, (void *)e->rest, 8);
        TB_TRY(dev_alloc(e, &e->left, (size_t)a1)); TB_TRY(dev_alloc(e, &e->right, (size_t)a1)); TB_TRY(dev_alloc(e, &e->k, (size_t)a1)); TB_TRY(dev_alloc(e, &e->rest, (size_t)a1));
        TB_TRY(fill<double>(e, e->k, 10000.0, n1)); TB_TRY(fill<double>(e, e->rest, 1.0, n1));
    } else {
        TB_TRY(regrow<int32_t>(e, &e->left, n0, n1, 1, -1, a0, a1)); TB_TRY(regrow<int32_t>(e, &e->right, n0, n1, 1, -1, a0, a1));
        TB_TRY(regrow<double>(e, &e->k, n0, n1, 1, 10000.0, a0, a1));               // Spring default, include/Titan/spring.h:25
        TB_TRY(regrow<double>(e, &e->rest, n0, n1, 1, 1.0, a0, a1));
    }
    TB_TRY(regrow<int32_t>(e, &e->type, n0, n1, 1, 0, a0, a1));
    TB_TRY(regrow<double>(e, &e->damping, n0, n1, 1, 0.0, a0, a1)); TB_TRY(regrow<double>(e, &e->omega, n0, n1, 1, 0.0, a0, a1));
    TB_TRY(regrow<double>(e, &e->l_max, n0, n1, 1, 0.0, a0, a1)); TB_TRY(regrow<double>(e, &e->l_min, n0, n1, 1, 0.0, a0, a1));
    TB_TRY(regrow<double>(e, &e->rate, n0, n1, 1, 0.0, a0, a1));
    TB_TRY(regrow<uint8_t>(e, &e->removed_dev, n0, n1, 1, (uint8_t)0, a0, a1));
    e->cap_s = a1;
    dev_free(e, (void *)e->export_idx, (size_t)n0 * 8); e->export_idx = nullptr;
    TB_CUDA(e, cudaMemcpy(e->left + n0, left, (size_t)n * 4, cudaMemcpyHostToDevice));
    TB_CUDA(e, cudaMemcpy(e->right + n0, right, (size_t)n * 4, cudaMemcpyHostToDevice));
    e->h_left.insert(e->h_left.end(), left, left + n); e->h_right.insert(e->h_right.end(), right, right + n);
    e->h_type.resize((size_t)n1, 0);
    if (!e->h_omega.empty()) e->h_omega.resize((size_t)n1, 0.0);
    if (!e->h_removed.empty()) e->h_removed.resize((size_t)n1, 0);
    for (int64_t j = 0; j < n; j++) {
        const int32_t s = (int32_t)(n0 + j);
        e->patch_springs.push_back(s);
        if (left[j] < 0 || right[j] < 0) continue;
        e->extra_adj[right[j]].push_back(2 * s + 1);
        e->extra_adj[left[j]].push_back(2 * s);
    }
    e->ns = n1;
    e->patch_needs_layout = true;
    e->tiled_stale |= TS_PATCH | TS_PARAMS;
    e->graph_dirty = true;
    return TB_OK;
}

int tb_get_spring_rest(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, double *rest) {
    E_CHECK(e);
    TB_TRY(check_range(e, e->ns, first, count, idx, "tb_get_spring_rest"));
    if (count == 0 || !rest) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return down_scalar<double>(e, rest, e->rest, idx, first, count);
}

static void refresh_globals(tb_engine *e) {
    e->G.n_planes = (int32_t)e->planes.size();
    e->G.n_balls = (int32_t)e->balls.size();
    for (int j = 0; j < TB_INLINE_C; j++) {
        if (j < (int)e->planes.size()) e->G.planes[j] = e->planes[j];
        if (j < (int)e->balls.size()) e->G.balls[j] = e->balls[j];
    }
    e->G.more_planes = e->more_planes;
    e->G.more_balls = e->more_balls;
    e->graph_dirty = true;
}

int tb_set_planes(tb_engine *e, int32_t n, const tb_plane *planes) {
    E_CHECK(e);
    if (n < 0 || (n > 0 && !planes)) return fail(e, TB_ERR_INVALID, "tb_set_planes: bad arguments");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    e->planes.assign(planes, planes + n);
    if (e->more_planes) { cudaFree(e->more_planes); e->more_planes = nullptr; }
    if (n > TB_INLINE_C) {
        TB_CUDA(e, cudaMalloc((void **)&e->more_planes, (size_t)(n - TB_INLINE_C) * sizeof(tb_plane)));
        TB_CUDA(e, cudaMemcpy(e->more_planes, planes + TB_INLINE_C, (size_t)(n - TB_INLINE_C) * sizeof(tb_plane), cudaMemcpyHostToDevice));
    }
    refresh_globals(e);
    return TB_OK;
}

int tb_set_balls(tb_engine *e, int32_t n, const tb_ball *balls) {
    E_CHECK(e);
    if (n < 0 || (n > 0 && !balls)) return fail(e, TB_ERR_INVALID, "tb_set_balls: bad arguments");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    e->balls.assign(balls, balls + n);
    if (e->more_balls) { cudaFree(e->more_balls); e->more_balls = nullptr; }
    if (n > TB_INLINE_C) {
        TB_CUDA(e, cudaMalloc((void **)&e->more_balls, (size_t)(n - TB_INLINE_C) * sizeof(tb_ball)));
        TB_CUDA(e, cudaMemcpy(e->more_balls, balls + TB_INLINE_C, (size_t)(n - TB_INLINE_C) * sizeof(tb_ball), cudaMemcpyHostToDevice));
    }
    refresh_globals(e);
    return TB_OK;
}

int tb_set_local_constraints(tb_engine *e, const int64_t *row_ptr, const tb_local *records) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_set_local_constraints before tb_set_topology");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->lc_ptr) { cudaFree(e->lc_ptr); e->lc_ptr = nullptr; }
    if (e->lc) { cudaFree(e->lc); e->lc = nullptr; }
    int64_t total = row_ptr ? row_ptr[e->nm] : 0;
    e->any_local = total > 0;
    bool changed = false;
    std::vector<int32_t> ptr32;
    if (total > 0) {
        if (!records || total >= (int64_t)1 << 31) return fail(e, TB_ERR_INVALID, "tb_set_local_constraints: bad records");
        ptr32.resize((size_t)e->nm + 1);
        for (int64_t i = 0; i <= e->nm; i++) ptr32[(size_t)i] = (int32_t)row_ptr[i];
        for (int64_t j = 0; j < total; j++)
            if (records[j].kind < 0 || records[j].kind > 3) return fail(e, TB_ERR_INVALID, "tb_set_local_constraints: bad kind");
        TB_CUDA(e, cudaMalloc((void **)&e->lc_ptr, ((size_t)e->nm + 1) * 4));
        TB_CUDA(e, cudaMalloc((void **)&e->lc, (size_t)total * sizeof(tb_local)));
        TB_CUDA(e, cudaMemcpy(e->lc_ptr, ptr32.data(), ((size_t)e->nm + 1) * 4, cudaMemcpyHostToDevice));
        TB_CUDA(e, cudaMemcpy(e->lc, records, (size_t)total * sizeof(tb_local), cudaMemcpyHostToDevice));
    }
    for (int64_t i = 0; i < e->nm; i++) {
        bool has = total > 0 && ptr32[(size_t)i + 1] > ptr32[(size_t)i];
        uint8_t f = e->h_flags[(size_t)i], nf = (uint8_t)((f & ~F_LOCAL) | (has ? F_LOCAL : 0));
        if (nf != f) { e->h_flags[(size_t)i] = nf; changed = true; }
    }
    if (changed) TB_CUDA(e, cudaMemcpy(e->mflags, e->h_flags.data(), (size_t)e->nm, cudaMemcpyHostToDevice));
    e->graph_dirty = true; e->tiled_stale |= TS_FLAGS;
    return TB_OK;
}

int tb_set_globals(tb_engine *e, const double g[3], double dt) {
    E_CHECK(e);
    if (!(dt > 0)) return fail(e, TB_ERR_INVALID, "Cannot set time step to negative or zero value.");  // src/sim.cu:803-805
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (g) { e->G.g[0] = g[0]; e->G.g[1] = g[1]; e->G.g[2] = g[2]; }
    if (dt != e->G.dt) {
        e->G.dt = dt;
        if (e->dt_log.back().first_step == e->steps_done) e->dt_log.back().dt = dt;
        else e->dt_log.push_back({e->steps_done, dt});
    }
    e->graph_dirty = true;
    return TB_OK;
}

}  // extern "C"

// ---- run loop ---------------------------------------------------------------------------------------------------
namespace tb {

__global__ void set_t_kernel(double *t, double v, long long *step, long long n) { *t = v; *step = n; }

// ACTUATED_EXPAND / ACTUATED_CONTRACT mutate _rest on the device before the force is evaluated
// (src/sim.cu:1174-1181); springs with a null or invalid endpoint are skipped first (:1163).
// One thread per ACTUATED_* spring (the list e->act_idx): the generic kernel's pre-pass.  The tiled kernels advance the
// records of their tiles themselves (tiled.cu: advance_actuated_records) and keep e->rest current.
__global__ void advance_rest_kernel(double *rest, const int32_t *type, const double *l_max, const double *l_min,
                                    const double *rate, const int32_t *left, const int32_t *right,
                                    const uint8_t *flags, double dt, const int32_t *act_idx, int64_t n_act) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_act) return;
    const int32_t s = act_idx[j];
    int ty = type[s];
    if (ty < TB_ACTUATED_EXPAND) return;
    int l = left[s], r = right[s];
    if (l < 0 || r < 0 || !(flags[l] & F_VALID) || !(flags[r] & F_VALID)) return;
    double x = rest[s];
    if (ty == TB_ACTUATED_EXPAND) { if (x < l_max[s]) x = __dadd_rn(x, __dmul_rn(rate[s], dt)); }
    else { if (x > l_min[s]) x = __dsub_rn(x, __dmul_rn(rate[s], dt)); }
    rest[s] = x;
}

// CSR incidence from the host copy of the topology: ascending spring index per mass; right entry before left entry (the
// reference adds to the right endpoint first, src/sim.cu:1187-1193).  Springs with a null endpoint have no entry.
int ensure_csr(Engine *e) {
    if (e->csr_nm == e->nm && e->csr_ns == e->ns && e->extra_adj.empty()) return TB_OK;
    const int64_t nm = e->nm, ns = e->ns;
    const std::vector<int32_t> &left = e->h_left, &right = e->h_right;
    std::vector<int32_t> &ptr = e->h_row_ptr;
    ptr.assign((size_t)nm + 1, 0);
    for (int64_t s = 0; s < ns; s++) {
        if (left[(size_t)s] < 0 || right[(size_t)s] < 0) continue;
        ptr[(size_t)right[(size_t)s] + 1]++; ptr[(size_t)left[(size_t)s] + 1]++;
    }
    for (int64_t i = 0; i < nm; i++) ptr[(size_t)i + 1] += ptr[(size_t)i];
    std::vector<int32_t> &adj = e->h_adj;
    adj.assign((size_t)std::max<int64_t>(1, ptr[(size_t)nm]), 0);
    {
        std::vector<int32_t> cursor(ptr.begin(), ptr.end() - 1);
        for (int64_t s = 0; s < ns; s++) {
            if (left[(size_t)s] < 0 || right[(size_t)s] < 0) continue;
            adj[(size_t)cursor[(size_t)right[(size_t)s]]++] = (int32_t)(2 * s + 1);
            adj[(size_t)cursor[(size_t)left[(size_t)s]]++] = (int32_t)(2 * s);
        }
    }
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->row_ptr_cap < nm + 1) {
        dev_free(e, (void *)e->row_ptr, (size_t)e->row_ptr_cap * 4); e->row_ptr = nullptr;
        TB_TRY(dev_alloc(e, &e->row_ptr, (size_t)nm + 1)); e->row_ptr_cap = nm + 1;
    }
    if (e->adj_cap < (int64_t)adj.size()) {
        dev_free(e, (void *)e->adj, (size_t)e->adj_cap * 4); e->adj = nullptr;
        TB_TRY(dev_alloc(e, &e->adj, adj.size())); e->adj_cap = (int64_t)adj.size();
    }
    TB_CUDA(e, cudaMemcpy(e->row_ptr, ptr.data(), ((size_t)nm + 1) * 4, cudaMemcpyHostToDevice));
    TB_CUDA(e, cudaMemcpy(e->adj, adj.data(), adj.size() * 4, cudaMemcpyHostToDevice));
    e->csr_nm = nm; e->csr_ns = ns; e->extra_adj.clear();
    e->graph_dirty = true;
    return TB_OK;
}

// (re)builds the device list of actuated springs; outside any stream capture (it allocates)
int prepare_actuated(Engine *e) {
    if (!e->act_dirty) return TB_OK;
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    dev_free(e, (void *)e->act_idx, (size_t)e->n_act * 4); e->act_idx = nullptr; e->n_act = 0;
    e->h_act.clear();
    for (int64_t s = 0; s < (int64_t)e->h_type.size(); s++) if (e->h_type[(size_t)s] >= TB_ACTUATED_EXPAND) e->h_act.push_back((int32_t)s);
    e->n_act = (int64_t)e->h_act.size();
    if (e->n_act) {
        TB_TRY(dev_alloc(e, &e->act_idx, (size_t)e->n_act));
        TB_CUDA(e, cudaMemcpy(e->act_idx, e->h_act.data(), (size_t)e->n_act * 4, cudaMemcpyHostToDevice));
    }
    e->act_dirty = false;
    e->graph_dirty = true;
    return TB_OK;
}

// the tiled kernel covers both integrators and takes the magnet pass's forces like an uploaded external force; a
// partitioned scene (ghost masses) with magnets is not supported (the pass would need the peers' positions before the
// step's halo exchange)
static bool use_tiled(Engine *e) {
    return !(e->flags & TB_KERNEL_CSR) && !(e->any_magnet && e->any_ghost) && tiled_supported(e);
}

// enqueue ONE step (capturable): [advance_rest] + [occupancy grid + magnet pass] + fused spring-gather/mass-update kernel
static int enqueue_step(Engine *e, bool use_ext, bool write_acc) {
    if (e->any_actuated) {
        if (!e->l_max || !e->l_min || !e->rate) return fail(e, TB_ERR_INVALID, "actuated springs need l_max, l_min and rate");
        if (!use_tiled(e)) {     // the tiled kernels advance the records of their own tiles
            advance_rest_kernel<<<blocks_for(e->n_act), 256, 0, e->stream>>>(
                e->rest, e->type, e->l_max, e->l_min, e->rate, e->left, e->right, e->mflags, e->G.dt, e->act_idx, e->n_act);
            e->launches++;
        }
    }
    if (e->flags & TB_INTEGRATOR_RK2) {
        // two force evaluations per step (src/sim.cu:1778-1799): half step into the other buffer, then the full step from the
        // backup -- which is what the first buffer still holds when the second launch overwrites it, slot by slot
        if (e->any_magnet || e->any_ghost) return fail(e, TB_ERR_UNSUPPORTED, "the RK2 integrator is not built for magnetic masses or partitioned systems");
        for (int stage = 1; stage <= 2; stage++) {
            if (stage == 2 && e->any_actuated && !use_tiled(e)) {     // computeSpringForces runs (and advances _rest) in both halves
                advance_rest_kernel<<<blocks_for(e->n_act), 256, 0, e->stream>>>(
                    e->rest, e->type, e->l_max, e->l_min, e->rate, e->left, e->right, e->mflags, e->G.dt, e->act_idx, e->n_act);
                e->launches++;
            }
            int rc = use_tiled(e) ? launch_tiled_step(e, use_ext && stage == 1, write_acc && stage == 2, stage)
                                  : launch_csr_step(e, use_ext && stage == 1, write_acc && stage == 2, stage);
            if (rc != TB_OK) return rc;
            e->launches++;
            e->cur ^= 1;
        }
        return TB_OK;
    }
    if (e->any_magnet) TB_TRY(launch_magnet_pass(e, use_ext, use_tiled(e)));   // recomputeOG + magnet forces, src/sim.cu:1775,1298
    int rc = use_tiled(e) ? launch_tiled_step(e, use_ext, write_acc) : launch_csr_step(e, use_ext, write_acc);
    if (rc != TB_OK) return rc;
    e->launches++;
    e->cur ^= 1;
    return TB_OK;
}

static const int GRAPH_STEPS = 50;  // even: a graph launch leaves the ping-pong parity unchanged

static int build_graph(Engine *e, int parity) {
    if (e->graph[parity]) { cudaGraphExecDestroy(e->graph[parity]); e->graph[parity] = nullptr; }
    int saved_cur = e->cur;
    int64_t saved_launches = e->launches;
    e->cur = parity;
    cudaGraph_t g = nullptr;
    TB_CUDA(e, cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
    int rc = TB_OK;
    for (int s = 0; s < GRAPH_STEPS && rc == TB_OK; s++) rc = enqueue_step(e, false, false);
    cudaError_t err = cudaStreamEndCapture(e->stream, &g);
    e->cur = saved_cur;
    e->launches = saved_launches;
    if (rc != TB_OK) { if (g) cudaGraphDestroy(g); return rc; }
    if (err != cudaSuccess) return fail(e, TB_ERR_CUDA, "stream capture failed: %s", cudaGetErrorString(err));
    err = cudaGraphInstantiate(&e->graph[parity], g, 0);
    cudaGraphDestroy(g);
    if (err != cudaSuccess) return fail(e, TB_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(err));
    e->graph_steps = GRAPH_STEPS;
    return TB_OK;
}

}  // namespace tb

extern "C" {

int tb_run_async(tb_engine *e, int64_t nsteps, double t0) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "No masses have been added. Please add masses before starting the simulation.");  // src/sim.cu:1552-1554
    if (nsteps < 0) return fail(e, TB_ERR_INVALID, "tb_run_async: negative step count");
    if ((e->flags & TB_INTEGRATOR_RK2) && e->any_local && !e->rk2_bvel) TB_TRY(dev_alloc(e, &e->rk2_bvel, (size_t)e->m_alloc() * 3));
    if (e->any_actuated) {
        if (!e->l_max || !e->l_min || !e->rate) return fail(e, TB_ERR_INVALID, "actuated springs need l_max, l_min and rate");
        TB_TRY(prepare_actuated(e));   // before tiled_prepare: the tile builder records where their rest lengths live
    }
    if (use_tiled(e)) TB_TRY(tiled_prepare(e));       // may discover that the graph does not fit the tiled format
    if (!use_tiled(e) && e->canon_stale) TB_TRY(tiled_to_canonical(e));
    if (!use_tiled(e)) TB_TRY(ensure_csr(e));
    if (nsteps == 0) return TB_OK;                    // validation + device layout brought up to date, nothing launched
    if (!use_tiled(e) && e->any_ghost) return fail(e, TB_ERR_UNSUPPORTED, "ghost masses (halo exchange) need the tiled step kernel, which does not support this scene");
    TB_CUDA(e, cudaEventRecord(e->ev0, e->stream));
    set_t_kernel<<<1, 1, 0, e->stream>>>(e->t_dev + e->cur, t0, e->step_dev + e->cur, (long long)e->steps_done);
    int64_t left_steps = nsteps;
    if (e->ext_pending) {  // the one-shot external force acts on exactly the first step (src/sim.cu:1300,1365)
        TB_TRY(enqueue_step(e, true, left_steps == 1));
        clear_ext_kernel<<<(unsigned)((e->nm + 255) / 256), 256, 0, e->stream>>>(e->ext, e->mflags, e->nm);
        e->ext_pending = false;
        left_steps--;
    }
    // Small systems (<= 8 tiles): every remaining step in one cluster launch, state resident in shared memory
    while (left_steps > 0 && use_tiled(e) && tiled_resident_ok(e)) {
        const int chunk = (int)std::min<int64_t>(left_steps, (int64_t)1 << 20);
        TB_TRY(launch_tiled_resident(e, chunk, chunk == left_steps));
        e->launches++;
        e->cur ^= (chunk & 1);
        left_steps -= chunk;
    }
    bool graphs = !(e->flags & TB_NO_GRAPH);
    if (graphs && left_steps - 1 >= GRAPH_STEPS) {
        if (e->graph_dirty) {
            for (int p = 0; p < 2; p++) if (e->graph[p]) { cudaGraphExecDestroy(e->graph[p]); e->graph[p] = nullptr; }
            e->graph_dirty = false;
        }
        if (!e->graph[e->cur]) TB_TRY(build_graph(e, e->cur));
        while (left_steps - 1 >= GRAPH_STEPS) {
            TB_CUDA(e, cudaGraphLaunch(e->graph[e->cur], e->stream));
            e->launches += (int64_t)GRAPH_STEPS * (((e->flags & TB_INTEGRATOR_RK2) ? 2 : 1) * (1 + ((e->any_actuated && !use_tiled(e)) ? 1 : 0)) + (e->any_magnet ? 5 : 0));
            left_steps -= GRAPH_STEPS;
        }
    }
    while (left_steps > 0) { TB_TRY(enqueue_step(e, false, left_steps == 1)); left_steps--; }
    TB_CUDA(e, cudaEventRecord(e->ev1, e->stream));
    TB_CUDA(e, cudaGetLastError());
    e->steps_done += nsteps;
    {
        const int slot = (int)(e->run_count++ % Engine::RUN_RING);
        TB_CUDA(e, cudaEventRecord(e->run_ev[slot], e->stream));
        e->run_ev_steps[slot] = e->steps_done;
    }
    if (use_tiled(e)) e->canon_stale = true;
    return TB_OK;
}

int tb_sync(tb_engine *e) {
    E_CHECK(e);
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    TB_CUDA(e, cudaGetLastError());
    return tiled_halo_check(e);
}

int tb_wait_steps(tb_engine *e, int64_t steps) {
    E_CHECK(e);
    if (steps > e->steps_done) return fail(e, TB_ERR_INVALID, "tb_wait_steps: %lld steps were never enqueued (%lld so far)", (long long)steps, (long long)e->steps_done);
    // the earliest recorded run that reaches `steps`; runs older than the ring have completed if any ring entry has
    int best = -1;
    const int64_t live = std::min<int64_t>(e->run_count, Engine::RUN_RING);
    for (int64_t r = 0; r < live; r++) {
        const int slot = (int)((e->run_count - 1 - r) % Engine::RUN_RING);
        if (e->run_ev_steps[slot] >= steps) best = slot; else break;    // newest to oldest: step counts descend
    }
    if (best < 0) return TB_OK;                   // nothing enqueued yet
    TB_CUDA(e, cudaEventSynchronize(e->run_ev[best]));
    return TB_OK;
}

}  // extern "C"

// ---- download without any allocation: gather scratch in a file-scope device array ------------------------------------
// tb_get_mass_state stages its gather through e->stage_dev, which grows by cudaMalloc on demand.  Callers that must not
// allocate get this variant: the same gather kernel, its scratch in the array below.  Every GPU that loads the library
// holds these 1.75 MiB until the process exits.  ONE array per device: every call on that device shares it (calls that
// overlap on the GPU -- two engines of one device used from two threads -- race), and it keeps the previous call's
// contents; nothing in it has to start at zero, a call reads back only what it has written first.
namespace tb {
constexpr int64_t STATIC_MAX_MASSES = 65536;
constexpr size_t STATIC_IDX_BYTES = (size_t)STATIC_MAX_MASSES * 4;
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_IDX_BYTES + (size_t)STATIC_MAX_MASSES * 24];

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current.
static int static_scratch_ptr(Engine *e, unsigned char **out) {
    constexpr int MAXD = 64;
    static std::mutex mu;
    static unsigned char *cached[MAXD] = {};
    std::lock_guard<std::mutex> lock(mu);
    int cur = -1;
    TB_CUDA(e, cudaGetDevice(&cur));
    if (cur != e->device) TB_CUDA(e, cudaSetDevice(e->device));
    if (e->device < MAXD && cached[e->device]) { *out = cached[e->device]; return TB_OK; }
    void *p = nullptr;
    TB_CUDA(e, cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) return fail(e, TB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
    if (e->device < MAXD) cached[e->device] = (unsigned char *)p;
    *out = (unsigned char *)p;
    return TB_OK;
}
}  // namespace tb

extern "C" {

int64_t tb_static_scratch_max_masses(void) { return tb::STATIC_MAX_MASSES; }

int tb_get_mass_state_static(tb_engine *e, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel) {
    E_CHECK(e);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_get_mass_state_static before tb_set_topology");
    if (count > STATIC_MAX_MASSES)
        return fail(e, TB_ERR_INVALID, "tb_get_mass_state_static: %lld masses exceed the static scratch (at most %lld per call; use "
                    "tb_get_mass_state, or several calls)", (long long)count, (long long)STATIC_MAX_MASSES);
    TB_TRY(check_range(e, e->nm, first, count, idx, "tb_get_mass_state_static"));
    if (count == 0 || (!pos && !vel)) return TB_OK;
    unsigned char *scratch = nullptr;
    TB_TRY(static_scratch_ptr(e, &scratch));
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (e->canon_stale) TB_TRY(tiled_to_canonical(e));       // kernel + sync, no allocation
    const int32_t *didx = nullptr;
    if (idx) {
        TB_CUDA(e, cudaMemcpyAsync(scratch, idx, (size_t)count * 4, cudaMemcpyHostToDevice, e->stream));
        didx = (const int32_t *)scratch;
    }
    double *tmp = (double *)(scratch + STATIC_IDX_BYTES);
    const double *src[2] = {e->pos[e->cur], e->vel[e->cur]};
    double *dst[2] = {pos, vel};
    for (int f = 0; f < 2; f++) {
        if (!dst[f]) continue;
        gather_vec3<<<blocks_for(count), 256, 0, e->stream>>>(tmp, src[f], e->nm, didx, first, count);
        TB_CUDA(e, cudaGetLastError());
        TB_CUDA(e, cudaMemcpyAsync(dst[f], tmp, (size_t)count * 24, cudaMemcpyDeviceToHost, e->stream));
        TB_CUDA(e, cudaStreamSynchronize(e->stream));        // `tmp` is reused by the next field
    }
    return TB_OK;
}

int tb_steps_done(tb_engine *e, int64_t *steps) {
    if (!e || !steps) return TB_ERR_INVALID;
    *steps = e->steps_done;
    return TB_OK;
}

int tb_last_run_ms(tb_engine *e, double *ms) {
    E_CHECK(e);
    if (!ms) return TB_ERR_INVALID;
    float f = 0;
    if (e->run_count == 0) { *ms = 0.0; return TB_OK; }   // nothing timed yet: the events were never recorded (cudaEventElapsedTime would fail)
    TB_CUDA(e, cudaEventSynchronize(e->ev1));
    TB_CUDA(e, cudaEventElapsedTime(&f, e->ev0, e->ev1));
    *ms = f;
    return TB_OK;
}

int tb_get_info(tb_engine *e, tb_info *info) {
    if (!e || !info) return TB_ERR_INVALID;
    memset(info, 0, sizeof *info);
    info->nm = e->nm; info->ns = e->ns;
    info->kernel = use_tiled(e) ? 1 : 0;
    info->launches = e->launches;
    info->device_bytes = e->device_bytes;
    // 104 B per mass (read pos,vel,m; write pos,vel) + per spring 24 B (+8 damping, +12 actuated)
    double bs = 24.0 + (e->any_damping ? 8.0 : 0.0) + ((e->any_active || e->any_actuated) ? 12.0 : 0.0);
    info->algorithmic_bytes_per_step = 104.0 * (double)e->nm + bs * (double)(e->ns - e->n_removed);
    info->sm_count = e->sm_count;
    info->graph_steps = (e->flags & TB_NO_GRAPH) ? 0 : GRAPH_STEPS;
    info->ns_removed = e->n_removed;
    tiled_info(e, info);
    return TB_OK;
}

int tb_export_positions_f32_dev(tb_engine *e, const float **out_dev) {
    E_CHECK(e);
    if (!out_dev || e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_export_positions_f32_dev: nothing to export");
    if (e->canon_stale) TB_TRY(tiled_to_canonical(e));
    if (!e->export_f32) TB_TRY(dev_alloc(e, &e->export_f32, (size_t)e->nm * 3));
    export_f32_kernel<<<(unsigned)((e->nm + 255) / 256), 256, 0, e->stream>>>(e->export_f32, e->pos[e->cur], e->nm);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    *out_dev = e->export_f32;
    return TB_OK;
}

int tb_export_indices_dev(tb_engine *e, const uint32_t **out_dev) {
    E_CHECK(e);
    if (!out_dev || e->ns == 0) return fail(e, TB_ERR_INVALID, "tb_export_indices_dev: nothing to export");
    if (!e->export_idx) TB_TRY(dev_alloc(e, &e->export_idx, (size_t)e->ns * 2));
    export_indices_kernel<<<blocks_for(e->ns), 256, 0, e->stream>>>(e->export_idx, e->left, e->right, e->mflags, e->removed_dev, e->ns);
    TB_CUDA(e, cudaGetLastError());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    *out_dev = e->export_idx;
    return TB_OK;
}

int tb_read_device(tb_engine *e, void *dst_host, const void *src_dev, size_t bytes) {
    E_CHECK(e);
    if (!dst_host || !src_dev) return fail(e, TB_ERR_INVALID, "tb_read_device: null pointer");
    TB_CUDA(e, cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, e->stream));
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return TB_OK;
}

}  // extern "C"

extern "C" int tb_alloc_host(tb_engine *e, size_t bytes, void **out) {
    E_CHECK(e);
    if (!out) return TB_ERR_INVALID;
    *out = nullptr;
    void *p = nullptr;
    TB_CUDA(e, cudaMallocHost(&p, bytes ? bytes : 1));
    e->pinned.push_back(p);
    *out = p;
    return TB_OK;
}
