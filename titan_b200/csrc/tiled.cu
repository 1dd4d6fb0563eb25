// tiled.cu -- the fast-path step kernel: one CTA per TILE of masses, everything a tile needs staged in
// shared memory, forces gathered per mass in ascending spring order in registers (deterministic), the
// whole mass update fused behind it.  One launch == one time step (Jacobi ping-pong, as csr_step.cu).
//
// Layout in HBM (built once per topology by the host tile builder below, refreshed by small gather kernels
// when only parameters / state change):
//   masses are re-ordered tile-major: slot = tile*TM + t.  Tiles come from a recursive coordinate bisection of
//   the current positions (bricks for lattices), or from a caller hint (tb_set_tile_hint).
//   pos[2], vel[2] : 3 planes of n_slots doubles, ping-pong           m, flags : per slot
//   per tile:  halo[]  : slots of the foreign masses its springs reach (gathered into smem every step)
//              prm[]   : (k, rest) pairs of every spring touching the tile, one record per spring and tile,
//                        contiguous -> streamed into smem with one bulk copy; [dmp[], omg[] when present]
//              ell[]   : per own mass, one u32 per incident spring in ascending spring order (ELL, TM-strided
//                        so a warp reads consecutive words): neighbour's local index | record slot | side | type
// Per spring this reads 16 B of parameters (+ a second copy for springs cut by a tile face) and 2 x 4 B of
// ELL entries, against the algorithmic 24 B; endpoint gathers hit shared memory.
// Replaces computeSpringForces + massForcesAndUpdate, src/sim.cu:1157-1200,1285-1367.
#include "engine.h"

#include <algorithm>
#include <cstring>
#include <numeric>

namespace tb {

#define TB_TRY(x) do { int _rc = (x); if (_rc != TB_OK) return _rc; } while (0)

constexpr int TM = 128;                 // masses (threads) per tile
constexpr int ELL_NB_BITS = 11;         // neighbour local index  (own: 0..TM-1, halo: TM..)
constexpr int ELL_REC_BITS = 15;        // record slot in the tile's parameter arrays
constexpr uint32_t ELL_SIDE = 1u << 26; // this mass is the spring's RIGHT endpoint
constexpr int ELL_TYPE_SHIFT = 27;      // 0 passive, 1 contract-then-expand, 2 expand-then-contract
constexpr uint32_t ELL_VALID = 1u << 31;
constexpr int MAX_DEG = 32;             // ELL rows are held in registers; more incident springs -> CSR kernel
constexpr int MAX_LOCAL = 1 << ELL_NB_BITS;
constexpr int MAX_REC = 1 << ELL_REC_BITS;
constexpr size_t SMEM_BUDGET = 100 * 1024;

struct TileMeta {
    int32_t ell_off;    // first ELL word of this tile (deg * TM words follow)
    int32_t deg;        // ELL rows
    int32_t halo_off, n_halo;
    int32_t rec_off, n_rec;   // rec_off is a multiple of 8 records (128 B)
};

struct TiledState {
    int64_t n_tiles = 0, n_slots = 0, n_rec = 0, n_ell = 0, n_halo = 0;
    int max_halo = 0, max_rec = 0, max_deg = 0;
    size_t smem_bytes = 0;
    bool has_damping = false, has_active = false;
    // device
    double *pos[2] = {nullptr, nullptr}, *vel[2] = {nullptr, nullptr}, *m = nullptr;
    uint8_t *flags = nullptr;
    int32_t *slot_to_api = nullptr, *halo = nullptr, *rec_spring = nullptr;
    uint32_t *ell = nullptr;
    double *prm = nullptr, *dmp = nullptr, *omg = nullptr;   // prm: 2 doubles per record
    TileMeta *meta = nullptr;
    // host
    std::vector<int32_t> h_slot_to_api, h_api_to_slot;
    bool supported_checked = false, supported = false;
    std::string why_not;
};

// ---------------------------------------------------------------- tile assignment (host) -----------------
// Recursive coordinate bisection: split the mass set at a multiple of TM along its longest axis until one
// tile is left.  Ties (lattice planes) are broken by API index, so the result is deterministic.
static void rcb(std::vector<int32_t> &order, const double *pos, int64_t nm, int64_t lo, int64_t hi, int64_t tiles) {
    if (tiles <= 1) return;
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (int64_t q = lo; q < hi; q++)
        for (int c = 0; c < 3; c++) {
            double v = pos[c * nm + (int64_t)order[q]];
            if (v < mn[c]) mn[c] = v;
            if (v > mx[c]) mx[c] = v;
        }
    int ax = 0;
    for (int c = 1; c < 3; c++) if (mx[c] - mn[c] > mx[ax] - mn[ax]) ax = c;
    int64_t left_tiles = tiles / 2;
    int64_t mid = lo + left_tiles * TM;
    if (mid >= hi) return;
    std::nth_element(order.begin() + lo, order.begin() + mid, order.begin() + hi, [&](int32_t a, int32_t b) {
        double va = pos[ax * nm + (int64_t)a], vb = pos[ax * nm + (int64_t)b];
        return va < vb || (va == vb && a < b);
    });
    rcb(order, pos, nm, lo, mid, left_tiles);
    rcb(order, pos, nm, mid, hi, tiles - left_tiles);
}

// Fills slot_to_api (n_tiles*TM entries, -1 = padding slot) from an RCB of the current positions (SoA, 3 planes
// of nm) or from the caller's hint.  Inside a tile masses are sorted by API index.
static void assign_tiles(Engine *e, const std::vector<double> &pos_soa, std::vector<int32_t> &slot_to_api) {
    int64_t nm = e->nm;
    std::vector<int32_t> order((size_t)nm);
    std::iota(order.begin(), order.end(), 0);
    slot_to_api.clear();
    if (!e->tile_hint.empty()) {
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return e->tile_hint[a] < e->tile_hint[b]; });
        int in_tile = 0;
        for (int64_t q = 0; q < nm; q++) {
            bool new_tile = q > 0 && (in_tile == TM || e->tile_hint[order[q]] != e->tile_hint[order[q - 1]]);
            if (new_tile) { slot_to_api.resize(slot_to_api.size() + (size_t)(TM - in_tile), -1); in_tile = 0; }
            slot_to_api.push_back(order[q]);
            in_tile++;
        }
        slot_to_api.resize(slot_to_api.size() + (size_t)(TM - in_tile), -1);
        return;
    }
    int64_t tiles = (nm + TM - 1) / TM;
    rcb(order, pos_soa.data(), nm, 0, nm, tiles);
    for (int64_t t = 0; t < tiles; t++)
        std::sort(order.begin() + t * TM, order.begin() + std::min<int64_t>(nm, (t + 1) * TM));
    slot_to_api.assign(order.begin(), order.end());
    slot_to_api.resize((size_t)(tiles * TM), -1);
}


// ---------------------------------------------------------------- structure build (host) -----------------
template <class T> static int upload_vec(Engine *e, T **dst, const std::vector<T> &src) {
    TB_TRY(dev_alloc(e, dst, src.size()));
    if (!src.empty()) TB_CUDA(e, cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return TB_OK;
}
template <class T> static void release(Engine *e, T *&p, size_t count) { dev_free(e, (void *)p, count * sizeof(T)); p = nullptr; }

static void free_structure(Engine *e, TiledState *S) {
    for (int p = 0; p < 2; p++) { release(e, S->pos[p], (size_t)S->n_slots * 3); release(e, S->vel[p], (size_t)S->n_slots * 3); }
    release(e, S->m, (size_t)S->n_slots); release(e, S->flags, (size_t)S->n_slots);
    release(e, S->slot_to_api, (size_t)S->n_slots); release(e, S->halo, (size_t)S->n_halo);
    release(e, S->rec_spring, (size_t)S->n_rec); release(e, S->ell, (size_t)S->n_ell);
    release(e, S->prm, (size_t)S->n_rec * 2); release(e, S->dmp, (size_t)S->n_rec); release(e, S->omg, (size_t)S->n_rec);
    if (S->meta) { cudaFree(S->meta); S->meta = nullptr; }
    S->n_tiles = S->n_slots = S->n_rec = S->n_ell = S->n_halo = 0;
}

static size_t smem_need(int max_halo, int max_rec, bool damping, bool active) {
    size_t nl = (size_t)((TM + max_halo + 1) & ~1);
    size_t b = nl * 24 * (damping ? 2 : 1) + (size_t)max_rec * 16;
    if (damping) b += (size_t)max_rec * 8;
    if (active) b += (size_t)max_rec * 8;
    return b + 16;
}

// Builds tiles, ELL rows, halo lists and record tables on the host and uploads them.  Returns TB_OK with
// S->supported == false (and a reason) when the graph does not fit the tiled format.
static int build_structure(Engine *e, TiledState *S) {
    free_structure(e, S);
    S->supported = false;
    const int64_t nm = e->nm;
    // current positions (SoA) drive the bisection
    std::vector<double> pos((size_t)nm * 3);
    TB_CUDA(e, cudaMemcpy(pos.data(), e->pos[e->cur], (size_t)nm * 24, cudaMemcpyDeviceToHost));
    assign_tiles(e, pos, S->h_slot_to_api);
    std::vector<double>().swap(pos);
    const int64_t n_slots = (int64_t)S->h_slot_to_api.size(), n_tiles = n_slots / TM;
    S->h_api_to_slot.assign((size_t)nm, -1);
    for (int64_t s = 0; s < n_slots; s++) if (S->h_slot_to_api[(size_t)s] >= 0) S->h_api_to_slot[(size_t)S->h_slot_to_api[(size_t)s]] = (int32_t)s;

    std::vector<TileMeta> meta((size_t)n_tiles);
    std::vector<uint32_t> ell;
    std::vector<int32_t> halo, rec_spring;
    std::vector<int32_t> halo_stamp((size_t)nm, -1), halo_local((size_t)nm, 0);
    std::vector<int32_t> rec_stamp((size_t)e->ns, -1), rec_local((size_t)e->ns, 0);
    struct Half { int32_t s; int32_t o; uint8_t side; };
    std::vector<std::vector<Half>> rows(TM);
    int max_halo = 0, max_rec = 0, max_deg = 0;
    const std::vector<int32_t> &rp = e->h_row_ptr, &adj = e->h_adj;
    for (int64_t T = 0; T < n_tiles; T++) {
        int deg = 0;
        for (int t = 0; t < TM; t++) {
            rows[t].clear();
            int32_t i = S->h_slot_to_api[(size_t)(T * TM + t)];
            if (i < 0 || !(e->h_flags[(size_t)i] & F_VALID)) continue;
            for (int32_t q = rp[(size_t)i]; q < rp[(size_t)i + 1]; q++) {
                int32_t a = adj[(size_t)q], s = a >> 1;
                uint8_t side = (uint8_t)(a & 1);
                int32_t o = side ? e->h_left[(size_t)s] : e->h_right[(size_t)s];
                if (!(e->h_flags[(size_t)o] & F_VALID)) continue;   // spring skipped, src/sim.cu:1163
                rows[t].push_back({s, o, side});
            }
            deg = std::max(deg, (int)rows[t].size());
        }
        if (deg > MAX_DEG) { S->why_not = "a mass has more incident springs than the ELL format holds"; return TB_OK; }
        TileMeta &M = meta[(size_t)T];
        M.ell_off = (int32_t)ell.size(); M.deg = deg;
        M.halo_off = (int32_t)halo.size();
        while (rec_spring.size() % 8) rec_spring.push_back(-1);   // 128-byte aligned record blocks
        M.rec_off = (int32_t)rec_spring.size();
        if ((int64_t)ell.size() + (int64_t)deg * TM >= ((int64_t)1 << 31) || rec_spring.size() >= ((size_t)1 << 31) - MAX_REC) {
            S->why_not = "tiled tables exceed int32 offsets"; return TB_OK;
        }
        ell.resize(ell.size() + (size_t)deg * TM, 0u);
        int n_halo = 0, n_rec = 0;
        for (int j = 0; j < deg; j++) for (int t = 0; t < TM; t++) {   // j-major walk: record slots of one row are consecutive
            if (j >= (int)rows[t].size()) continue;
            const Half &h = rows[t][(size_t)j];
            int32_t oslot = S->h_api_to_slot[(size_t)h.o];
            int nb;
            if (oslot / TM == T) nb = oslot % TM;
            else {
                if (halo_stamp[(size_t)h.o] != (int32_t)T) { halo_stamp[(size_t)h.o] = (int32_t)T; halo_local[(size_t)h.o] = n_halo++; halo.push_back(oslot); }
                nb = TM + halo_local[(size_t)h.o];
            }
            if (rec_stamp[(size_t)h.s] != (int32_t)T) { rec_stamp[(size_t)h.s] = (int32_t)T; rec_local[(size_t)h.s] = n_rec++; rec_spring.push_back(h.s); }
            int rec = rec_local[(size_t)h.s];
            if (nb >= MAX_LOCAL || rec >= MAX_REC) { S->why_not = "a tile reaches too many foreign masses or springs"; return TB_OK; }
            uint32_t ty = 0;
            if (!e->h_type.empty()) { uint8_t x = e->h_type[(size_t)h.s]; ty = (x == 2) ? 1u : (x == 3) ? 2u : 0u; }
            ell[(size_t)M.ell_off + (size_t)j * TM + (size_t)t] =
                (uint32_t)nb | ((uint32_t)rec << ELL_NB_BITS) | (h.side ? ELL_SIDE : 0u) | (ty << ELL_TYPE_SHIFT) | ELL_VALID;
        }
        M.n_halo = n_halo; M.n_rec = n_rec;
        max_halo = std::max(max_halo, n_halo); max_rec = std::max(max_rec, n_rec); max_deg = std::max(max_deg, deg);
    }
    while (rec_spring.size() % 8) rec_spring.push_back(-1);
    S->has_damping = e->any_damping; S->has_active = e->any_active;
    S->smem_bytes = smem_need(max_halo, max_rec, S->has_damping || (e->flags & TB_MATH_STRICT), S->has_active);
    if (S->smem_bytes > SMEM_BUDGET) { S->why_not = "a tile's working set exceeds the shared-memory budget"; return TB_OK; }
    // padding: the ELL rows of all tiles together must not dwarf the real incidence (skewed degree distributions)
    if ((double)ell.size() > 1.6 * (double)std::max<int64_t>(1, rp[(size_t)nm]) + 64.0 * TM) { S->why_not = "ELL padding too large for this degree distribution"; return TB_OK; }

    S->n_tiles = n_tiles; S->n_slots = n_slots; S->n_rec = (int64_t)rec_spring.size(); S->n_ell = (int64_t)ell.size();
    S->n_halo = (int64_t)halo.size(); S->max_halo = max_halo; S->max_rec = max_rec; S->max_deg = max_deg;
    for (int p = 0; p < 2; p++) { TB_TRY(dev_alloc(e, &S->pos[p], (size_t)n_slots * 3)); TB_TRY(dev_alloc(e, &S->vel[p], (size_t)n_slots * 3)); }
    TB_TRY(dev_alloc(e, &S->m, (size_t)n_slots)); TB_TRY(dev_alloc(e, &S->flags, (size_t)n_slots));
    TB_TRY(upload_vec(e, &S->slot_to_api, S->h_slot_to_api));
    TB_TRY(upload_vec(e, &S->halo, halo)); TB_TRY(upload_vec(e, &S->rec_spring, rec_spring)); TB_TRY(upload_vec(e, &S->ell, ell));
    TB_TRY(dev_alloc(e, &S->prm, (size_t)S->n_rec * 2));
    if (S->has_damping) TB_TRY(dev_alloc(e, &S->dmp, (size_t)S->n_rec));
    if (S->has_active) TB_TRY(dev_alloc(e, &S->omg, (size_t)S->n_rec));
    TB_CUDA(e, cudaMalloc((void **)&S->meta, (size_t)n_tiles * sizeof(TileMeta)));
    TB_CUDA(e, cudaMemcpy(S->meta, meta.data(), (size_t)n_tiles * sizeof(TileMeta), cudaMemcpyHostToDevice));
    S->supported = true;
    return TB_OK;
}

// ---------------------------------------------------------------- refresh kernels -------------------------
__global__ void tiled_gather_state(double *pos_t, double *vel_t, double *m_t, const double *pos_c, const double *vel_c,
                                   const double *m_c, const int32_t *slot_to_api, int64_t n_slots, int64_t nm) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int32_t i = slot_to_api[s];
    for (int c = 0; c < 3; c++) {
        pos_t[c * n_slots + s] = i >= 0 ? pos_c[c * nm + i] : 0.0;
        vel_t[c * n_slots + s] = i >= 0 ? vel_c[c * nm + i] : 0.0;
    }
    m_t[s] = i >= 0 ? m_c[i] : 1.0;
}
__global__ void tiled_scatter_state(double *pos_c, double *vel_c, const double *pos_t, const double *vel_t,
                                    const int32_t *slot_to_api, int64_t n_slots, int64_t nm) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int32_t i = slot_to_api[s];
    if (i < 0) return;
    for (int c = 0; c < 3; c++) { pos_c[c * nm + i] = pos_t[c * n_slots + s]; vel_c[c * nm + i] = vel_t[c * n_slots + s]; }
}
__global__ void tiled_gather_flags(uint8_t *flags_t, const uint8_t *flags_c, const int32_t *slot_to_api, int64_t n_slots) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int32_t i = slot_to_api[s];
    flags_t[s] = i >= 0 ? flags_c[i] : (uint8_t)0;
}
__global__ void tiled_gather_params(double *prm, double *dmp, double *omg, const int32_t *rec_spring, const double *k,
                                    const double *rest, const double *damping, const double *omega, int64_t n_rec) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rec) return;
    int32_t s = rec_spring[r];
    prm[2 * r] = s >= 0 ? k[s] : 0.0;
    prm[2 * r + 1] = s >= 0 ? rest[s] : 0.0;
    if (dmp) dmp[r] = (s >= 0 && damping) ? damping[s] : 0.0;
    if (omg) omg[r] = (s >= 0 && omega) ? omega[s] : 0.0;
}


// ---------------------------------------------------------------- the step kernel -------------------------
struct TiledArgs {
    int64_t n_slots;
    int32_t nl;                      // smem stride of one coordinate plane (TM + max_halo, even)
    int32_t max_rec;
    const double *pos_in, *vel_in;
    double *pos_out, *vel_out;
    const double *m;
    const uint8_t *flags;
    const int32_t *slot_to_api, *halo;
    const uint32_t *ell;
    const double *prm, *dmp, *omg;
    const TileMeta *meta;
    double *acc;                     // canonical (API-ordered) acceleration, stride X.nm
    const double *t_in;
    double *t_out;
};

template <class P, bool DAMPING, bool ACTIVE, bool WRITE_ACC>
__global__ void __launch_bounds__(TM) tiled_step_kernel(TiledArgs A, Globals G, MassExtras X) {
    extern __shared__ __align__(16) double smem[];
    const int nl = A.nl;
    double *px = smem, *py = px + nl, *pz = py + nl;
    double *vx = pz + nl, *vy = vx + nl, *vz = vy + nl;               // only when staging velocities
    constexpr bool STAGE_VEL = DAMPING || P::strict;
    double2 *prm = reinterpret_cast<double2 *>(STAGE_VEL ? vz + nl : pz + nl);
    double *dmp = reinterpret_cast<double *>(prm + A.max_rec);
    double *omg = DAMPING ? dmp + A.max_rec : dmp;

    const TileMeta M = A.meta[blockIdx.x];
    const int t = threadIdx.x;
    const int64_t n = A.n_slots, slot = (int64_t)blockIdx.x * TM + t;
    const double t_sim = *A.t_in;
    if (blockIdx.x == 0 && t == 0) *A.t_out = __dadd_rn(t_sim, G.dt);   // T += dt, src/sim.cu:1811

    // ---- phase A: everything the tile needs, all loads in flight together ------------------------------
    uint32_t ent[MAX_DEG];
#pragma unroll
    for (int j = 0; j < MAX_DEG; j++) ent[j] = (j < M.deg) ? A.ell[(size_t)M.ell_off + (size_t)j * TM + t] : 0u;
    {
        const double2 *src = reinterpret_cast<const double2 *>(A.prm) + M.rec_off;
        for (int r = t; r < M.n_rec; r += TM) prm[r] = src[r];
        if (DAMPING) for (int r = t; r < M.n_rec; r += TM) dmp[r] = A.dmp[M.rec_off + r];
        if (ACTIVE) for (int r = t; r < M.n_rec; r += TM) omg[r] = A.omg[M.rec_off + r];
    }
    V3 pos = {A.pos_in[slot], A.pos_in[n + slot], A.pos_in[2 * n + slot]};
    V3 vel = {A.vel_in[slot], A.vel_in[n + slot], A.vel_in[2 * n + slot]};
    const uint8_t fl = A.flags[slot];
    const double mass = A.m[slot];
    px[t] = pos.x; py[t] = pos.y; pz[t] = pos.z;
    if (STAGE_VEL) { vx[t] = vel.x; vy[t] = vel.y; vz[t] = vel.z; }
    for (int h = t; h < M.n_halo; h += TM) {
        int64_t o = A.halo[M.halo_off + h];
        px[TM + h] = A.pos_in[o]; py[TM + h] = A.pos_in[n + o]; pz[TM + h] = A.pos_in[2 * n + o];
        if (STAGE_VEL) { vx[TM + h] = A.vel_in[o]; vy[TM + h] = A.vel_in[n + o]; vz[TM + h] = A.vel_in[2 * n + o]; }
    }
    __syncthreads();

    if (!(fl & F_VALID) || (fl & F_FIXED)) {      // padding / deleted / fixed: frozen (src/sim.cu:1292-1295)
        A.pos_out[slot] = pos.x; A.pos_out[n + slot] = pos.y; A.pos_out[2 * n + slot] = pos.z;
        A.vel_out[slot] = vel.x; A.vel_out[n + slot] = vel.y; A.vel_out[2 * n + slot] = vel.z;
        return;
    }

    // ---- phase B: gather the incident springs in ascending spring order, registers only -----------------
    V3 force = {0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < MAX_DEG; j++) {
        if (j < M.deg) {
            const uint32_t en = ent[j];
            if (en & ELL_VALID) {
                const int nb = en & (MAX_LOCAL - 1), rec = (en >> ELL_NB_BITS) & (MAX_REC - 1);
                const V3 po = {px[nb], py[nb], pz[nb]};
                const double2 kr = prm[rec];
                SpringArgs sp;
                sp.k = kr.x;
                sp.rest_scaled = kr.y;
                sp.damping = DAMPING ? dmp[rec] : 0.0;
                if (ACTIVE) {
                    const int ty = (en >> ELL_TYPE_SHIFT) & 3;
                    if (ty) sp.rest_scaled = P::mul(kr.y, actuation_scale<P>(ty + 1, omg[rec], t_sim));
                }
                if (P::strict) {
                    const V3 vo = {vx[nb], vy[nb], vz[nb]};
                    if (en & ELL_SIDE) acc3<P>(force, spring_force<P, DAMPING>(po, pos, vo, vel, sp));
                    else { V3 f = spring_force<P, DAMPING>(pos, po, vel, vo, sp); acc3<P>(force, V3{-f.x, -f.y, -f.z}); }
                } else {
                    // force on this mass = (p_me - p_other) * coef whichever endpoint it is (see physics.cuh)
                    V3 vo = {0.0, 0.0, 0.0};
                    if (DAMPING) vo = V3{vx[nb], vy[nb], vz[nb]};
                    acc3<P>(force, spring_force<P, DAMPING>(po, pos, vo, vel, sp));
                }
            }
        }
    }

    // ---- phase C: mass update ----------------------------------------------------------------------------
    const bool extras = X.ext || X.drag || (fl & F_LOCAL) || WRITE_ACC;
    const int64_t i_api = extras ? A.slot_to_api[slot] : 0;
    V3 acc;
    mass_update<P>(G, X, i_api, fl & F_LOCAL, mass, pos, vel, force, acc);
    A.pos_out[slot] = pos.x; A.pos_out[n + slot] = pos.y; A.pos_out[2 * n + slot] = pos.z;
    A.vel_out[slot] = vel.x; A.vel_out[n + slot] = vel.y; A.vel_out[2 * n + slot] = vel.z;
    if (WRITE_ACC) { A.acc[i_api] = acc.x; A.acc[X.nm + i_api] = acc.y; A.acc[2 * X.nm + i_api] = acc.z; }
}

// ---------------------------------------------------------------- glue ------------------------------------
template <class P, bool D, bool AC, bool W>
static int launch_v(Engine *e, const TiledArgs &A, const MassExtras &X, size_t smem, unsigned grid) {
    static bool configured = false;     // per template instantiation
    if (!configured) {
        TB_CUDA(e, cudaFuncSetAttribute(tiled_step_kernel<P, D, AC, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BUDGET));
        configured = true;
    }
    tiled_step_kernel<P, D, AC, W><<<grid, TM, smem, e->stream>>>(A, e->G, X);
    return TB_OK;
}
template <class P, bool D, bool AC>
static int launch_w(Engine *e, const TiledArgs &A, const MassExtras &X, size_t smem, unsigned grid, bool w) {
    return w ? launch_v<P, D, AC, true>(e, A, X, smem, grid) : launch_v<P, D, AC, false>(e, A, X, smem, grid);
}
template <class P> static int launch_p(Engine *e, TiledState *S, const TiledArgs &A, const MassExtras &X, unsigned grid, bool w) {
    bool d = S->has_damping, a = S->has_active;
    size_t smem = S->smem_bytes;
    if (d && a) return launch_w<P, true, true>(e, A, X, smem, grid, w);
    if (d) return launch_w<P, true, false>(e, A, X, smem, grid, w);
    if (a) return launch_w<P, false, true>(e, A, X, smem, grid, w);
    return launch_w<P, false, false>(e, A, X, smem, grid, w);
}

bool tiled_supported(const Engine *e) {
    if (e->any_actuated || e->nm == 0) return false;   // ACTUATED_* springs mutate _rest: generic kernel
    const TiledState *S = e->tiled;
    if (S && !(e->tiled_stale & TS_STRUCT) && !S->supported) return false;
    return true;
}

int tiled_prepare(Engine *e) {
    if (!e->tiled) e->tiled = new TiledState();
    TiledState *S = e->tiled;
    uint32_t stale = e->tiled_stale;
    if (S->n_tiles > 0 && (S->has_damping != e->any_damping || S->has_active != e->any_active)) stale |= TS_STRUCT;
    if (S->n_tiles == 0 && !(stale & TS_STRUCT) && S->supported) stale |= TS_STRUCT;
    if (stale & TS_STRUCT) {
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
        if (e->canon_stale) TB_TRY(tiled_to_canonical(e));
        TB_TRY(build_structure(e, S));
        e->tiled_stale &= ~TS_STRUCT;
        e->graph_dirty = true;
        if (!S->supported) return TB_OK;
        stale |= TS_PARAMS | TS_STATE | TS_FLAGS;
    }
    if (!S->supported) return TB_OK;
    unsigned gs = (unsigned)((S->n_slots + 255) / 256), gr = (unsigned)((S->n_rec + 255) / 256);
    if (stale & TS_STATE) {
        if (e->canon_stale && !(stale & TS_STRUCT)) TB_TRY(tiled_to_canonical(e));  // partial upload landed in canonical
        tiled_gather_state<<<gs, 256, 0, e->stream>>>(S->pos[e->cur], S->vel[e->cur], S->m, e->pos[e->cur], e->vel[e->cur], e->m,
                                                       S->slot_to_api, S->n_slots, e->nm);
    }
    if (stale & TS_FLAGS) tiled_gather_flags<<<gs, 256, 0, e->stream>>>(S->flags, e->mflags, S->slot_to_api, S->n_slots);
    if (stale & TS_PARAMS)
        tiled_gather_params<<<gr, 256, 0, e->stream>>>(S->prm, S->dmp, S->omg, S->rec_spring, e->k, e->rest, e->damping, e->omega, S->n_rec);
    TB_CUDA(e, cudaGetLastError());
    e->tiled_stale = 0;
    return TB_OK;
}

int launch_tiled_step(Engine *e, bool use_ext, bool write_acc) {
    TiledState *S = e->tiled;
    if (!S || !S->supported) return fail(e, TB_ERR_INVALID, "tiled step requested without a tiled layout");
    TiledArgs A;
    A.n_slots = S->n_slots;
    A.nl = (TM + S->max_halo + 1) & ~1;
    A.max_rec = S->max_rec;
    A.pos_in = S->pos[e->cur]; A.vel_in = S->vel[e->cur];
    A.pos_out = S->pos[e->cur ^ 1]; A.vel_out = S->vel[e->cur ^ 1];
    A.m = S->m; A.flags = S->flags; A.slot_to_api = S->slot_to_api; A.halo = S->halo; A.ell = S->ell;
    A.prm = S->prm; A.dmp = S->dmp; A.omg = S->omg; A.meta = S->meta; A.acc = e->acc;
    A.t_in = e->t_dev + e->cur; A.t_out = e->t_dev + (e->cur ^ 1);
    MassExtras X;
    X.ext = (use_ext && e->ext) ? e->ext : nullptr;
    X.drag = e->drag;
    X.lc_ptr = e->any_local ? e->lc_ptr : nullptr;
    X.lc = e->lc;
    X.nm = e->nm;
    unsigned grid = (unsigned)S->n_tiles;
    int rc = (e->flags & TB_MATH_STRICT) ? launch_p<Strict>(e, S, A, X, grid, write_acc) : launch_p<Fast>(e, S, A, X, grid, write_acc);
    if (rc != TB_OK) return rc;
    TB_CUDA(e, cudaGetLastError());
    return TB_OK;
}

int tiled_to_canonical(Engine *e) {
    TiledState *S = e->tiled;
    if (S && S->supported && S->n_slots > 0 && e->canon_stale) {
        tiled_scatter_state<<<(unsigned)((S->n_slots + 255) / 256), 256, 0, e->stream>>>(
            e->pos[e->cur], e->vel[e->cur], S->pos[e->cur], S->vel[e->cur], S->slot_to_api, S->n_slots, e->nm);
        TB_CUDA(e, cudaGetLastError());
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    e->canon_stale = false;
    return TB_OK;
}

void tiled_destroy(Engine *e) {
    if (!e->tiled) return;
    free_structure(e, e->tiled);
    delete e->tiled;
    e->tiled = nullptr;
}

int tiled_info(const Engine *e, int32_t *n_tiles) {
    *n_tiles = (e->tiled && e->tiled->supported) ? (int32_t)e->tiled->n_tiles : 0;
    return TB_OK;
}

}  // namespace tb
