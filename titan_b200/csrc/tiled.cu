// tiled.cu -- the fast-path step kernel: PERSISTENT CTAs stream TILES of masses through a double-buffered
// shared-memory pipeline fed by bulk async copies (cp.async.bulk + mbarrier, SASS UBLKCP) and per-element
// cp.async gathers (LDGSTS) for the halo, so the next tile's bytes are in flight while this tile computes.
// One launch == one time step (Jacobi ping-pong between two state buffers, like csr_step.cu), many launches
// per CUDA graph.  Replaces computeSpringForces + massForcesAndUpdate, src/sim.cu:1157-1200,1285-1367.
//
// Layout in HBM (built per topology by the host tile builder below; refreshed by small kernels when only
// parameters / state / flags change):
//   masses are re-ordered tile-major: slot = tile*TM + t.  Tiles come from the caller's hint (lattice bricks,
//   tb_set_tile_hint) or from a recursive coordinate bisection of the current positions.
//   state[2]  : per tile one contiguous block [px py pz vx vy vz][TM] of doubles, ping-pong
//   blob      : per tile one contiguous block  header | flags[TM] | m[TM] | k[] | rest[] | dmp[] | omg[] | byte map
//               -- one record per spring touching the tile -- fetched with ONE bulk copy
//   halo      : per tile the state indices of the foreign masses its springs reach (bulk-copied two tiles
//               ahead, then used as the address list of the cp.async position gather one tile ahead)
//   ell       : per tile, per own mass, one u32 per incident spring in ascending spring order (ELL, TM-strided
//               so a warp reads consecutive words): neighbour's local index | record slot | side | type.
//               Identical blocks are stored once (content dedupe): congruent lattice bricks share their ELL
//               block, which then stays L2-resident instead of costing 8 B per spring of HBM traffic.
//   meta      : 32 B per tile (offsets and counts), prefetched through a small shared-memory ring
// Forces are gathered per mass in registers; with TPM threads per mass each thread takes every TPM-th ELL row
// and the partial sums are combined in a fixed order (deterministic).  TPM == 1 reproduces the oracle's
// ascending-spring-index summation exactly (TB_MATH_STRICT).
#include "engine.h"

#include <cooperative_groups.h>
#include <type_traits>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <chrono>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <unordered_map>

#include <unistd.h>

namespace tb {


constexpr int TM = 128;                 // masses per tile
constexpr int MAX_DEVICES = 64;         // per-device launch configuration caches (beyond that: configured on every launch)
// ELL word: bits 3..11 neighbour local index (so `word & 0xff8` is its byte offset in a position array; own
// masses 0..TM-1, halo TM..), bits 16..27 record slot (`(word >> 12) & 0xfff0` = byte offset of the (k,rest)
// record), bits 28..29 actuation (0 passive, 1 contract-then-expand, 2 expand-then-contract), bit 30 this mass
// is the spring's RIGHT endpoint, bit 31 valid.
constexpr int ELL_NB_SHIFT = 3, ELL_REC_SHIFT = 16, ELL_TYPE_SHIFT = 28;
constexpr uint32_t ELL_SIDE = 1u << 30, ELL_VALID = 1u << 31;
constexpr int MAX_DEG = 32;             // ELL rows are held in registers; more incident springs -> CSR kernel
constexpr int MAX_LOCAL = 512;
constexpr int MAX_REC = 4096;
constexpr int CS_SMALL = 384, CS_BIG = MAX_LOCAL;   // shared-memory stride of one position component (see plan_smem)
constexpr int ELL_CHUNK = 4;            // springs evaluated together, branch-free, by one thread (ILP)
// Local indices [cs - N_DUMMY, cs) of every position array are DUMMY masses (far away, at rest, never written by a copy):
// padding incidence words point at one of them and at record 0 (k = 0), and so contribute exactly zero.  Sixteen of them,
// one per 8-byte bank pair, so that a padding lane never collides with the live lanes of its half-warp.
constexpr int N_DUMMY = 16;
constexpr uint32_t PAD_NB = 512 - N_DUMMY;   // provisional dummy index in a padding word (rebased once cs is known)

// Which of a mass's TPM threads takes which ELL rows, in chunks of ELL_CHUNK rows (blocks are padded with zero
// rows to a whole number of chunks).  Thread 0 also runs the mass update (worth about MASS_UPDATE_ROWS springs),
// so it takes fewer chunks: chunks [0, c0) go to thread 0, the rest in equal contiguous runs to threads 1..TPM-1.
constexpr int MASS_UPDATE_ROWS = 5;
__host__ __device__ constexpr int ell_chunks(int deg) { return (deg + ELL_CHUNK - 1) / ELL_CHUNK; }
__host__ __device__ constexpr int ell_chunks0(int n, int tpm) {          // chunks of thread 0 out of n
    if (tpm == 1) return n;
    int r0 = (n * ELL_CHUNK + MASS_UPDATE_ROWS) / tpm - MASS_UPDATE_ROWS;
    int c0 = r0 > 0 ? (r0 + ELL_CHUNK / 2) / ELL_CHUNK : 0;
    return c0 < n ? c0 : n;
}
__host__ __device__ constexpr int ell_chunks_q(int n, int tpm) {         // chunks per thread 1..TPM-1
    return tpm == 1 ? 0 : (n - ell_chunks0(n, tpm) + tpm - 2) / (tpm - 1);
}
__host__ __device__ constexpr int ell_words_per_thread(int tpm) {
    int m = 0;
    for (int n = 0; n <= ell_chunks(MAX_DEG); n++) {
        int a = ell_chunks0(n, tpm), b = ell_chunks_q(n, tpm);
        if (a > m) m = a;
        if (b > m) m = b;
    }
    return m * ELL_CHUNK;
}
constexpr size_t SMEM_LIMIT = 227 * 1024;
constexpr int HALO_RING = 3;
constexpr int MAX_OMEGA = 64;           // distinct actuation frequencies served from the per-CTA sine table
constexpr int BLOB_HDR = 16, BLOB_FLAGS = BLOB_HDR, BLOB_M = BLOB_FLAGS + TM, BLOB_REC = BLOB_M + TM * 8;
constexpr int STATE_BLOCK = 6 * TM;     // doubles per tile in a state buffer

struct __align__(16) TileMeta {
    uint32_t blob_off16;  // blob offset in 16-byte units
    uint32_t blob_bytes;  // multiple of 16
    uint32_t halo_off;    // first halo entry (multiple of 4 entries)
    uint32_t n_halo;
    uint32_t ell_off;     // first ELL word (deg rounded up to ELL_CHUNK rows, TM words each)
    uint32_t deg;         // ELL rows
    uint32_t n_rec;       // spring records in the blob
    uint32_t rec_off;     // first entry of rec_spring (refresh kernels only)
    uint32_t exp_off;     // first entry of the export table (halo exchange): exp_rows rows of TM entries
    uint32_t exp_rows;    // 0 = no mass of this tile is a ghost on a peer
    uint32_t tflags;      // TF_*
    uint32_t map_off16;   // dictionary mode: the tile's slot -> record byte map in rec_map, in units of 16 bytes
};
static_assert(sizeof(TileMeta) == 48, "TileMeta is copied as three 16-byte words");
enum : uint32_t { TF_BOUNDARY = 1, TF_ACTUATED = 2 };   // ACTUATED: records of this tile are advanced before it is staged
//   // the tile's springs reach ghost masses: wait for the peers before reading, signal after

// a parameter class of the dictionary modes: bit patterns of (k, rest, damping, omega) + actuation code / private tag
struct ClassKey { uint64_t a, b, c, d, t; bool operator==(const ClassKey &o) const { return a == o.a && b == o.b && c == o.c && d == o.d && t == o.t; } };
struct ClassKeyHash { size_t operator()(const ClassKey &x) const {
    uint64_t h = x.a * 0x9E3779B97F4A7C15ull; h ^= (x.b + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full; h ^= h >> 29;
    h ^= x.c * 0x165667B19E3779F9ull; h ^= (x.d + x.t) * 0x27D4EB2F165667C5ull; return (size_t)(h ^ (h >> 32)); } };

// Host copies of the tables the tiled kernel reads, kept so that single tiles can be rebuilt later (merge_tile).
struct HostTables {
    std::unordered_map<ClassKey, int32_t, ClassKeyHash> class_seen;   // class key -> dense id
    std::vector<int32_t> class_first;                            // class id -> first spring (the representative)
    std::vector<uint32_t> ell;                                   // deduplicated incidence blocks (without the prefetch tail)
    std::map<uint64_t, std::vector<uint32_t>> ell_index;         // content hash -> offsets of stored blocks
    int64_t unique_blocks = 0;
    std::vector<int32_t> halo, rec_spring, kslot_spring;
    std::vector<uint32_t> krec_off;                              // split dictionary: (first slot, slots) per tile in kslot_spring
    std::vector<uint8_t> rec_map;
    std::vector<int32_t> class_of, rep_of;                       // parameter classes (dictionary modes)
    std::vector<std::vector<int4>> act;                          // per tile: actuated records
    std::vector<int32_t> tile_slots;                             // per tile: record slots
    size_t blob_total = 0;
    size_t base_ell = 0, base_halo = 0, base_rec = 0, base_map = 0, base_blob = 0;   // sizes right after the last full build
};

// elements allocated for the device tables (>= what is in use: the ends are room for in-place edits)
struct TableCaps { int64_t tiles = 0, st_tiles = 0, halo = 0, ell = 0, rec = 0, rec_map = 0, krec = 0, cls = 0; size_t blob = 0; };

struct TiledState {
    HostTables H;
    TableCaps cap;
    int64_t n_halo_up = 0, n_ell_up = 0, n_rec_map_up = 0;   // host entries already on the device (halo / ell / rec_map carry a tail)
    uint32_t ell_pad = 0;                 // the padding word the free end of the ELL table is filled with
    int n_uploads = 0, n_patches = 0;     // whole uploads of the tables / edits written into their free ends
    bool exports_set = false;             // the tile table carries export fields (tb_halo_set_exports) since its last full build
    int build_mode = 1;                   // record mode of the last build (1 full dictionary, 2 split, 0 plain)
    int64_t n_tiles = 0, n_slots = 0, n_rec = 0, n_ell = 0, n_halo = 0;
    size_t blob_bytes = 0;
    int max_halo = 0, max_rec = 0, max_deg = 0, max_blob = 0;
    int64_t ell_unique_tiles = 0;
    bool has_damping = false, has_active = false, stage_vel = false;
    int tpm = 4, grid = 0, cs = CS_BIG, nst = 2;
    size_t smem_bytes = 0;
    // device
    double *st[2] = {nullptr, nullptr};   // tile-blocked state, ping-pong
    uint8_t *blob = nullptr;
    int32_t *slot_to_api = nullptr, *api_to_slot = nullptr, *halo = nullptr, *rec_spring = nullptr;
    // Dictionary mode (see build_structure): springs with bit-identical parameters share one record per tile.
    bool dict = false;                    // the blobs hold distinct records + a byte map (record slot -> record)
    bool dict_split = false;              // ... for (rest, damping, omega) only: the stiffness stays one value per record
                                          // slot (systems whose k is edited per spring, e.g. topology optimisation)
    int32_t *krec_spring = nullptr; int64_t n_krec = 0;   // split mode: record slot -> spring, all tiles
    uint32_t *krec_off = nullptr;                         // split mode: [n_tiles + 1] first slot of each tile in krec_spring
    // ACTUATED_* springs: per tile the records whose rest length the step kernel advances before it stages the tile
    // (x = byte offset of the rest slot inside the tile's blob, y = spring, z = 1 on the copy that also keeps the canonical
    // rest array current -- a cut spring has a record in two tiles, both advance theirs identically)
    uint32_t *act_tile_off = nullptr; int4 *act_rec = nullptr; int64_t n_act_rec = 0;
    int dict_failures = 0;                // parameter uploads that broke the classes; after 2 the mode stays off
    int split_failures = 0;               // the same for the split dictionary
    int64_t n_class_rep = 0;
    int32_t *class_rep = nullptr;         // [ns] a spring with the same parameters (the class representative)
    uint8_t *rec_map = nullptr; int64_t n_rec_map = 0;
    int *dict_bad = nullptr;              // set by dict_check_kernel
    int res_cluster = 0;                  // tiles per self-contained group (independent systems of <= 8 tiles each, aligned
                                          // in the tile order), 0 = none: what the resident kernel runs as one cluster
    int max_dict = 0;
    uint32_t *ell = nullptr;
    TileMeta *meta = nullptr;
    // host
    std::vector<int32_t> h_slot_to_api, h_api_to_slot;
    size_t n_api = 0;                       // entries of the device copy api_to_slot (made when a magnet pass needs it)
    std::vector<TileMeta> h_meta;
    bool supported = false;
    std::string why_not;
    // distinct actuation frequencies: sin(omega t) is evaluated once per CTA and step instead of once per spring
    double *omega_tab = nullptr; int n_omega = 0;
    // ---- halo exchange (partitioned simulation) ----
    int64_t n_active = 0;                 // tiles the step kernel processes (ghost-only tiles come last and are skipped)
    int64_t n_boundary = 0;               // tiles [0, n_boundary) reach ghosts
    bool halo_prepared = false, halo_connected = false;
    int n_peers = 0;
    unsigned long long *flags_dev = nullptr;          // [TB_MAX_PEERS] step counters released by the peers
    unsigned *done_dev = nullptr;                     // boundary-tile completion counter of the running step
    int *err_dev = nullptr;                           // set when a wait for a peer timed out
    double *peer_st[TB_MAX_PEERS][2] = {};            // the peers' state buffers (peer memory)
    unsigned long long *peer_flag[TB_MAX_PEERS] = {}; // this engine's word in each peer's flags array
    void *ipc_opened[TB_MAX_PEERS][3] = {};           // bases opened with cudaIpcOpenMemHandle (closed on destroy)
    int2 *exp_tab = nullptr; int64_t n_exp_tab = 0;   // per boundary tile: rows of TM (peer, remote offset) entries
    int4 *exp_flat = nullptr; int64_t n_exp_flat = 0; // (local state offset, peer, remote offset, 0) for tb_halo_push_now
};

static inline size_t up16(size_t x) { return (x + 15) & ~(size_t)15; }

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

// Fills slot_to_api (n_tiles*TM entries, -1 = padding slot) from the caller's hint or from an RCB of the
// current positions (SoA, 3 planes of nm).  Inside a tile masses are sorted by API index.
// With ghost masses (halo exchange) own masses and ghosts never share a tile, and tiles are ordered
//   [tiles whose springs reach a ghost] [interior tiles] [ghost-only tiles]
// so that the step kernel handles (and exports) the boundary first and skips the ghost-only tail.
static void assign_tiles(Engine *e, const std::vector<double> &pos_soa, std::vector<int32_t> &slot_to_api,
                         int64_t *n_boundary, int64_t *n_active) {
    int64_t nm = e->nm;
    auto ghost = [&](int32_t i) { return (e->h_flags[(size_t)i] & F_GHOST) != 0; };
    std::vector<int32_t> order((size_t)nm);
    std::iota(order.begin(), order.end(), 0);
    std::stable_partition(order.begin(), order.end(), [&](int32_t i) { return !ghost(i); });
    const int64_t n_own = e->any_ghost ? (int64_t)(std::find_if(order.begin(), order.end(), ghost) - order.begin()) : nm;
    slot_to_api.clear();
    auto close_tile = [&](int in_tile) { if (in_tile) slot_to_api.resize(slot_to_api.size() + (size_t)(TM - in_tile), -1); };
    for (int part = 0; part < 2; part++) {      // own masses, then ghosts
        const int64_t lo = part ? n_own : 0, hi = part ? nm : n_own;
        if (lo == hi) continue;
        if (!e->tile_hint.empty()) {
            std::stable_sort(order.begin() + lo, order.begin() + hi, [&](int32_t a, int32_t b) { return e->tile_hint[a] < e->tile_hint[b]; });
            int in_tile = 0;
            for (int64_t q = lo; q < hi; q++) {
                bool new_tile = q > lo && (in_tile == TM || e->tile_hint[order[q]] != e->tile_hint[order[q - 1]]);
                if (new_tile) { close_tile(in_tile); in_tile = 0; }
                slot_to_api.push_back(order[q]);
                in_tile++;
            }
            close_tile(in_tile);
        } else {
            int64_t tiles = (hi - lo + TM - 1) / TM;
            rcb(order, pos_soa.data(), nm, lo, hi, tiles);
            for (int64_t t = 0; t < tiles; t++)
                std::sort(order.begin() + lo + t * TM, order.begin() + std::min<int64_t>(hi, lo + (t + 1) * TM));
            slot_to_api.insert(slot_to_api.end(), order.begin() + lo, order.begin() + hi);
            close_tile((int)((hi - lo) % TM));
        }
    }
    const int64_t n_tiles = (int64_t)slot_to_api.size() / TM;
    *n_active = n_tiles; *n_boundary = 0;
    if (!e->any_ghost) return;
    // classify: 0 = reaches a ghost, 1 = interior, 2 = ghost-only; stable reorder of whole tiles
    std::vector<uint8_t> cls((size_t)n_tiles, 1);
    for (int64_t T = 0; T < n_tiles; T++) {
        bool any_own = false, reaches = false;
        for (int t = 0; t < TM; t++) {
            int32_t i = slot_to_api[(size_t)(T * TM + t)];
            if (i < 0) continue;
            if (!ghost(i)) any_own = true;
            for (int32_t q = e->h_row_ptr[(size_t)i]; q < e->h_row_ptr[(size_t)i + 1] && !reaches; q++) {
                int32_t a = e->h_adj[(size_t)q], sp = a >> 1;
                int32_t o = (a & 1) ? e->h_left[(size_t)sp] : e->h_right[(size_t)sp];
                reaches = ghost(o);
            }
        }
        cls[(size_t)T] = !any_own ? 2 : (reaches ? 0 : 1);
    }
    std::vector<int64_t> tiles((size_t)n_tiles);
    std::iota(tiles.begin(), tiles.end(), 0);
    std::stable_sort(tiles.begin(), tiles.end(), [&](int64_t a, int64_t b) { return cls[(size_t)a] < cls[(size_t)b]; });
    std::vector<int32_t> re(slot_to_api.size());
    for (int64_t T = 0; T < n_tiles; T++)
        std::copy(slot_to_api.begin() + tiles[(size_t)T] * TM, slot_to_api.begin() + (tiles[(size_t)T] + 1) * TM, re.begin() + T * TM);
    slot_to_api.swap(re);
    *n_boundary = std::count(cls.begin(), cls.end(), (uint8_t)0);
    *n_active = n_tiles - std::count(cls.begin(), cls.end(), (uint8_t)2);
}

// ---------------------------------------------------------------- shared-memory plan -----------------------
struct SmemPlan {            // byte offsets inside the dynamic shared memory of one CTA
    int32_t cs;              // stride of one position component: TM own masses + halo, even
    int32_t vs;              // stride of one velocity component (cs when halo velocities are staged, else TM)
    int32_t hring_stride;    // bytes of one halo-index ring entry
    int32_t off_ctrl, off_sintab, off_hring, off_part, part_bytes, off_stage, stage_bytes;
    int32_t st_pos, st_vel;  // offsets inside a stage (the blob sits at 0)
    int32_t total;
    int32_t nst;             // stages of the tile ring (2..MAX_STAGES)
};
constexpr int MAX_STAGES = 4;
// mbarriers at offset 0: full[MAX_STAGES], empty[MAX_STAGES], hfull[HALO_RING]; tile control words at 128
// Position component stride is a compile-time constant of the kernel (CS_SMALL or CS_BIG doubles) so that the
// y and z reads are immediate offsets from the x address.  The last N_DUMMY local indices are the dummy masses.
static int choose_cs(int max_halo, bool strict) { return (!strict && TM + max_halo + N_DUMMY <= CS_SMALL) ? CS_SMALL : CS_BIG; }
static SmemPlan plan_smem(int cs, int max_halo, int max_blob, bool stage_vel, int tpm, int nst) {
    SmemPlan p;
    p.nst = nst;
    int hp = (max_halo + 3) & ~3;
    p.cs = cs;
    p.vs = stage_vel ? p.cs : TM;
    p.hring_stride = (int32_t)up16((size_t)hp * 4 + 16);
    p.off_ctrl = 128;
    p.off_sintab = 128 + MAX_STAGES * 32;
    p.off_hring = p.off_sintab + MAX_OMEGA * 8;
    p.off_part = p.off_hring + HALO_RING * p.hring_stride;
    p.part_bytes = (tpm - 1) * 3 * TM * 8;
    p.off_stage = p.off_part + nst * p.part_bytes;
    p.st_pos = (int32_t)up16((size_t)max_blob);
    p.st_vel = p.st_pos + 3 * p.cs * 8;
    p.stage_bytes = p.st_vel + 3 * p.vs * 8;
    p.total = p.off_stage + nst * p.stage_bytes;
    return p;
}

// ---------------------------------------------------------------- structure build (host) -----------------
template <class T> static int upload_vec(Engine *e, T **dst, const std::vector<T> &src) {
    TB_TRY(dev_alloc(e, dst, src.size()));
    if (!src.empty()) TB_CUDA(e, cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return TB_OK;
}
template <class T> static void release(Engine *e, T *&p, size_t count) { dev_free(e, (void *)p, count * sizeof(T)); p = nullptr; }

static void free_structure(Engine *e, TiledState *S, bool keep_state = false) {
    TableCaps &C = S->cap;
    if (!keep_state) { for (int p = 0; p < 2; p++) release(e, S->st[p], (size_t)C.st_tiles * STATE_BLOCK); C.st_tiles = 0; }
    release(e, S->class_rep, (size_t)C.cls); S->n_class_rep = 0; release(e, S->rec_map, (size_t)C.rec_map);
    S->n_rec_map = 0; S->dict = false; S->dict_split = false;
    release(e, S->krec_spring, (size_t)C.krec); S->n_krec = 0;
    if (S->act_tile_off) { cudaFree(S->act_tile_off); S->act_tile_off = nullptr; }
    if (S->act_rec) { cudaFree(S->act_rec); S->act_rec = nullptr; }
    S->n_act_rec = 0;
    if (S->krec_off) { cudaFree(S->krec_off); S->krec_off = nullptr; }
    release(e, S->blob, C.blob);
    release(e, S->slot_to_api, (size_t)C.tiles * TM); release(e, S->halo, (size_t)C.halo);
    release(e, S->api_to_slot, S->n_api); S->n_api = 0;
    release(e, S->rec_spring, (size_t)C.rec); release(e, S->ell, (size_t)C.ell);
    if (S->meta) { cudaFree(S->meta); S->meta = nullptr; }
    S->n_tiles = S->n_slots = S->n_rec = S->n_ell = S->n_halo = 0;
    S->n_halo_up = S->n_ell_up = S->n_rec_map_up = 0;
    S->blob_bytes = 0;
    const int64_t st_tiles = C.st_tiles;
    C = TableCaps(); C.st_tiles = st_tiles;
}

// n_rec records (k | rest | damping | omega arrays of n_rec slots each, rounded up to even), then the byte map of the
// dictionary modes; n_k >= 0 (split dictionary): the k array has n_k slots of its own (one per record slot)
static size_t blob_size(int n_rec, bool damping, bool active, int n_map = 0, int n_k = -1) {
    size_t np = (size_t)((n_rec + 1) & ~1), nk = n_k >= 0 ? (size_t)((n_k + 1) & ~1) : np;
    return (size_t)BLOB_REC + nk * 8 + np * 8 + (damping ? np * 8 : 0) + (active ? np * 8 : 0) + up16((size_t)n_map);
}

// Threads per mass.  Strict arithmetic needs the oracle's summation order (1).  Otherwise: big problems keep
// every SM busy with one thread per mass (fewest instructions and hand-offs; measured best on C2), small ones
// (fewer tiles than resident CTAs) are latency bound and split each mass's springs over 4 threads.
static bool resident_env();
static int choose_tpm(const Engine *e, int64_t n_tiles) {
    if ((e->flags & TB_MATH_STRICT) || e->any_ghost) return 1;
    // a system of <= 8 tiles runs on the resident cluster kernel: one thread per mass (which also allows dictionary records)
    if (n_tiles <= 8 && !(e->flags & TB_NO_RESIDENT) && resident_env() && !getenv("TB_TPM")) return 1;
    int tpm = n_tiles >= 2 * (int64_t)e->sm_count ? 1 : 4;
    if (const char *s = getenv("TB_TPM")) { int v = atoi(s); if (v == 1 || v == 2 || v == 4) tpm = v; }
    return tpm;
}

// Parameter classes for the dictionary mode: springs whose (k, rest, damping, omega) are bit-identical.  A lattice
// built by createLattice has a few hundred classes for millions of springs (one k, rest lengths that differ only by
// the rounding of the coordinates), so a tile's ~2,000 records collapse to a few dozen: the blob shrinks from ~35 KB
// to ~4 KB (less HBM traffic, less shared memory per CTA => more resident CTAs).  class_of[s] = dense class id,
// rep_of[s] = the first spring of the class.  Returns false (mode off) when there are too many classes.
#ifndef TB_DICT_CTAS
#define TB_DICT_CTAS 3
#endif
constexpr int DICT_CTAS = TB_DICT_CTAS;   // resident CTAs per SM the dictionary-mode kernels are compiled for (register cap)
constexpr int DICT_MAX_TILE = 256;        // records a tile's byte map can address
constexpr int DICT_MAX_GLOBAL = 1 << 16;
static int param_classes(Engine *e, bool with_k, bool damping, bool active, HostTables &H, int64_t s0, bool *ok) {
    *ok = false;
    const size_t ns = (size_t)e->ns, n = ns - (size_t)s0;
    std::vector<double> hk(n), hr(n), hd, ho;
    if (n) {
        TB_CUDA(e, cudaMemcpy(hk.data(), e->k + s0, n * 8, cudaMemcpyDeviceToHost));
        TB_CUDA(e, cudaMemcpy(hr.data(), e->rest + s0, n * 8, cudaMemcpyDeviceToHost));
        if (damping && e->damping) { hd.resize(n); TB_CUDA(e, cudaMemcpy(hd.data(), e->damping + s0, n * 8, cudaMemcpyDeviceToHost)); }
        if (active && e->omega) { ho.resize(n); TB_CUDA(e, cudaMemcpy(ho.data(), e->omega + s0, n * 8, cudaMemcpyDeviceToHost)); }
    }
    const bool typed = active && e->h_type.size() == ns;    // the actuation type is part of a record (sine-table mode scales per record)
    if (s0 == 0) { H.class_seen.clear(); H.class_first.clear(); }
    if (H.class_of.capacity() < ns) { H.class_of.reserve(ns + ns / 8 + 4096); H.rep_of.reserve(ns + ns / 8 + 4096); }   // springs added later: no copy
    H.class_of.resize(ns); H.rep_of.resize(ns);
    ClassKey last = {0, 0, 0, 0, 0}; int32_t last_class = -1;
    auto bits = [](double x) { uint64_t u; memcpy(&u, &x, 8); return u; };
    for (size_t sp = (size_t)s0; sp < ns; sp++) {
        const size_t q = sp - (size_t)s0;
        // an ACTUATED_* spring's rest length changes every step (the step kernel advances it in its record): a class of its own
        uint64_t own = (e->h_type.size() == ns && e->h_type[sp] >= TB_ACTUATED_EXPAND) ? ((uint64_t)(sp + 1) << 8) : 0;
        if (!e->h_removed.empty() && e->h_removed[sp]) own |= 1ull << 48;     // tombstones (k, damping zeroed) never represent live springs
        ClassKey key = {with_k ? bits(hk[q]) : 0, bits(hr[q]), hd.empty() ? 0 : bits(hd[q]), ho.empty() ? 0 : bits(ho[q]),
                        own | (typed ? (uint64_t)(e->h_type[sp] == 2 ? 1 : e->h_type[sp] == 3 ? 2 : 0) : 0)};
        int32_t c;
        if (last_class >= 0 && key == last) c = last_class;
        else {
            auto it = H.class_seen.find(key);
            if (it == H.class_seen.end()) {
                if ((int)H.class_first.size() >= DICT_MAX_GLOBAL) return TB_OK;
                c = (int32_t)H.class_first.size(); H.class_first.push_back((int32_t)sp); H.class_seen.emplace(key, c);
            } else c = it->second;
            last = key; last_class = c;
        }
        H.class_of[sp] = c; H.rep_of[sp] = H.class_first[(size_t)c];
    }
    *ok = true;
    return TB_OK;
}

// Distinct actuation frequencies of the ACTIVE_* springs (bit patterns).  At most MAX_OMEGA of them: sin(omega t) is
// evaluated once per CTA and step and the records carry a table index; more: per-spring sin() in the kernel (false).
static bool omega_table(const Engine *e, std::vector<double> &tab) {
    tab.clear();
    if (e->h_omega.size() != (size_t)e->ns || e->h_type.size() != (size_t)e->ns) return false;
    std::map<uint64_t, int> seen;
    for (int64_t sp = 0; sp < e->ns; sp++) {
        if (e->h_type[(size_t)sp] != 2 && e->h_type[(size_t)sp] != 3) continue;
        uint64_t bits; memcpy(&bits, &e->h_omega[(size_t)sp], 8);
        if (seen.emplace(bits, (int)tab.size()).second) { tab.push_back(e->h_omega[(size_t)sp]); if (tab.size() > (size_t)MAX_OMEGA) return false; }
    }
    return !tab.empty();
}

// ---------------------------------------------------------------- structure build (host), phase 1: one tile ------------
// What one tile contributes to the tables of the tiled layout.  Pure function of the topology, the mass -> slot assignment and
// the parameter classes: tiles are built independently (all host cores at start(); a handful of tiles after tb_add_springs).
struct TileOut {
    int deg = 0;                          // ELL rows
    int n_slots = 1;                      // record slots an incidence word can name (slot 0 = the k = 0 record of the padding words)
    int n_final = 1;                      // records stored in the blob: n_slots (one per spring) or the tile's distinct classes
    std::vector<uint32_t> block;          // incidence words [rows rounded up to ELL_CHUNK][TM], provisional dummy indices
    std::vector<int32_t> th;              // halo list: state index of every foreign mass, -1 = hole
    std::vector<int32_t> recs;            // spring behind every final record (-1 for record 0)
    std::vector<int32_t> kslots;          // split dictionary: spring behind every record SLOT
    std::vector<uint8_t> map;             // dictionary modes: slot -> record
    std::vector<std::pair<int32_t, int>> act;   // (ACTUATED_* spring, its final record)
    const char *why_not = nullptr;        // the tile does not fit the tiled format
    bool dict_overflow = false;           // more distinct records than a tile's byte map addresses: try the next record mode
    uint64_t hash = 0;                    // content hash of `block` once its dummy indices are final (seal_tile)
};
struct BuildCfg {
    bool dict = false, split = false, align_rows = true;
    bool drop_removed = false;            // leave the incidence words of removed springs out (tile patches, tb_compact)
    const std::vector<int32_t> *class_of = nullptr, *rep_of = nullptr;
};

// int32 key (>= 0) -> int32 value, open addressing; clear() undoes only what was inserted (the tables are per-thread and reused)
struct SmallMap {
    std::vector<int32_t> key, val;
    std::vector<uint32_t> used;
    uint32_t mask;
    explicit SmallMap(int bits) : key((size_t)1 << bits, -1), val((size_t)1 << bits, 0), mask((1u << bits) - 1) {}
    void clear() { for (uint32_t q : used) key[q] = -1; used.clear(); }
    int32_t *find(int32_t k) {
        for (uint32_t q = ((uint32_t)k * 2654435761u >> 7) & mask;; q = (q + 1) & mask) {
            if (key[q] == k) return &val[q];
            if (key[q] < 0) return nullptr;
        }
    }
    int32_t *insert(int32_t k, int32_t v) {            // k must not be present; the tables are sized for twice a tile's worst case
        uint32_t q = ((uint32_t)k * 2654435761u >> 7) & mask;
        while (key[q] >= 0) q = (q + 1) & mask;
        key[q] = k; val[q] = v; used.push_back(q);
        return &val[q];
    }
};
struct Half { int32_t s, o, os; uint8_t side; };      // spring, the other endpoint, that mass's slot, which end this mass is
struct TileScratch {
    std::vector<Half> rows = std::vector<Half>((size_t)TM * MAX_DEG);
    int cnt[TM];
    SmallMap halo_local{13}, rec_local{13}, cls_local{10};   // <= TM * MAX_DEG = 4096 foreign masses / springs, <= 256 classes
    std::vector<int32_t> slot_spring;
};

static void build_tile(const Engine *e, const TiledState *S, int64_t T, const BuildCfg &cfg, TileOut &o) {
    static thread_local std::unique_ptr<TileScratch> scratch;
    if (!scratch) scratch.reset(new TileScratch());        // lives as long as its thread (the builder's threads are short-lived)
    TileScratch &W = *scratch;
    Half *const rows = W.rows.data();
    int *const cnt = W.cnt;
    auto row = [&](int t) { return rows + (size_t)t * MAX_DEG; };
    Half aligned[MAX_DEG];
    const std::vector<int32_t> &rp = e->h_row_ptr, &adj = e->h_adj;
    const bool any_removed = cfg.drop_removed && !e->h_removed.empty();
    int deg = 0;
    bool overflow = false;
    for (int t = 0; t < TM; t++) {
        cnt[t] = 0;
        int32_t i = S->h_slot_to_api[(size_t)(T * TM + t)];
        if (i < 0 || !(e->h_flags[(size_t)i] & F_VALID) || (e->h_flags[(size_t)i] & F_GHOST)) continue;   // ghosts are updated by their owner
        Half *r = row(t);
        auto take = [&](int32_t a) {
            const int32_t s = a >> 1;
            const uint8_t side = (uint8_t)(a & 1);
            const int32_t other = side ? e->h_left[(size_t)s] : e->h_right[(size_t)s];
            if (other < 0 || !(e->h_flags[(size_t)other] & F_VALID)) return;   // spring skipped, src/sim.cu:1163
            // a tombstone (tb_remove_springs) keeps its incidence word (it reads k = 0) unless this build compacts: dropping the
            // words of scattered removals makes every tile's block unique, and the content dedupe that keeps the blocks of a
            // lattice L2-resident is worth far more than 0.1 % of dead words (C5: 168 -> 151 G spring-steps/s when dropped)
            if (any_removed && e->h_removed[(size_t)s]) return;
            if (cnt[t] == MAX_DEG) { overflow = true; return; }
            r[cnt[t]++] = Half{s, other, S->h_api_to_slot[(size_t)other], side};
        };
        if ((int64_t)i < e->csr_nm) for (int32_t q = rp[(size_t)i]; q < rp[(size_t)i + 1]; q++) take(adj[(size_t)q]);
        if (!e->extra_adj.empty()) {        // springs added since the CSR was built: higher indices, so they come last
            auto it = e->extra_adj.find(i);
            if (it != e->extra_adj.end()) for (int32_t a : it->second) take(a);
        }
        deg = std::max(deg, cnt[t]);
    }
    o.deg = overflow ? MAX_DEG + 1 : deg;
    if (overflow) { o.why_not = "a mass has more incident springs than the ELL format holds"; return; }
    // Row alignment: a mass with fewer springs than the tile's fullest one (a lattice surface) gets HOLES so that its
    // springs sit in the rows of the same springs of a full mass -- "same" = same offset in API index, which on a
    // lattice is the direction.  Order per mass is unchanged (ascending spring index), only padding moves; every row
    // then holds one direction, its in-tile neighbours are t + const, and a half-warp's reads spread over all banks.
    // A mass whose offsets are no subsequence of the template's keeps its compact rows.
    if (deg > 1 && cfg.align_rows) {
        int tmpl = -1;
        for (int t = 0; t < TM && tmpl < 0; t++) if (cnt[t] == deg) tmpl = t;
        const int32_t i_tmpl = S->h_slot_to_api[(size_t)(T * TM + tmpl)];
        const Half *rt = row(tmpl);
        for (int t = 0; t < TM; t++) {
            const int d = cnt[t];
            if (d == 0 || d == deg) continue;
            const int32_t i = S->h_slot_to_api[(size_t)(T * TM + t)];
            Half *r = row(t);
            for (int q = 0; q < deg; q++) aligned[q] = Half{-1, 0, 0, 0};
            int p = 0;
            bool ok = true;
            for (int q = 0; q < d && ok; q++) {
                const int32_t key = r[q].o - i;
                while (p < deg && rt[p].o - i_tmpl != key) p++;
                if (p == deg) ok = false; else aligned[p++] = r[q];
            }
            if (ok) { for (int q = 0; q < deg; q++) r[q] = aligned[q]; cnt[t] = deg; }
        }
    }
    auto has = [&](int t, int j) { return j < cnt[t] && row(t)[j].s >= 0; };
    o.block.assign((size_t)ell_chunks(deg) * ELL_CHUNK * TM, 0u);   // zero rows pad the last chunk
    // Halo slots are chosen so that the 16 lanes of a half-warp reading row j hit 16 different shared-memory
    // bank pairs (an LDS.64 conflict costs a replay).  Own neighbours sit at local index t + c_j (c_j = the
    // row's dominant offset), so a foreign mass met by lane t in row j wants a slot = t + c_j (mod 16): on a
    // lattice brick every row that reaches the mass asks for the same residue (the brick's linear index
    // extended over its shell), which makes every row conflict-free.  Other graphs fall back to the lowest
    // free slot whose residue no lane of the half-warp uses.  Unused slots (holes) are not gathered.
    std::vector<int32_t> &th = o.th;
    th.clear();
    SmallMap &halo_local = W.halo_local;                  // foreign mass -> its halo slot in this tile
    halo_local.clear();
    int cjs[MAX_DEG];
    for (int j = 0; j < deg; j++) {
        int votes[16] = {0}, cj = -1, best = 0;
        for (int t = 0; t < TM; t++) {
            if (!has(t, j)) continue;
            const int32_t oslot = row(t)[j].os;
            if (oslot / TM == T) votes[((oslot % TM) - t) & 15]++;
        }
        for (int r = 0; r < 16; r++) if (votes[r] > best) { best = votes[r]; cj = r; }
        cjs[j] = cj;
        for (int hw = 0; hw < TM / 16; hw++) {
            uint32_t used = 0;
            for (int t = hw * 16; t < hw * 16 + 16; t++) {
                if (!has(t, j)) continue;
                const Half &h = row(t)[j];
                if (h.os / TM == T) used |= 1u << (h.os % 16);
                else if (const int32_t *v = halo_local.find(h.o)) used |= 1u << (*v % 16);
            }
            for (int t = hw * 16; t < hw * 16 + 16; t++) {
                if (!has(t, j)) continue;
                const Half &h = row(t)[j];
                const int32_t oslot = h.os;
                if (oslot / TM == T || halo_local.find(h.o)) continue;
                int pick = -1;
                if (cj >= 0) {
                    // preferred residue first, but never at the price of a halo list that no longer fits the
                    // small shared-memory variant: beyond that budget any free slot will do (a bank conflict
                    // costs a replay, an oversized list costs occupancy or the tiled format altogether)
                    const int want = (t + cj) & 15, budget = CS_SMALL - TM - N_DUMMY;
                    for (int x = want; pick < 0 && x < budget; x += 16) if (x >= (int)th.size() || th[(size_t)x] < 0) pick = x;
                    for (int x = 0; pick < 0; x++) if (x >= (int)th.size() || th[(size_t)x] < 0) pick = x;
                } else {
                    for (int x = 0; x < (int)th.size() + 16 && pick < 0; x++) {
                        bool free_slot = x >= (int)th.size() || th[(size_t)x] < 0;
                        if (free_slot && (!((used >> (x % 16)) & 1u) || used == 0xffffu)) pick = x;
                    }
                }
                if (pick >= (int)th.size()) th.resize((size_t)pick + 1, -1);
                th[(size_t)pick] = (int32_t)((int64_t)(oslot / TM) * STATE_BLOCK + oslot % TM);
                halo_local.insert(h.o, pick);
                used |= 1u << (pick % 16);
            }
        }
    }
    // padding words: the dummy mass in the bank pair the lane's own neighbour would occupy, record 0
    for (size_t w = 0; w < o.block.size(); w++) {
        const int j = (int)(w / TM), t = (int)(w % TM);
        o.block[w] = (PAD_NB + (uint32_t)((t + ((j < deg && cjs[j] >= 0) ? cjs[j] : 0)) & 15)) << ELL_NB_SHIFT;
    }
    std::vector<int32_t> &slot_spring = W.slot_spring;    // record slot -> spring; slot 0: k = 0, read by padding words
    slot_spring.assign(1, -1);
    SmallMap &rec_local = W.rec_local;                    // spring -> its record slot in this tile
    rec_local.clear();
    const bool typed = !e->h_type.empty();
    for (int j = 0; j < deg; j++) for (int t = 0; t < TM; t++) {   // j-major walk: record slots of one row are consecutive
        if (!has(t, j)) continue;
        const Half &h = row(t)[j];
        const int32_t oslot = h.os;
        const int nb = (oslot / TM == T) ? oslot % TM : TM + *halo_local.find(h.o);
        int32_t *it = rec_local.find(h.s);
        if (!it) { it = rec_local.insert(h.s, (int32_t)slot_spring.size()); slot_spring.push_back(h.s); }
        const int rec = *it;
        if (nb >= MAX_LOCAL - N_DUMMY || rec >= MAX_REC) { o.why_not = "a tile reaches too many foreign masses or springs"; return; }
        uint32_t ty = 0;
        if (typed) { uint8_t x = e->h_type[(size_t)h.s]; ty = (x == 2) ? 1u : (x == 3) ? 2u : 0u; }
        o.block[(size_t)j * TM + (size_t)t] =
            ((uint32_t)nb << ELL_NB_SHIFT) | ((uint32_t)rec << ELL_REC_SHIFT) | (h.side ? ELL_SIDE : 0u) | (ty << ELL_TYPE_SHIFT) | ELL_VALID;
    }
    const int n_rec = (int)slot_spring.size();
    o.n_slots = n_rec;
    const bool actuated = e->any_actuated && !e->h_type.empty();
    if (!cfg.dict) {
        o.n_final = n_rec;
        o.recs.swap(slot_spring);
        for (int r = 1; r < n_rec && actuated; r++)
            if (e->h_type[(size_t)o.recs[(size_t)r]] >= TB_ACTUATED_EXPAND) o.act.push_back({o.recs[(size_t)r], r});
        return;
    }
    // dictionary modes: record slots -> the tile's distinct records; slot 0 (padding) -> record 0 (k = 0)
    SmallMap &cls_local = W.cls_local;                    // parameter class -> its record in this tile
    cls_local.clear();
    o.recs.assign(1, -1);
    o.map.assign(1, 0);
    o.map.reserve((size_t)n_rec);
    for (int r = 1; r < n_rec; r++) {
        const int32_t sp = slot_spring[(size_t)r], c = (*cfg.class_of)[(size_t)sp];
        int32_t *it = cls_local.find(c);
        if (!it) {
            if ((int)o.recs.size() >= DICT_MAX_TILE) { o.dict_overflow = true; return; }
            it = cls_local.insert(c, (int32_t)o.recs.size());
            o.recs.push_back((*cfg.rep_of)[(size_t)sp]);
        }
        o.map.push_back((uint8_t)*it);
    }
    o.n_final = (int)o.recs.size();
    for (int r = 1; r < o.n_final && actuated; r++)
        if (e->h_type[(size_t)o.recs[(size_t)r]] >= TB_ACTUATED_EXPAND) o.act.push_back({o.recs[(size_t)r], r});
    if (cfg.split) o.kslots.swap(slot_spring);             // slot -> spring, for the per-slot stiffness
}

// ---------------------------------------------------------------- structure build (host), phase 2: merge + upload -----
// Host copies of the tables the tiled kernel reads (TiledState::H), kept so that single tiles can be rebuilt later: a rebuilt
// tile's lists are appended and its TileMeta redirected (the old entries stay behind until the next full build).

// Once the shared-memory stride cs is known (it depends on the longest halo list of all tiles): the provisional dummy index of
// the padding words becomes the stride's own dummies [cs - N_DUMMY, cs), and the block gets its content hash (FNV-1a over the
// words).  Per tile, so it runs on all cores like build_tile.
static void seal_tile(TileOut &o, int cs) {
    const uint32_t base = (uint32_t)(cs - N_DUMMY);
    uint64_t hsh = 1469598103934665603ull ^ (uint64_t)o.deg;
    for (uint32_t &w : o.block) {
        if (!(w & ELL_VALID)) { const uint32_t nb = (w >> ELL_NB_SHIFT) & 0x1ffu; w = (base + (nb >= PAD_NB ? nb - PAD_NB : 0u)) << ELL_NB_SHIFT; }
        hsh ^= w; hsh *= 1099511628211ull;
    }
    o.hash = hsh;
}

// appends tile T's contribution (sealed: seal_tile)
static const char *merge_tile(Engine *e, TiledState *S, int64_t T, TileOut &o, bool keep_exchange) {
    HostTables &H = S->H;
    TileMeta &M = S->h_meta[(size_t)T];
    const TileMeta old = M;
    const bool dict = S->dict, split = S->dict_split;
    M.deg = (uint32_t)o.deg;
    while (H.halo.size() % 4) H.halo.push_back(0);               // 16-byte aligned halo lists (bulk copy)
    M.halo_off = (uint32_t)H.halo.size();
    M.n_halo = (uint32_t)o.th.size();
    H.halo.insert(H.halo.end(), o.th.begin(), o.th.end());      // -1 = hole, skipped by the gather
    // content dedupe of the ELL block (hash from seal_tile, verified by comparison)
    const uint64_t hsh = o.hash;
    bool found = false;
    std::vector<uint32_t> &cands = H.ell_index[hsh];
    for (uint32_t off : cands)
        if (off + o.block.size() <= H.ell.size() && std::equal(o.block.begin(), o.block.end(), H.ell.begin() + off)) { M.ell_off = off; found = true; break; }
    if (!found) {
        if (H.ell.size() + o.block.size() >= ((size_t)1 << 31)) return "ELL table exceeds int32 offsets";
        M.ell_off = (uint32_t)H.ell.size();
        cands.push_back(M.ell_off);
        H.ell.insert(H.ell.end(), o.block.begin(), o.block.end());
        H.unique_blocks++;
    }
    M.n_rec = (uint32_t)o.n_final;
    M.rec_off = (uint32_t)H.rec_spring.size();
    H.rec_spring.insert(H.rec_spring.end(), o.recs.begin(), o.recs.end());
    M.map_off16 = 0;
    size_t bb = blob_size(o.n_final, S->has_damping, S->has_active);
    if (dict) {
        while (H.rec_map.size() % 16) H.rec_map.push_back(0);
        M.map_off16 = (uint32_t)(H.rec_map.size() / 16);
        H.rec_map.insert(H.rec_map.end(), o.map.begin(), o.map.end());
        if (split) {
            H.krec_off[2 * (size_t)T] = (uint32_t)H.kslot_spring.size();
            H.krec_off[2 * (size_t)T + 1] = (uint32_t)o.kslots.size();
            H.kslot_spring.insert(H.kslot_spring.end(), o.kslots.begin(), o.kslots.end());
        }
        bb = blob_size(o.n_final, S->has_damping, S->has_active, o.n_slots, split ? o.n_slots : -1);
    }
    M.exp_off = 0; M.exp_rows = 0;
    M.tflags = (T < S->n_boundary) ? (uint32_t)TF_BOUNDARY : 0u;
    if (keep_exchange) { M.exp_off = old.exp_off; M.exp_rows = old.exp_rows; M.tflags = old.tflags; }
    if (H.blob_total / 16 >= ((size_t)1 << 32)) return "blob exceeds 64 GB";
    M.blob_off16 = (uint32_t)(H.blob_total / 16); M.blob_bytes = (uint32_t)bb;
    H.blob_total += bb;
    {   // where the rest lengths of this tile's actuated springs sit: after the k array (blob_size / tiled_refresh_params)
        const size_t np = (size_t)((M.n_rec + 1) & ~1u), nk = (dict && split) ? (size_t)((o.n_slots + 1) & ~1) : np;
        std::vector<int4> &list = H.act[(size_t)T];
        list.clear();
        for (const std::pair<int32_t, int> &a : o.act) list.push_back(make_int4((int)(BLOB_REC + nk * 8 + (size_t)a.second * 8), a.first, 0, 0));
        M.tflags = (M.tflags & ~(uint32_t)TF_ACTUATED) | (o.act.empty() ? 0u : (uint32_t)TF_ACTUATED);
    }
    H.tile_slots[(size_t)T] = o.n_slots;
    (void)e;
    return nullptr;
}

// maxima, shared-memory plan, cluster grouping, then every table to the device.  The state buffers survive (and grow with
// the tile count); the blobs are re-allocated, so the caller refreshes parameters and flags afterwards.
static int finalize_and_upload(Engine *e, TiledState *S, int64_t old_tiles, bool incremental = false) {
    HostTables &H = S->H;
    const int64_t n_tiles = (int64_t)S->h_meta.size();
    int max_halo = 0, max_rec = 0, max_deg = 0, max_dict = 0;
    size_t max_blob = 0;
    for (int64_t T = 0; T < n_tiles; T++) {
        const TileMeta &M = S->h_meta[(size_t)T];
        max_halo = std::max(max_halo, (int)M.n_halo); max_deg = std::max(max_deg, (int)M.deg);
        max_rec = std::max(max_rec, H.tile_slots[(size_t)T]); max_blob = std::max(max_blob, (size_t)M.blob_bytes);
        if (S->dict) max_dict = std::max(max_dict, (int)M.n_rec);
    }
    // stages of the tile ring (TB_STAGES): the small dictionary blobs leave room for more than two
    int nst = 2;      // measured on B200 (C2, dictionary mode): 3 stages 183.7 vs 2 stages 187.9 G spring-steps/s
    if (const char *v = getenv("TB_STAGES")) nst = std::min(MAX_STAGES, std::max(2, atoi(v)));
    if (S->tpm > 1) nst = 2;    // named barriers of the partial-sum hand-off are numbered for two stages
    while (nst > 2 && (size_t)plan_smem(S->cs, max_halo, (int)max_blob, S->stage_vel, S->tpm, nst).total > SMEM_LIMIT / 3) nst--;
    S->nst = nst;
    SmemPlan plan = plan_smem(S->cs, max_halo, (int)max_blob, S->stage_vel, S->tpm, nst);
    if ((size_t)plan.total > SMEM_LIMIT) { S->why_not = "a tile's working set exceeds the shared-memory budget"; S->supported = false; return TB_OK; }
    S->smem_bytes = (size_t)plan.total;
    // smallest group size c <= 8 such that every tile's foreign masses live in its own aligned group of c tiles
    S->res_cluster = 0;
    for (int c = 1; c <= 8 && !S->res_cluster; c++) {
        if (n_tiles % c) continue;
        bool ok = true;
        for (int64_t T = 0; T < n_tiles && ok; T++) {
            const TileMeta &M = S->h_meta[(size_t)T];
            const int64_t base = T - T % c;
            for (uint32_t h = 0; h < M.n_halo && ok; h++) {
                const int32_t hi = H.halo[(size_t)M.halo_off + h];
                if (hi >= 0 && (hi / STATE_BLOCK < base || hi / STATE_BLOCK >= base + c)) ok = false;
            }
        }
        if (ok) S->res_cluster = c;
    }
    // the primary copy of an actuated spring (the one that keeps the canonical rest array current): its first record in tile order
    std::vector<int4> act_rec;
    std::vector<uint32_t> act_tile_off;
    if (e->any_actuated) {
        std::vector<uint8_t> seen(e->h_act.size(), 0);
        for (int64_t T = 0; T < n_tiles; T++) {
            act_tile_off.push_back((uint32_t)act_rec.size());
            for (int4 r : H.act[(size_t)T]) {
                const size_t j = (size_t)(std::lower_bound(e->h_act.begin(), e->h_act.end(), r.y) - e->h_act.begin());
                const bool known = j < e->h_act.size() && e->h_act[j] == r.y;
                r.z = (known && !seen[j]) ? 1 : 0;
                if (known) seen[j] = 1;
                act_rec.push_back(r);
            }
        }
        act_tile_off.push_back((uint32_t)act_rec.size());
    }
    const int64_t n_slots = (int64_t)S->h_slot_to_api.size();
    const size_t ell_tail = (size_t)(MAX_DEG + 2 * ELL_CHUNK) * TM;      // the kernel prefetches whole register sets past a block
    const uint32_t ell_pad = (uint32_t)(S->cs - N_DUMMY) << ELL_NB_SHIFT;
    TableCaps &C = S->cap;
    const bool dict_mode = S->dict, split_mode = S->dict_split;
    if (!incremental) {                              // the host tables were rebuilt from nothing: every device table goes
        free_structure(e, S, /*keep_state=*/true);   // (device tables only; the host tables above stay)
        S->dict = dict_mode; S->dict_split = split_mode;
    }
    if (S->ell && S->ell_pad != ell_pad) { release(e, S->ell, (size_t)C.ell); C.ell = 0; }
    auto room = [](size_t used, size_t least) { return (int64_t)(used + used / 8 + least); };
    bool moved = false;                              // a table was (re-)allocated
    // Each table on its own: the host copy only grew at its end since the last call (merge_tile appends) -> when the device
    // table has room, the new end is copied behind what is already there; otherwise the table moves to a larger allocation
    // (room of an eighth for the next edits) and is uploaded whole.  `tail` elements of `fill` must follow the used part.
    auto put = [&](auto *&dst, int64_t &cap, int64_t &up, const auto &src, size_t tail, auto fill, size_t least) -> int {
        typedef typename std::decay<decltype(src)>::type Vec;
        if (dst && (int64_t)(src.size() + tail) <= cap) {
            if ((int64_t)src.size() > up) TB_CUDA(e, cudaMemcpy(dst + up, src.data() + up, (src.size() - (size_t)up) * sizeof(src[0]), cudaMemcpyHostToDevice));
        } else {
            release(e, dst, (size_t)cap);
            cap = room(src.size() + tail, least);
            Vec padded(src);
            padded.resize((size_t)cap, (typename Vec::value_type)fill);
            TB_TRY(upload_vec(e, &dst, padded));
            moved = true;
        }
        up = (int64_t)src.size();
        return TB_OK;
    };
    S->max_halo = max_halo; S->max_rec = max_rec; S->max_deg = max_deg; S->max_dict = max_dict;
    S->max_blob = (int)max_blob; S->ell_unique_tiles = H.unique_blocks;
    // state buffers and the per-tile tables (tile table, slots, split-dictionary offsets) share one capacity in tiles
    if (n_tiles > C.st_tiles || !S->st[0]) {         // first build, or tiles appended for new masses beyond the room
        const int64_t want = room((size_t)n_tiles, 64);
        if (want * (int64_t)STATE_BLOCK >= ((int64_t)1 << 31)) { S->why_not = "state exceeds int32 halo indices"; S->supported = false; return TB_OK; }
        for (int p = 0; p < 2; p++) {
            double *bigger = nullptr;
            TB_TRY(dev_alloc(e, &bigger, (size_t)want * STATE_BLOCK));
            TB_CUDA(e, cudaMemset(bigger, 0, (size_t)want * STATE_BLOCK * 8));
            if (S->st[p] && old_tiles > 0) TB_CUDA(e, cudaMemcpy(bigger, S->st[p], (size_t)std::min(old_tiles, n_tiles) * STATE_BLOCK * 8, cudaMemcpyDeviceToDevice));
            release(e, S->st[p], (size_t)C.st_tiles * STATE_BLOCK);
            S->st[p] = bigger;
        }
        C.st_tiles = want;
        moved = true;
    }
    if (!S->meta || C.tiles < n_tiles) {
        if (S->meta) { cudaFree(S->meta); S->meta = nullptr; }
        if (S->krec_off) { cudaFree(S->krec_off); S->krec_off = nullptr; }
        release(e, S->slot_to_api, (size_t)C.tiles * TM);
        C.tiles = C.st_tiles; S->n_slots = 0;
        TB_CUDA(e, cudaMalloc((void **)&S->meta, (size_t)C.tiles * sizeof(TileMeta)));
        std::vector<int32_t> slots(S->h_slot_to_api);
        slots.resize((size_t)C.tiles * TM, -1);
        TB_TRY(upload_vec(e, &S->slot_to_api, slots));
        moved = true;
    } else if (n_slots > S->n_slots)
        TB_CUDA(e, cudaMemcpy(S->slot_to_api + S->n_slots, S->h_slot_to_api.data() + S->n_slots, (size_t)(n_slots - S->n_slots) * 4, cudaMemcpyHostToDevice));
    TB_CUDA(e, cudaMemcpy(S->meta, S->h_meta.data(), (size_t)n_tiles * sizeof(TileMeta), cudaMemcpyHostToDevice));
    int64_t up_rec = S->n_rec, up_krec = S->n_krec, up_cls = S->n_class_rep;
    TB_TRY(put(S->halo, C.halo, S->n_halo_up, H.halo, 8, 0, (size_t)1 << 16));
    TB_TRY(put(S->ell, C.ell, S->n_ell_up, H.ell, ell_tail, ell_pad, (size_t)(n_tiles / 8 + 64) * MAX_DEG * TM));   // (a lattice shares its blocks: room by tiles)
    TB_TRY(put(S->rec_spring, C.rec, up_rec, H.rec_spring, 0, -1, (size_t)1 << 18));
    S->ell_pad = ell_pad;
    S->n_halo = S->n_halo_up + 4; S->n_ell = S->n_ell_up + (int64_t)ell_tail; S->n_rec = up_rec;
    if (S->dict) {
        TB_TRY(put(S->rec_map, C.rec_map, S->n_rec_map_up, H.rec_map, 32, 0, (size_t)1 << 18));
        TB_TRY(put(S->class_rep, C.cls, up_cls, H.rep_of, 0, 0, (size_t)1 << 16));
        S->n_class_rep = up_cls;
        if (!S->dict_bad) TB_CUDA(e, cudaMalloc((void **)&S->dict_bad, sizeof(int)));
        if (S->dict_split) {
            TB_TRY(put(S->krec_spring, C.krec, up_krec, H.kslot_spring, 0, -1, (size_t)1 << 18));
            S->n_krec = up_krec;
            if (!S->krec_off) TB_CUDA(e, cudaMalloc((void **)&S->krec_off, 2 * (size_t)C.tiles * sizeof(uint32_t)));
            TB_CUDA(e, cudaMemcpy(S->krec_off, H.krec_off.data(), H.krec_off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        }
    }
    {   // actuated records: small, rebuilt every time
        if (S->act_tile_off) { cudaFree(S->act_tile_off); S->act_tile_off = nullptr; }
        if (S->act_rec) { cudaFree(S->act_rec); S->act_rec = nullptr; }
        S->n_act_rec = 0;
        if (!act_rec.empty()) {
            S->n_act_rec = (int64_t)act_rec.size();
            TB_CUDA(e, cudaMalloc((void **)&S->act_tile_off, act_tile_off.size() * sizeof(uint32_t)));
            TB_CUDA(e, cudaMemcpy(S->act_tile_off, act_tile_off.data(), act_tile_off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
            TB_CUDA(e, cudaMalloc((void **)&S->act_rec, act_rec.size() * sizeof(int4)));
            TB_CUDA(e, cudaMemcpy(S->act_rec, act_rec.data(), act_rec.size() * sizeof(int4), cudaMemcpyHostToDevice));
        }
    }
    // the blobs: rebuilt tiles got fresh places at the end (zeroed room); the caller refreshes masses, flags and parameters
    if (!S->blob || H.blob_total > C.blob) {
        release(e, S->blob, C.blob);
        C.blob = (size_t)room(H.blob_total, 4u << 20);
        TB_TRY(dev_alloc(e, &S->blob, C.blob));
        TB_CUDA(e, cudaMemset(S->blob, 0, C.blob));
        moved = true;
    }
    S->n_tiles = n_tiles; S->n_slots = n_slots; S->blob_bytes = H.blob_total;
    S->supported = true;
    if (moved) S->n_uploads++; else S->n_patches++;
    return TB_OK;
}

template <class F> static void parallel_tiles(int64_t n, F f) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = n < 64 ? 1u : std::min(32u, hw);
    if (nt == 1) { for (int64_t T = 0; T < n; T++) f(T); return; }
    std::atomic<int64_t> next{0};
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([&] { for (int64_t T = next.fetch_add(16); T < n; T = next.fetch_add(16)) for (int64_t q = T; q < std::min(n, T + 16); q++) f(q); });
    for (std::thread &t : th) t.join();
}

// Builds tiles, ELL rows, halo lists and record tables on the host and uploads them.  Returns TB_OK with
// S->supported == false (and a reason) when the graph does not fit the tiled format.
// keep_slots: the mass -> slot assignment, the state buffers and the halo-exchange fields of the tile table stay
// (a rebuild after a parameter upload that broke the dictionary classes; peers keep their pointers and offsets).
// mode: 1 = try the full dictionary, then (2) the split one -- classes over (rest, damping, omega), k per record slot --
// then (0) one record per spring.
static int build_structure(Engine *e, TiledState *S, bool keep_slots = false, int mode = 1) {
    const int64_t old_tiles = keep_slots ? S->n_tiles : 0;
    if (!keep_slots) free_structure(e, S, false);
    S->supported = false;
    TB_TRY(ensure_csr(e));                        // masses / springs added since the last build join the CSR
    const int64_t nm = e->nm;
    if (!keep_slots) {
        std::vector<double> pos;
        if (e->tile_hint.empty()) {     // current positions (SoA) drive the bisection
            pos.resize((size_t)nm * 3);
            TB_CUDA(e, cudaMemcpy(pos.data(), e->pos[e->cur], (size_t)nm * 24, cudaMemcpyDeviceToHost));
        }
        assign_tiles(e, pos, S->h_slot_to_api, &S->n_boundary, &S->n_active);
    }
    const int64_t n_slots = (int64_t)S->h_slot_to_api.size(), n_tiles = n_slots / TM;
    if (n_tiles * (int64_t)STATE_BLOCK >= ((int64_t)1 << 31)) { S->why_not = "state exceeds int32 halo indices"; return TB_OK; }
    S->h_api_to_slot.assign((size_t)nm, -1);
    for (int64_t s = 0; s < n_slots; s++) if (S->h_slot_to_api[(size_t)s] >= 0) S->h_api_to_slot[(size_t)S->h_slot_to_api[(size_t)s]] = (int32_t)s;

    S->has_damping = e->any_damping; S->has_active = e->any_active;
    S->stage_vel = S->has_damping || (e->flags & TB_MATH_STRICT);
    if (!keep_slots) S->tpm = choose_tpm(e, n_tiles);     // a records-only rebuild keeps the summation order

    // dictionary mode: fast arithmetic, one thread per mass (the big-problem configuration)
    const int dict_env = getenv("TB_DICT") ? atoi(getenv("TB_DICT")) : 1;   // TB_DICT=0: one record per spring, 2: split only
    if (mode == 1 && (dict_env == 2 || S->dict_failures >= 2)) mode = 2;
    if (mode == 2 && S->split_failures >= 2) mode = 0;
    bool dict = mode != 0 && dict_env != 0 && !(e->flags & TB_MATH_STRICT) && S->tpm == 1 && e->ns > 0;
    if (dict && S->has_active) {      // dictionary records scale their rest length per record from the sine table
        std::vector<double> tab;
        if (!omega_table(e, tab)) dict = false;
    }
    const bool split = dict && mode == 2;
    const int next_mode = mode == 1 ? 2 : 0;          // what to try when this mode does not fit
    HostTables &H = S->H;
    H = HostTables();
    if (dict) {
        TB_TRY(param_classes(e, !split, S->has_damping, S->has_active, H, 0, &dict));
        if (!dict && next_mode) return build_structure(e, S, keep_slots, next_mode);
    }
    BuildCfg cfg;
    cfg.dict = dict; cfg.split = split;
    cfg.align_rows = !(getenv("TB_ALIGN_ROWS") && atoi(getenv("TB_ALIGN_ROWS")) == 0);   // measurement aid
    cfg.class_of = &H.class_of; cfg.rep_of = &H.rep_of;
    cfg.drop_removed = e->compact_requested;

    std::vector<TileOut> outs((size_t)n_tiles);
    parallel_tiles(n_tiles, [&](int64_t T) { build_tile(e, S, T, cfg, outs[(size_t)T]); });
    int max_halo = 0;
    for (const TileOut &o : outs) {
        if (o.why_not) { S->why_not = o.why_not; return TB_OK; }
        if (o.dict_overflow) return build_structure(e, S, keep_slots, next_mode);   // a tile with too many classes
        max_halo = std::max(max_halo, (int)o.th.size());
    }
    if (TM + max_halo + N_DUMMY > CS_BIG) { S->why_not = "a tile reaches too many foreign masses"; return TB_OK; }
    S->cs = choose_cs(max_halo, (e->flags & TB_MATH_STRICT) != 0);
    parallel_tiles(n_tiles, [&](int64_t T) { seal_tile(outs[(size_t)T], S->cs); });
    S->dict = dict; S->dict_split = split; S->build_mode = mode;
    S->h_meta.resize((size_t)n_tiles);            // keep_slots: the old entries hand over their exchange fields
    if (!keep_slots) { std::fill(S->h_meta.begin(), S->h_meta.end(), TileMeta()); S->exports_set = false; }
    H.act.assign((size_t)n_tiles, {});
    H.tile_slots.assign((size_t)n_tiles, 1);
    H.krec_off.assign(split ? 2 * (size_t)n_tiles : 0, 0u);
    for (int64_t T = 0; T < n_tiles; T++) {
        const char *why = merge_tile(e, S, T, outs[(size_t)T], keep_slots);
        if (why) { S->why_not = why; return TB_OK; }
        TileOut().block.swap(outs[(size_t)T].block);      // release as we go
    }
    e->compact_requested = false;
    H.base_ell = H.ell.size(); H.base_halo = H.halo.size(); H.base_rec = H.rec_spring.size(); H.base_map = H.rec_map.size(); H.base_blob = H.blob_total;
    return finalize_and_upload(e, S, old_tiles);
}

// Masses / springs were appended (tb_add_masses / tb_add_springs): new masses get tiles of their own at the end, and only the
// tiles holding an endpoint of a new spring are rebuilt.  *done = false: the layout cannot absorb the edit (a tile outgrows the
// format or the record mode) -- the caller rebuilds everything.
static int patch_structure(Engine *e, TiledState *S, bool *done) {
    *done = false;
    if (!S->supported || S->n_tiles == 0 || e->any_ghost || S->halo_connected || S->n_active != S->n_tiles) return TB_OK;
    if ((int64_t)S->h_meta.size() != S->n_tiles || (int64_t)S->H.tile_slots.size() != S->n_tiles) return TB_OK;
    HostTables &H = S->H;
    const int64_t old_tiles = S->n_tiles;
    // The lists of rebuilt tiles are appended and the old ones left behind: once that garbage is about one system's worth,
    // build everything afresh (half a second per million masses) instead of letting host and device tables grow for ever.
    if (H.ell.size() > 2 * H.base_ell + (size_t)old_tiles * MAX_DEG * TM || H.blob_total > 2 * H.base_blob + ((size_t)64 << 20) ||
        H.halo.size() > 2 * H.base_halo + ((size_t)1 << 22) || H.rec_spring.size() > 2 * H.base_rec + ((size_t)1 << 22) ||
        H.rec_map.size() > 2 * H.base_map + ((size_t)1 << 24))
        return TB_OK;
    const auto w0 = std::chrono::steady_clock::now();
    // new masses: tiles of their own, in index order
    std::vector<int64_t> touched;
    S->h_api_to_slot.resize((size_t)e->nm, -1);
    for (size_t q = 0; q < e->patch_masses.size(); q++) {
        if (q % TM == 0) { touched.push_back((int64_t)S->h_slot_to_api.size() / TM); S->h_slot_to_api.resize(S->h_slot_to_api.size() + TM, -1); }
        const size_t slot = S->h_slot_to_api.size() - TM + q % TM;
        S->h_slot_to_api[slot] = e->patch_masses[q];
        S->h_api_to_slot[(size_t)e->patch_masses[q]] = (int32_t)slot;
    }
    const int64_t n_tiles = (int64_t)S->h_slot_to_api.size() / TM;
    if (n_tiles * (int64_t)STATE_BLOCK >= ((int64_t)1 << 31)) return TB_OK;
    for (int32_t sp : e->patch_springs)
        for (int32_t end : {e->h_left[(size_t)sp], e->h_right[(size_t)sp]})
            if (end >= 0 && S->h_api_to_slot[(size_t)end] >= 0) touched.push_back(S->h_api_to_slot[(size_t)end] / TM);
    std::sort(touched.begin(), touched.end());
    touched.erase(std::unique(touched.begin(), touched.end()), touched.end());
    if (S->dict && (int64_t)H.class_of.size() < e->ns) {
        bool ok = false;
        TB_TRY(param_classes(e, !S->dict_split, S->has_damping, S->has_active, H, (int64_t)H.class_of.size(), &ok));
        if (!ok) return TB_OK;
    }
    BuildCfg cfg;
    cfg.dict = S->dict; cfg.split = S->dict_split;
    cfg.align_rows = !(getenv("TB_ALIGN_ROWS") && atoi(getenv("TB_ALIGN_ROWS")) == 0);
    cfg.class_of = &H.class_of; cfg.rep_of = &H.rep_of;
    cfg.drop_removed = true;
    const auto wa = std::chrono::steady_clock::now();
    std::vector<TileOut> outs(touched.size());
    parallel_tiles((int64_t)touched.size(), [&](int64_t q) { build_tile(e, S, touched[(size_t)q], cfg, outs[(size_t)q]); seal_tile(outs[(size_t)q], S->cs); });
    const auto wb = std::chrono::steady_clock::now();
    for (size_t q = 0; q < touched.size(); q++)
        if (outs[q].why_not || outs[q].dict_overflow || TM + (int)outs[q].th.size() + N_DUMMY > S->cs) return TB_OK;
    S->h_meta.resize((size_t)n_tiles, TileMeta());
    H.act.resize((size_t)n_tiles);
    H.tile_slots.resize((size_t)n_tiles, 1);
    if (S->dict_split) H.krec_off.resize(2 * (size_t)n_tiles, 0u);
    S->n_active = n_tiles;
    for (size_t q = 0; q < touched.size(); q++) {
        const char *why = merge_tile(e, S, touched[q], outs[q], touched[q] < old_tiles);
        if (why) return TB_OK;
    }
    const auto w1 = std::chrono::steady_clock::now();
    TB_TRY(finalize_and_upload(e, S, old_tiles, /*incremental=*/true));
    if (getenv("TB_TIMING"))
        fprintf(stderr, "[tb timing]   %zu tiles: classes %.2f ms, built %.2f ms, merged %.2f ms, device tables %.2f ms\n", touched.size(),
                std::chrono::duration<double, std::milli>(wa - w0).count(), std::chrono::duration<double, std::milli>(wb - wa).count(),
                std::chrono::duration<double, std::milli>(w1 - wb).count(), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w1).count());
    *done = S->supported;
    return TB_OK;
}

// ---------------------------------------------------------------- refresh kernels -------------------------
// canonical (API order, 3 planes of stride nm) <-> tile-blocked state
// skip_ghosts: once the halo exchange is connected a ghost's slot belongs to its owner (the peer pushes into it)
__global__ void tiled_gather_state(double *st, const double *pos_c, const double *vel_c, const int32_t *slot_to_api,
                                   int64_t n_slots, int64_t nm, const uint8_t *flags_c, int skip_ghosts) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int32_t i = slot_to_api[s];
    if (skip_ghosts && i >= 0 && (flags_c[i] & F_GHOST)) return;
    double *b = st + (s / TM) * STATE_BLOCK + (s % TM);
    for (int c = 0; c < 3; c++) {
        b[c * TM] = i >= 0 ? pos_c[c * nm + i] : 0.0;
        b[(3 + c) * TM] = i >= 0 ? vel_c[c * nm + i] : 0.0;
    }
}
__global__ void tiled_scatter_state(double *pos_c, double *vel_c, const double *st, const int32_t *slot_to_api,
                                    int64_t n_slots, int64_t nm) {
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    int32_t i = slot_to_api[s];
    if (i < 0) return;
    const double *b = st + (s / TM) * STATE_BLOCK + (s % TM);
    for (int c = 0; c < 3; c++) { pos_c[c * nm + i] = b[c * TM]; vel_c[c * nm + i] = b[(3 + c) * TM]; }
}
// one CTA of TM threads per tile: header, flags and masses of the tile's blob
__global__ void tiled_refresh_mass(uint8_t *blob, const TileMeta *meta, const int32_t *slot_to_api, const double *m_c,
                                   const uint8_t *flags_c) {
    const TileMeta M = meta[blockIdx.x];
    uint8_t *b = blob + (size_t)M.blob_off16 * 16;
    int t = threadIdx.x;
    int32_t i = slot_to_api[(int64_t)blockIdx.x * TM + t];
    if (t == 0) { int32_t *h = (int32_t *)b; h[0] = (int32_t)M.deg; h[1] = (int32_t)M.n_halo; h[2] = (int32_t)M.n_rec; }   // h[3]: tiled_refresh_params
    b[BLOB_FLAGS + t] = i >= 0 ? flags_c[i] : (uint8_t)0;
    ((double *)(b + BLOB_M))[t] = i >= 0 ? m_c[i] : 1.0;
}
// dictionary mode: do all springs still carry the parameters of their class representative (bit for bit)?
__device__ __forceinline__ int actuation_code(int type) { return type == TB_ACTIVE_CONTRACT_THEN_EXPAND ? 1 : type == TB_ACTIVE_EXPAND_THEN_CONTRACT ? 2 : 0; }
__global__ void dict_check_kernel(int64_t ns, const int32_t *class_rep, const double *k, const double *rest, const double *damping,
                                  const double *omega, const int32_t *type, const uint8_t *removed, int rep_matters, int *bad) {
    const int64_t sp = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (sp >= ns) return;
    const int32_t r = class_rep[sp];
    if (r == sp) return;
    if (removed && removed[sp]) return;           // a tombstone is in no tile's rows any more (or reads k = 0 from its own slot)
    // its class lost its representative, whose k and damping are zero now: matters where records take those from it (full
    // dictionary, or damped springs); the rest length and omega of a tombstone are still those of its class
    if (removed && removed[r]) { if (rep_matters) atomicOr(bad, 2); return; }
    bool same = __double_as_longlong(rest[sp]) == __double_as_longlong(rest[r]);
    if (k) same = same && __double_as_longlong(k[sp]) == __double_as_longlong(k[r]);      // null: split dictionary, k is per slot
    if (damping) same = same && __double_as_longlong(damping[sp]) == __double_as_longlong(damping[r]);
    if (omega) same = same && __double_as_longlong(omega[sp]) == __double_as_longlong(omega[r]);
    if (type) same = same && actuation_code(type[sp]) == actuation_code(type[r]);
    if (!same) atomicOr(bad, 1);
}
// one CTA per tile: the spring records of the tile's blob
__global__ void tiled_refresh_params(uint8_t *blob, const TileMeta *meta, const int32_t *rec_spring, const double *k,
                                     const double *rest, const double *damping, const double *omega, int has_damping,
                                     int has_active, const double *omega_tab, int n_omega, const uint8_t *rec_map,
                                     const int32_t *krec_spring, const uint32_t *krec_off, const int32_t *type) {
    const TileMeta M = meta[blockIdx.x];
    uint8_t *b = blob + (size_t)M.blob_off16 * 16;
    const uint32_t np = (M.n_rec + 1) & ~1u;      // slots per parameter array: k | rest | damping | omega | byte map
    // split dictionary: the k array has one slot per record slot of the tile, the other arrays one per distinct record
    const uint32_t k0 = krec_off ? krec_off[2 * blockIdx.x] : 0u, n_k = krec_off ? krec_off[2 * blockIdx.x + 1] : 0u;
    const uint32_t nk = krec_off ? ((n_k + 1) & ~1u) : np;
    double *kk = (double *)(b + BLOB_REC), *rr = kk + nk;
    double *dmp = rr + np;
    double *omg = dmp + (has_damping ? np : 0);
    if (rec_map) {   // dictionary mode: the slot -> record byte map follows the records (16-byte units)
        uint4 *dst = (uint4 *)(omg + (has_active ? np : 0));
        const uint4 *src = (const uint4 *)rec_map + M.map_off16;
        const uint32_t n16 = (uint32_t)((b + M.blob_bytes - (uint8_t *)dst) / 16);
        for (uint32_t r = threadIdx.x; r < n16; r += blockDim.x) dst[r] = src[r];
    }
    // one stiffness for the whole tile?  (record 0 is the k = 0 record of the padding words and does not count)
    const int32_t s1 = M.n_rec > 1 ? rec_spring[M.rec_off + 1] : -1;
    const long long k1 = s1 >= 0 ? __double_as_longlong(k[s1]) : 0;
    int same = M.n_rec > 1 && s1 >= 0 && !krec_off;
    for (uint32_t r = threadIdx.x; r < n_k; r += blockDim.x) {
        int32_t s = krec_spring[k0 + r];
        kk[r] = s >= 0 ? k[s] : 0.0;
    }
    for (uint32_t r = threadIdx.x; r < M.n_rec; r += blockDim.x) {
        int32_t s = rec_spring[M.rec_off + r];
        if (!krec_off) kk[r] = s >= 0 ? k[s] : 0.0;
        rr[r] = s >= 0 ? rest[s] : 0.0;
        if (r > 0 && (s < 0 || __double_as_longlong(k[s]) != k1)) same = 0;
        if (has_damping) dmp[r] = (s >= 0 && damping) ? damping[s] : 0.0;
        if (has_active) {
            double w = (s >= 0 && omega) ? omega[s] : 0.0;
            if (n_omega > 0) {      // table mode: the slot holds the index of omega in the distinct-value table (bits 0..15)
                int found = 0;      // and the record's actuation code (bits 16..17): the step kernel scales rest per record
                for (int q = 0; q < n_omega; q++) if (__double_as_longlong(omega_tab[q]) == __double_as_longlong(w)) found = q;
                const int code = (s >= 0 && type) ? actuation_code(type[s]) : 0;
                w = __longlong_as_double((long long)(found | (code << 16)));
            }
            omg[r] = w;
        }
    }
    same = __syncthreads_and(same);
    // header word 3: bit 0 = one stiffness for the whole tile (the step kernel keeps it in a register), bits 8.. = slots
    // of the k array in split-dictionary mode
    if (threadIdx.x == 0) ((int32_t *)b)[3] = same | (int32_t)(n_k << 8);
}

// ---------------------------------------------------------------- PTX helpers ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// global -> shared bulk copy completing on an mbarrier (SASS UBLKCP); 16-byte aligned, size multiple of 16
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// same, with an L2 eviction policy: spring records are streamed exactly once per step, so they should not push the
// mass state (which neighbouring tiles re-read as their halo) out of the 126 MB L2
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// 1/sqrt(x): MUFU.RSQ64H seed (about 2^-22 relative) + one cubically convergent correction; no special-case
// branch (x == 0 gives NaN exactly like the reference's temp / temp.norm(), include/Titan/vec.h:133-151)
__device__ __forceinline__ double rsqrt_fast(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-(x * y), y, 1.0);
    const double p = fma(0.375, e, 0.5) * e;
    return fma(y, p, y);
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// this thread's earlier cp.async copies arrive on the mbarrier when they have landed (count pre-charged at init)
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
constexpr long long HALO_TIMEOUT_CLOCKS = 1ll << 32;   // ~2 s: a peer that never signals must not hang the GPU

// ---------------------------------------------------------------- the step kernel -------------------------
struct TiledArgs {
    int32_t n_tiles;
    int32_t write_acc;               // 1: store acc (canonical array); 2: Verlet -- acc is state, read and stored every step
    int32_t rk2_stage;               // the reference's RK2 build: 1 = half step (massForcesAndUpdate<true>), 2 = full step from the backup
    int prof_lite;                   // TB_PROFILE=2: only the barrier waits and the loop totals
    unsigned long long *prof;        // TB_PROFILE=1: per-phase clock64 sums (0 wait full, 1 springs, 2 partial-sum handoff, 3 mass update, 4 producer wait empty, 5 producer wait halo list, 6 producer issue, 7 ell-copy wait)
    int32_t l2_hint;                 // 1: stream the spring-record blobs through L2 with evict_first
    int32_t debug;                   // TB_DEBUG_SKIP bits (timing experiments only): 1 no halo gather, 2 no springs, 4 no mass update, 8 no peer stores, 16 no peer wait
    SmemPlan sp;
    const double *st_in;
    double *st_out;
    const uint8_t *blob;
    const int32_t *slot_to_api, *halo;
    const uint32_t *ell;
    const TileMeta *meta;
    double *acc;                     // canonical (API-ordered) acceleration, stride X.nm
    const double *t_in;
    double *t_out;
    const double *omega_tab;         // [n_omega] distinct actuation frequencies (n_omega == 0: per-spring sin)
    int32_t n_omega;
    // ---- ACTUATED_* springs (tiles flagged TF_ACTUATED only)
    const uint32_t *act_tile_off; const int4 *act_rec;
    double *rest_canon; const int32_t *type; const double *l_max, *l_min, *rate;
    // ---- halo exchange (n_peers == 0: single-engine simulation, everything below unused) ----
    int32_t n_peers;
    int32_t n_boundary_units;        // boundary tiles x consumer warps: arrivals that complete a step's exports
    int32_t n_boundary_tiles;        // tiles [0, n_boundary_tiles) reach ghosts or are exported (they come first)
    const long long *step_in;        // steps completed before this launch (the exchange sequence number)
    long long *step_out;
    unsigned long long *my_flags;    // [n_peers] written by the peers: "my exports of step n have landed"
    unsigned long long *peer_flag[TB_MAX_PEERS];
    double *peer_out[TB_MAX_PEERS];  // the peers' state buffer this step writes (same ping-pong parity everywhere)
    const int2 *exp_tab;
    unsigned *boundary_done;
    int *halo_err;
};

// producer -> consumers, one per stage: where the NEXT tile's ELL words are and which chunks thread q takes
// (first chunk / chunk count, one byte per q)
struct TileCtrl { uint32_t next_ell_off, next_cb, next_nch, has_next, exp_off, exp_rows, tflags, pad; };
__host__ __device__ inline void chunk_table(int deg, int tpm, uint32_t &cb_packed, uint32_t &nch_packed) {
    const int n = ell_chunks(deg), c0 = ell_chunks0(n, tpm), cq = ell_chunks_q(n, tpm);
    cb_packed = 0; nch_packed = (uint32_t)c0;
    if (tpm == 1) nch_packed |= (uint32_t)deg << 8;      // one thread per mass: it also learns the exact row count
    for (int q = 1; q < tpm; q++) {
        int cb = c0 + (q - 1) * cq, nch = n - cb < cq ? n - cb : cq;
        if (nch < 0) nch = 0;
        cb_packed |= (uint32_t)cb << (8 * q); nch_packed |= (uint32_t)nch << (8 * q);
    }
}

// sin(omega t) of an ACTIVE_* spring: from the CTA's table of distinct frequencies (the record then holds the table
// index), else evaluated per spring, out of line.  Same argument, same sin(): bit-identical either way.
template <class P> static __device__ __noinline__ double sin_omega_t(double omega, double t) { return sin(P::mul(omega, t)); }
// ty: 1 = contract-then-expand (1 - 0.2 sin), 2 = expand-then-contract (1 + 0.2 sin), src/sim.cu:1168-1173
template <class P> __device__ __forceinline__ double actuation_from_sin(int ty, double sn) {
    return ty == 1 ? P::sub(1.0, P::mul(0.2, sn)) : P::add(1.0, P::mul(0.2, sn));
}

// The new state of the masses of one tile that are ghosts on a peer, stored straight into the peers' state buffers.
__device__ __noinline__ void halo_export(const TiledArgs &A, uint32_t exp_rows, uint32_t first, V3 pos, V3 vel) {
#pragma unroll 1
    for (uint32_t r = 0; r < exp_rows; r++) {
        const int2 x = __ldg(A.exp_tab + first + r * TM);
        if (x.x >= 0) {
            double *d = A.peer_out[x.x] + x.y;
            d[0] = pos.x; d[TM] = pos.y; d[2 * TM] = pos.z;
            d[3 * TM] = vel.x; d[4 * TM] = vel.y; d[5 * TM] = vel.z;
        }
    }
}

// ACTUATED_EXPAND / ACTUATED_CONTRACT (src/sim.cu:1174-1181): `if (_rest < _l_max) _rest += _rate * dt` /
// `if (_rest > _l_min) _rest -= _rate * dt`, applied to the spring's record before the force of the step is evaluated.
__device__ __forceinline__ double advance_rest(int ty, double x, double l_max, double l_min, double rate, double dt) {
    if (ty == TB_ACTUATED_EXPAND) { if (x < l_max) x = __dadd_rn(x, __dmul_rn(rate, dt)); }
    else if (ty == TB_ACTUATED_CONTRACT) { if (x > l_min) x = __dsub_rn(x, __dmul_rn(rate, dt)); }
    return x;
}
// The producer warp, before it stages a tile flagged TF_ACTUATED: advance the tile's actuated records in the blob (this
// CTA is the only reader and writer of this tile's blob), then order those stores before the bulk copy that follows.
template <class ARGS>
__device__ __noinline__ void advance_actuated_records(const ARGS &A, double dt, int64_t tile, uint8_t *blob_tile, int lane) {
    const uint32_t lo = __ldg(A.act_tile_off + tile), hi = __ldg(A.act_tile_off + tile + 1);
    for (uint32_t q = lo + (uint32_t)lane; q < hi; q += 32) {
        const int4 r = __ldg(A.act_rec + q);
        double *slot = reinterpret_cast<double *>(blob_tile + r.x);
        const double x = advance_rest(A.type[r.y], *slot, A.l_max[r.y], A.l_min[r.y], A.rate[r.y], dt);
        *slot = x;
        if (r.z) A.rest_canon[r.y] = x;
    }
    __syncwarp();
    asm volatile("fence.proxy.async;" ::: "memory");
}

// RK2 build (physics.cuh): stage 1 is a half step into the other state buffer; stage 2 steps from the BACKUP -- the state
// before stage 1, which still sits in the slot this thread is about to overwrite (ping-pong buffers; nobody else reads it).
// A mass with local constraints keeps its velocity backup in a side array: the constraints project the velocity in place before
// the backup is taken (MassExtras::rk2_bvel).
// The integrators other than the default Euler step, out of line AND by value: a reference parameter of a function that is
// not inlined makes the caller keep `force`, `pos` and `vel` in local memory for the whole tile loop (their addresses escape)
// -- measured: C2 196 -> 157 G for every integrator.  stage 0 = Verlet (acc = the previous acceleration), 1 / 2 = RK2.
struct TailOut { V3 pos, vel, acc; };
template <class P>
__device__ __noinline__ TailOut other_integrators(int stage, double dt, double mass, V3 force, const double *ob, V3 pos, V3 vel, V3 acc,
                                                  double *bvel, int64_t i_api, int64_t nm) {
    if (stage == 0) {
        verlet_tail<P>(dt, mass, force, pos, vel, acc);
    } else if (stage == 1) {
        if (bvel) { bvel[i_api] = vel.x; bvel[nm + i_api] = vel.y; bvel[2 * nm + i_api] = vel.z; }
        rk2_stage1_tail<P>(dt, mass, force, pos, vel, acc);
    } else {
        const V3 bv = bvel ? V3{bvel[i_api], bvel[nm + i_api], bvel[2 * nm + i_api]} : V3{ob[3 * TM], ob[4 * TM], ob[5 * TM]};
        rk2_stage2_tail<P>(dt, mass, force, V3{ob[0], ob[TM], ob[2 * TM]}, bv, pos, vel, acc);
    }
    return TailOut{pos, vel, acc};
}

// A consumer warp has finished `pending` boundary tiles: make its peer stores visible, count them, and -- as the
// last warp of the step -- release the step counter into every peer's flag word.  Out of line: runs once per warp
// and step, and the hot loop's instruction footprint matters.
__device__ __noinline__ void halo_report(const TiledArgs &A, unsigned pending) {
    const unsigned long long seq = (unsigned long long)*A.step_in;
    __threadfence_system();   // every lane: its own peer stores are performed system-wide ...
    __syncwarp();             // ... before lane 0 counts the warp's tiles as done
    if ((threadIdx.x & 31) == 0) {
        if (atomicAdd(A.boundary_done, pending) + pending == (unsigned)A.n_boundary_units) {
            *A.boundary_done = 0u;
            __threadfence_system();
#pragma unroll 1
            for (int p = 0; p < A.n_peers; p++) st_release_sys(A.peer_flag[p], seq + 1ull);
        }
    }
}

// Warp roles: warps 0 .. TM*TPM/32-1 are CONSUMERS (thread = (mass t, share q)); the last warp is the PRODUCER
// that keeps the two stages filled.  Synchronisation is mbarrier-only between the roles:
//   full[s]  : 1 expect_tx arrival (bulk copies: blob + own state) + 32 cp.async arrivals (halo gather)
//   empty[s] : one arrival per consumer warp when it has finished reading stage s
//   hfull[r] : bulk copy of a halo index list (producer-internal)
// plus one named barrier per 32-mass column and tile parity for handing the partial sums to thread q == 0.
// DICT: 0 = one record per spring, 1 = per-tile dictionary of (k, rest, damping, omega), 2 = dictionary of (rest, damping,
// omega) with the stiffness stored per record slot
template <class P, bool DAMPING, bool ACTIVE, int TPM, int CS, bool HALO, int DICT>
__global__ void __launch_bounds__(TM * TPM + 32, (TPM <= 2 ? (DICT ? DICT_CTAS : 2) : 1)) tiled_step_kernel(const __grid_constant__ TiledArgs A,
                                                                   const __grid_constant__ Globals G,
                                                                   const __grid_constant__ MassExtras X) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr bool STAGE_VEL = DAMPING || P::strict;
    constexpr int NT = TM * TPM;                       // consumer threads
    constexpr int NCW = NT / 32;                       // consumer warps
    const SmemPlan &sp = A.sp;
    const int tid = threadIdx.x;
    const uint32_t smem_base = smem_u32(smem);
    const int n_my = (A.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    auto tile_of = [&](int i) { return (int64_t)blockIdx.x + (int64_t)i * gridDim.x; };
    auto bar_full = [&](int s) { return smem_base + 8u * (uint32_t)s; };
    auto bar_empty = [&](int s) { return smem_base + 8u * (uint32_t)(MAX_STAGES + s); };
    auto bar_halo = [&](int r) { return smem_base + 8u * (uint32_t)(2 * MAX_STAGES + r); };
    static_assert(8 * (2 * MAX_STAGES + HALO_RING) <= 128 && sizeof(TileCtrl) == 32, "barriers, then the control words");
    const int NST = sp.nst;

    if (tid == 0) {
        for (int s = 0; s < NST; s++) { mbar_init(bar_full(s), 33); mbar_init(bar_empty(s), NCW); }
        for (int r = 0; r < HALO_RING; r++) mbar_init(bar_halo(r), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (tid < NST) {   // the dummy mass of every stage: far away, at rest
        double *pp = reinterpret_cast<double *>(smem + sp.off_stage + tid * sp.stage_bytes + sp.st_pos);
        for (int d = 1; d <= N_DUMMY; d++) {
            pp[CS - d] = 1e30; pp[2 * CS - d] = 1e30; pp[3 * CS - d] = 1e30;
            if (STAGE_VEL) { double *vp = pp + 3 * CS; vp[CS - d] = 0.0; vp[2 * CS - d] = 0.0; vp[3 * CS - d] = 0.0; }
        }
    }
    // Programmatic dependent launch: the next step's grid may become resident as this grid's CTAs retire.  Everything
    // up to griddepcontrol.wait touches only per-topology tables (meta, ELL words, halo lists), never the state or
    // the clocks the previous step is still writing.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // this thread's ELL words: chunks [cb, cb + nch) of the tile's block, rows of TM words
    const int t = tid % TM, q = tid / TM;
    constexpr int EPT = ell_words_per_thread(TPM);     // ELL words per thread
    auto load_ell = [&](uint32_t ell_off, int cb, int nch, uint32_t (&dst)[EPT]) {
        const uint32_t *src = A.ell + ell_off + (size_t)cb * ELL_CHUNK * TM + t;
#pragma unroll
        for (int g = 0; g < EPT / ELL_CHUNK; g++)
            if (g < nch) {
#pragma unroll
                for (int u = 0; u < ELL_CHUNK; u++) dst[g * ELL_CHUNK + u] = __ldg(src + (g * ELL_CHUNK + u) * TM);
            }
    };
    uint32_t ell_cur[EPT];
#pragma unroll
    for (int jj = 0; jj < EPT; jj++) ell_cur[jj] = 0u;
    int nch = 0, nch_nxt = 0, rows = 0, rows_nxt = 0;   // rows: the tile's ELL rows before padding (TPM == 1 only)
    if (tid < NT) {   // consumers: first tile's words are in flight while the previous step drains
        const TileMeta *m0 = A.meta + tile_of(0);
        uint32_t cbp, nchp;
        chunk_table((int)__ldg(&m0->deg), TPM, cbp, nchp);
        nch = (int)((nchp >> (8 * q)) & 0xff);
        rows = (int)((nchp >> 8) & 0xff);
        load_ell(__ldg(&m0->ell_off), (int)((cbp >> (8 * q)) & 0xff), nch, ell_cur);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");   // the previous step is complete and visible from here on
    if (ACTIVE && tid < A.n_omega)   // sin(omega t) once per CTA and step for every distinct omega (src/sim.cu:1168-1173)
        reinterpret_cast<double *>(smem + sp.off_sintab)[tid] = sin(P::mul(A.omega_tab[tid], *A.t_in));
    __syncthreads();

    if (tid >= NT) {
        // =========================================== PRODUCER warp ===========================================
        const int lane = tid - NT;
        if (blockIdx.x == 0 && lane == 0) *A.t_out = __dadd_rn(*A.t_in, A.rk2_stage ? __dmul_rn(0.5, G.dt) : G.dt);   // T += dt, src/sim.cu:1811 (RK2: T += 0.5 * dt per stage, :1789,1798)
        const int4 *mg = reinterpret_cast<const int4 *>(A.meta);
        auto load_meta = [&](int i, int4 &a, int4 &b, int4 &c) {
            if (i < n_my) { a = __ldg(mg + 3 * tile_of(i)); b = __ldg(mg + 3 * tile_of(i) + 1); c = __ldg(mg + 3 * tile_of(i) + 2); }
            else { a = make_int4(0, 0, 0, 0); b = a; c = a; }
        };
        // a = {blob_off16, blob_bytes, halo_off, n_halo}, b = {ell_off, deg, n_rec, rec_off}, c = {exp_off, exp_rows, tflags, -}
        auto issue_hidx = [&](int i, const int4 &a) {     // lane 0
            uint32_t bytes = ((uint32_t)a.w * 4u + 15u) & ~15u;
            if (bytes < 16u) bytes = 16u;
            const int r = i % HALO_RING;
            mbar_expect_tx(bar_halo(r), bytes);
            bulk_g2s(smem_base + (uint32_t)(sp.off_hring + r * sp.hring_stride), A.halo + (uint32_t)a.z, bytes, bar_halo(r));
        };
        int4 m0a, m0b, m0c, m1a, m1b, m1c, m2a, m2b, m2c, m3a, m3b, m3c;
        load_meta(0, m0a, m0b, m0c); load_meta(1, m1a, m1b, m1c); load_meta(2, m2a, m2b, m2c);
        if (blockIdx.x == 0 && lane == 0) *A.step_out = *A.step_in + 1;
        const unsigned long long seq = HALO ? (unsigned long long)*A.step_in : 0ull;
        bool peers_ready = !HALO || A.n_peers == 0;
        const uint64_t pol_stream = l2_policy_evict_first();
        if (lane == 0) { issue_hidx(0, m0a); if (n_my > 1) issue_hidx(1, m1a); }
        long long pp0 = 0, pp1 = 0, pp2 = 0;
        const long long pb = A.prof ? clock64() : 0;
        for (int k = 0, s = 0, ph = 1; k < n_my; k++) {   // s = k % NST, ph = parity of the stage's previous use
            __syncwarp();
            load_meta(k + 3, m3a, m3b, m3c);
            long long p0 = A.prof ? clock64() : 0;
            mbar_wait(bar_empty(s), (uint32_t)ph);    // consumers are done with tile k - NST
            long long p1 = A.prof ? clock64() : 0;
            const uint32_t stage = smem_base + (uint32_t)(sp.off_stage + s * sp.stage_bytes);
            if constexpr (HALO) if (!peers_ready && (m0c.z & TF_BOUNDARY)) {
                // this tile reads ghost slots: every peer must have landed its exports of the previous step
                // (acquire of the peer's step counter; the peers stored the state before releasing it)
                // warp-uniform spin (every lane polls one peer's word; the vote keeps the warp converged)
                const unsigned long long *word = A.my_flags + (lane < A.n_peers ? lane : 0);
                const long long t0 = clock64();
                while (!(A.debug & 16)) {
                    const bool ok = ld_acquire_sys(word) >= seq;
                    if (__all_sync(0xffffffffu, ok)) break;
                    if (__any_sync(0xffffffffu, clock64() - t0 > HALO_TIMEOUT_CLOCKS)) { if (lane == 0) atomicExch(A.halo_err, 1); break; }
                    __nanosleep(64);
                }
                __syncwarp();
                asm volatile("fence.proxy.async;" ::: "memory");   // the bulk copies below read what the peers wrote
                peers_ready = true;
            }
            if (m0c.z & TF_ACTUATED)
                advance_actuated_records(A, G.dt, tile_of(k), const_cast<uint8_t *>(A.blob) + (size_t)(uint32_t)m0a.x * 16, lane);
            if (lane == 0) {
                TileCtrl *c = reinterpret_cast<TileCtrl *>(smem + sp.off_ctrl) + s;
                if (HALO) { c->exp_off = (uint32_t)m0c.x; c->exp_rows = (uint32_t)m0c.y; c->tflags = (uint32_t)m0c.z; }
                uint32_t cbp, nchp;
                chunk_table(m1b.y, TPM, cbp, nchp);
                // has_next: bit 0 = another tile follows; bit 1 (halo exchange) = masses of THIS tile are ghosts on a peer
                c->next_ell_off = (uint32_t)m1b.x; c->next_cb = cbp; c->next_nch = nchp;
                c->has_next = ((k + 1 < n_my) ? 1u : 0u) | ((HALO && m0c.y) ? 2u : 0u);
                mbar_expect_tx(bar_full(s), (uint32_t)m0a.y + (uint32_t)(STATE_BLOCK * 8));
                if (A.l2_hint) bulk_g2s_hint(stage, A.blob + (size_t)(uint32_t)m0a.x * 16, (uint32_t)m0a.y, bar_full(s), pol_stream);
                else bulk_g2s(stage, A.blob + (size_t)(uint32_t)m0a.x * 16, (uint32_t)m0a.y, bar_full(s));
                const double *sb = A.st_in + tile_of(k) * STATE_BLOCK;
#pragma unroll
                for (int c3 = 0; c3 < 3; c3++) {
                    bulk_g2s(stage + (uint32_t)(sp.st_pos + c3 * CS * 8), sb + c3 * TM, TM * 8, bar_full(s));
                    bulk_g2s(stage + (uint32_t)(sp.st_vel + c3 * sp.vs * 8), sb + (3 + c3) * TM, TM * 8, bar_full(s));
                }
                if (k + 2 < n_my) issue_hidx(k + 2, m2a);
            }
            // halo gather of tile k: the index list was bulk-copied two tiles ago
            long long p2 = (A.prof && !A.prof_lite) ? clock64() : 0;
            mbar_wait(bar_halo(k % HALO_RING), (uint32_t)((k / HALO_RING) & 1));
            long long p3 = (A.prof && !A.prof_lite) ? clock64() : 0;
            const int n_halo = m0a.w;
            const int32_t *hl = reinterpret_cast<const int32_t *>(smem + sp.off_hring + (k % HALO_RING) * sp.hring_stride);
            for (int h = lane; h < ((A.debug & 1) ? 0 : n_halo); h += 32) {
                const int32_t hi = hl[h];
                if (hi < 0) continue;
                const double *src = A.st_in + hi;
                const uint32_t d = stage + (uint32_t)(sp.st_pos + (TM + h) * 8);
                cp_async8(d, src); cp_async8(d + (uint32_t)(CS * 8), src + TM); cp_async8(d + (uint32_t)(2 * CS * 8), src + 2 * TM);
                if (STAGE_VEL) {
                    const uint32_t dv = stage + (uint32_t)(sp.st_vel + (TM + h) * 8);
                    cp_async8(dv, src + 3 * TM); cp_async8(dv + (uint32_t)(sp.vs * 8), src + 4 * TM); cp_async8(dv + (uint32_t)(2 * sp.vs * 8), src + 5 * TM);
                }
            }
            cp_async_arrive_noinc(bar_full(s));
            if (A.prof && A.prof_lite) pp0 += p1 - p0;
            if (A.prof && !A.prof_lite) { long long p4 = clock64(); pp0 += p1 - p0; pp1 += p3 - p2; pp2 += (p2 - p1) + (p4 - p3); }
            m0a = m1a; m0b = m1b; m0c = m1c; m1a = m2a; m1b = m2b; m1c = m2c; m2a = m3a; m2b = m3b; m2c = m3c;
            if (++s == NST) { s = 0; ph ^= 1; }
        }
        if (A.prof && A.prof_lite) pp2 = clock64() - pb;   // lite: slot 6 = the whole producer loop
        if (A.prof && lane == 0) { atomicAdd(A.prof + 4, (unsigned long long)pp0); atomicAdd(A.prof + 5, (unsigned long long)pp1); atomicAdd(A.prof + 6, (unsigned long long)pp2); }
        return;
    }

    // ============================================= CONSUMER warps ============================================
    const double t_sim = *A.t_in;
    const double *sintab = reinterpret_cast<const double *>(smem + sp.off_sintab);
    // Boundary tiles come first in the tile order, so they are the first nb_my tiles of this CTA's sequence: each
    // warp reports them with ONE system-scope fence right after the last of them.
    const int nb_my = (HALO && A.n_peers > 0 && A.n_boundary_tiles > (int)blockIdx.x)
                          ? (A.n_boundary_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    long long pa0 = 0, pa1 = 0, pa2 = 0, pa3 = 0, pa4 = 0;
    const long long c_begin = A.prof ? clock64() : 0;
    for (int i = 0, s = 0, ph = 0; i < n_my; i++) {           // s = i % NST, ph = (i / NST) & 1
        long long c0 = A.prof ? clock64() : 0;
        mbar_wait(bar_full(s), (uint32_t)ph);     // blob, own state and halo of tile i have landed
        long long c1 = A.prof ? clock64() : 0;
        const unsigned char *stage = smem + sp.off_stage + s * sp.stage_bytes;
        const uint4 ctrl4 = *reinterpret_cast<const uint4 *>(smem + sp.off_ctrl + s * sizeof(TileCtrl));   // next_ell_off, next_cb, next_nch, has_next
        struct { uint32_t next_ell_off, next_cb, next_nch, has_next; } ctrl = {ctrl4.x, ctrl4.y, ctrl4.z, ctrl4.w};
        // The next tile's ELL words replace this tile's in the same registers, each right after its last use (the
        // loads are unconditional: the table is padded, words beyond a tile's rows are never evaluated).
        const uint32_t *nsrc = A.ell + t;
        if (ctrl.has_next & 1u) {
            nch_nxt = (int)((ctrl.next_nch >> (8 * q)) & 0xff);
            rows_nxt = (int)((ctrl.next_nch >> 8) & 0xff);
            nsrc += ctrl.next_ell_off + (size_t)((ctrl.next_cb >> (8 * q)) & 0xff) * ELL_CHUNK * TM;
        }
        const int n_rec = reinterpret_cast<const int32_t *>(stage)[2];
        const uint8_t fl = stage[BLOB_FLAGS + t];
        // record arrays (one 8-byte slot per record, np slots each): k | rest | damping | omega | byte map
        const int hdr3 = reinterpret_cast<const int32_t *>(stage)[3];
        const int np8 = ((n_rec + 1) & ~1) * 8;
        const int nk8 = DICT == 2 ? (((hdr3 >> 8) + 1) & ~1) * 8 : np8;
        const char *kb = reinterpret_cast<const char *>(stage + BLOB_REC);
        const char *rb = kb + nk8;
        const char *dmb = rb + np8;
        const char *omb = dmb + (DAMPING ? np8 : 0);
        // Tiles whose records all carry one stiffness (the header says so) keep it in a register: the shared-memory
        // pipe is the kernel's bottleneck and this takes one of the two parameter loads per spring off it.
        const bool k_uniform = DICT != 2 && (hdr3 & 1);
        const double k_tile = k_uniform ? *reinterpret_cast<const double *>(kb + 8) : 0.0;
        // dictionary mode: the ELL word's record slot goes through the tile's byte map to one of n_rec distinct records
        const uint8_t *mapb = reinterpret_cast<const uint8_t *>(omb + (ACTIVE ? np8 : 0));
        const char *pbx = reinterpret_cast<const char *>(stage + sp.st_pos), *pby = pbx + CS * 8, *pbz = pby + CS * 8;
        const char *vbx = reinterpret_cast<const char *>(stage + sp.st_vel), *vby = vbx + sp.vs * 8, *vbz = vby + sp.vs * 8;
        // Sine-table actuation: a record carries its actuation code, so rest * (1 -/+ 0.2 sin(omega t)) is evaluated once
        // per record and tile, in place in the staged copy (the same two operations the reference applies per spring,
        // src/sim.cu:1168-1173), and the spring loop below is the passive one.
        const bool act_table = ACTIVE && A.n_omega > 0;
        if (act_table) {
            double *rw = reinterpret_cast<double *>(const_cast<char *>(rb));
            for (int r = tid; r < n_rec; r += NT) {
                const int slot = *reinterpret_cast<const int *>(omb + 8 * r);
                const int ty = (slot >> 16) & 3;
                const double c = ty == 1 ? -0.2 : (ty == 2 ? 0.2 : 0.0);
                rw[r] = P::mul(rw[r], P::add(1.0, P::mul(c, sintab[slot & 0xffff])));
            }
            named_bar_sync(15, NT);
        }
        V3 pos = {reinterpret_cast<const double *>(pbx)[t], reinterpret_cast<const double *>(pby)[t], reinterpret_cast<const double *>(pbz)[t]};
        V3 vel = {0.0, 0.0, 0.0};
        if (STAGE_VEL || q == 0) vel = V3{reinterpret_cast<const double *>(vbx)[t], reinterpret_cast<const double *>(vby)[t], reinterpret_cast<const double *>(vbz)[t]};
        const bool live = (fl & F_VALID) && !(fl & (F_FIXED | F_GHOST));   // else padding / deleted / fixed / ghost: frozen (src/sim.cu:1292-1295)

        // ---- gather this thread's share of the incident springs, ascending row order ------------------------
        V3 force = {0.0, 0.0, 0.0};
        if (P::strict) {
            if (live) {
#pragma unroll
                for (int jj = 0; jj < EPT; jj++) {
                    const uint32_t en = (jj < nch * ELL_CHUNK) ? ell_cur[jj] : 0u;
                    if (en & ELL_VALID) {
                        const uint32_t nbo = en & 0xff8u, rco = (en >> 13) & 0x7ff8u;
                        const V3 po = {*(const double *)(pbx + nbo), *(const double *)(pby + nbo), *(const double *)(pbz + nbo)};
                        const V3 vo = {*(const double *)(vbx + nbo), *(const double *)(vby + nbo), *(const double *)(vbz + nbo)};
                        const double2 kr = make_double2(*(const double *)(kb + rco), *(const double *)(rb + rco));
                        SpringArgs spr;
                        spr.k = kr.x;
                        spr.rest_scaled = kr.y;
                        spr.damping = DAMPING ? *(const double *)(dmb + rco) : 0.0;
                        if (ACTIVE) {
                            const int ty = (en >> ELL_TYPE_SHIFT) & 3;
                            if (act_table) spr.rest_scaled = kr.y;       // scaled per record above
                            else if (ty) spr.rest_scaled = P::mul(kr.y, actuation_from_sin<P>(ty, sin_omega_t<P>(*(const double *)(omb + rco), t_sim)));
                        }
                        if (en & ELL_SIDE) acc3<P>(force, spring_force<P, DAMPING>(po, pos, vo, vel, spr));
                        else { V3 f = spring_force<P, DAMPING>(pos, po, vel, vo, spr); acc3<P>(force, V3{-f.x, -f.y, -f.z}); }
                    }
                }
            }
#pragma unroll
            for (int jj = 0; jj < EPT; jj++) ell_cur[jj] = __ldg(nsrc + jj * TM);
        } else {
            // ELL_CHUNK springs at a time without branches (invalid words evaluate a dummy spring whose
            // coefficient is then forced to zero), so their loads and fp64 chains interleave.
            // force on this mass = (p_me - p_other) * coef whichever endpoint it is (see physics.cuh)
            auto spring_t = [&](const uint32_t en, auto per_spring_sine) {
                constexpr bool SINE = decltype(per_spring_sine)::value;   // more distinct omegas than the table holds
                const uint32_t nbo = en & 0xff8u;
                uint32_t rco = (en >> 13) & 0x7ff8u;       // byte offset of the record's slot in each parameter array
                const uint32_t kco = rco;                  // DICT == 2: the stiffness stays with the slot
                if (DICT) rco = (uint32_t)mapb[(en >> 16) & 0xfffu] << 3;
                const char *pn = pbx + nbo;
                const double dx = pos.x - *(const double *)(pn);
                const double dy = pos.y - *(const double *)(pn + CS * 8);
                const double dz = pos.z - *(const double *)(pn + 2 * CS * 8);
                double2 kr;
                kr.y = *(const double *)(rb + rco);
                kr.x = k_uniform ? (rco ? k_tile : 0.0) : *(const double *)(kb + (DICT == 2 ? kco : rco));   // record 0 (padding words): k = 0
                double rest = kr.y;
                if (ACTIVE && SINE) {
                    const int ty = (en >> ELL_TYPE_SHIFT) & 3;
                    if (ty) rest = kr.y * actuation_from_sin<P>(ty, sin_omega_t<P>(*(const double *)(omb + rco), t_sim));
                }
                const double L2 = fma(dz, dz, fma(dy, dy, dx * dx));
                const double inv = rsqrt_fast(L2);
                // k (rest - L) / L = k (rest / L - 1); padding words read k = 0 and the far dummy mass
                double coef = kr.x * fma(rest, inv, -1.0);
                if (DAMPING) {   // dot(v_left - v_right, dir) * damping along dir; symmetric in the endpoints
                    const double dvx = *(const double *)(vbx + nbo) - vel.x, dvy = *(const double *)(vby + nbo) - vel.y,
                                 dvz = *(const double *)(vbz + nbo) - vel.z;
                    coef = fma(fma(dvz, dz, fma(dvy, dy, dvx * dx)) * (inv * inv), *(const double *)(dmb + rco), coef);
                }
                force.x = fma(dx, coef, force.x); force.y = fma(dy, coef, force.y); force.z = fma(dz, coef, force.z);
            };
            auto spring = [&](const uint32_t en) { spring_t(en, std::false_type{}); };       // passive, or scaled per record
            auto spring_sine = [&](const uint32_t en) { spring_t(en, std::true_type{}); };
            constexpr int FULL = 7;     // chunks of a tile whose masses have 25..28 incident springs (the lattice interior)
            if (TPM == 1 && !DAMPING && (!ACTIVE || act_table) && nch == FULL && !(A.debug & 2)) {
                // the common case as one straight line: no per-chunk branch, so the scheduler can start a chunk's
                // shared-memory loads while the previous chunk's fp64 chains are still running
                // A 3-D lattice mass has at most 26 springs: the block's last two rows are then padding, skipped.
                constexpr int LATTICE_DEG = 26;
#pragma unroll
                for (int jj = 0; jj < LATTICE_DEG; jj++) { spring(ell_cur[jj]); ell_cur[jj] = __ldg(nsrc + jj * TM); }
                if (rows > LATTICE_DEG) {
#pragma unroll
                    for (int jj = LATTICE_DEG; jj < FULL * ELL_CHUNK; jj++) spring(ell_cur[jj]);
                }
#pragma unroll
                for (int jj = LATTICE_DEG; jj < EPT; jj++) ell_cur[jj] = __ldg(nsrc + jj * TM);
            } else {
#pragma unroll
                for (int g = 0; g < EPT / ELL_CHUNK; g++) {
                    if (g < nch && !(A.debug & 2)) {
                        if (ACTIVE && !act_table) {
#pragma unroll
                            for (int u = 0; u < ELL_CHUNK; u++) spring_sine(ell_cur[g * ELL_CHUNK + u]);
                        } else {
#pragma unroll
                            for (int u = 0; u < ELL_CHUNK; u++) spring(ell_cur[g * ELL_CHUNK + u]);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < ELL_CHUNK; u++) ell_cur[g * ELL_CHUNK + u] = __ldg(nsrc + (g * ELL_CHUNK + u) * TM);
                }
            }
        }
        long long c2 = (A.prof && !A.prof_lite) ? clock64() : 0;
        if (TPM > 1) {   // hand the partial sums to thread q == 0 of the mass, which adds them in a fixed order
            double *part = reinterpret_cast<double *>(smem + sp.off_part + s * sp.part_bytes);
            const int bar_id = 1 + (t >> 5) + (TM / 32) * s;
            if (q > 0) {
                part[((q - 1) * 3 + 0) * TM + t] = force.x; part[((q - 1) * 3 + 1) * TM + t] = force.y; part[((q - 1) * 3 + 2) * TM + t] = force.z;
                named_bar_arrive(bar_id, 32 * TPM);
            } else {
                named_bar_sync(bar_id, 32 * TPM);
#pragma unroll
                for (int qq = 1; qq < TPM; qq++) {
                    force.x = P::add(force.x, part[((qq - 1) * 3 + 0) * TM + t]);
                    force.y = P::add(force.y, part[((qq - 1) * 3 + 1) * TM + t]);
                    force.z = P::add(force.z, part[((qq - 1) * 3 + 2) * TM + t]);
                }
            }
        }

        long long c3 = (A.prof && !A.prof_lite) ? clock64() : 0;
        // ---- mass update (one thread per mass) ---------------------------------------------------------------
        if (q == 0) {
            const int64_t tile = tile_of(i);
            double *ob = A.st_out + tile * STATE_BLOCK + t;
            if (live && !(A.debug & 4)) {
                const double mass = reinterpret_cast<const double *>(stage + BLOB_M)[t];
                const bool extras = X.ext || X.drag || (fl & F_LOCAL) || A.write_acc;
                const int64_t i_api = extras ? A.slot_to_api[tile * TM + t] : 0;
                V3 acc;
                force = mass_forces<P>(G, X, i_api, fl & F_LOCAL, mass, pos, vel, force);
                if (A.write_acc == 2 || A.rk2_stage) {   // the reference's VERLET / RK2 builds: rare, out of line
                    // Verlet: the previous acceleration is state, kept in the canonical array by this thread
                    V3 prev = {0.0, 0.0, 0.0};
                    if (A.write_acc == 2) prev = V3{A.acc[i_api], A.acc[X.nm + i_api], A.acc[2 * X.nm + i_api]};
                    const TailOut o = other_integrators<P>(A.rk2_stage, G.dt, mass, force, ob, pos, vel, prev,
                                                           (fl & F_LOCAL) ? X.rk2_bvel : nullptr, i_api, X.nm);
                    pos = o.pos; vel = o.vel; acc = o.acc;
                } else {
                    euler_tail<P>(G.dt, mass, force, pos, vel, acc);
                }
                if (A.write_acc) { A.acc[i_api] = acc.x; A.acc[X.nm + i_api] = acc.y; A.acc[2 * X.nm + i_api] = acc.z; }
            }
            // (a ghost's slot belongs to the peer that owns the mass -- but ghosts live in tiles of their own, which are
            // never processed: assign_tiles)
            ob[0] = pos.x; ob[TM] = pos.y; ob[2 * TM] = pos.z;
            ob[3 * TM] = vel.x; ob[4 * TM] = vel.y; ob[5 * TM] = vel.z;
            if (HALO && (ctrl.has_next & 2u)) {
                // halo exchange: the new state of a mass that is a ghost elsewhere goes straight into the peer's buffer;
                // the tile control word says whether this tile exports at all, its exchange fields are read only then
                const TileCtrl *cp = reinterpret_cast<const TileCtrl *>(smem + sp.off_ctrl) + s;
                const uint32_t exp_rows = (A.debug & 8) ? 0u : cp->exp_rows;
                if (exp_rows) halo_export(A, exp_rows, cp->exp_off + (uint32_t)t, pos, vel);   // boundary tiles only, out of line
            }
        }
        __syncwarp();
        long long c4 = (A.prof && !A.prof_lite) ? clock64() : 0;
        if (act_table) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the staged rest lengths were rewritten in place
        if ((tid & 31) == 0) mbar_arrive(bar_empty(s));     // this warp no longer reads stage s
        if (HALO && i + 1 == nb_my) halo_report(A, (unsigned)nb_my);   // this warp's exports of the step are all issued
        nch = nch_nxt; rows = rows_nxt;
        if (++s == NST) { s = 0; ph ^= 1; }
        if (A.prof && A.prof_lite) pa0 += c1 - c0;
        if (A.prof && !A.prof_lite) {
            unsigned sink = 0;
#pragma unroll
            for (int jj = 0; jj < EPT; jj++) sink ^= ell_cur[jj];
            if (sink == 0x12345u) c4++;                       // forces the pending ELL loads to land before c5
            long long c5 = clock64();
            pa0 += c1 - c0; pa1 += c2 - c1; pa2 += c3 - c2; pa3 += c4 - c3; pa4 += c5 - c4;
        }
    }
    if (A.prof && A.prof_lite) pa1 = clock64() - c_begin;   // lite: slot 1 = the whole tile loop
    if (A.prof && (tid & 31) == 0) {
        unsigned long long *pp = A.prof + 8 * (q > 0);
        atomicAdd(pp + 0, (unsigned long long)pa0); atomicAdd(pp + 1, (unsigned long long)pa1); atomicAdd(pp + 2, (unsigned long long)pa2);
        atomicAdd(pp + 3, (unsigned long long)pa3); atomicAdd(pp + 7, (unsigned long long)pa4);
    }
}

// ---------------------------------------------------------------- glue ------------------------------------
// -DTB_DEV_BUILD compiles only the variants a kernel experiment needs (fast arithmetic, one thread per mass, small stride,
// no damping, no halo exchange): the full set takes minutes to build.  Never used for a shipped library.
template <class P, bool D, bool AC, int TPM, int CS, bool HALO, int DICT> constexpr bool variant_built() {
#ifdef TB_DEV_BUILD
    return !P::strict && TPM == 1 && CS == CS_SMALL && !D && !HALO;
#else
    return true;
#endif
}
template <class P, bool D, bool AC, int TPM, int CS, bool HALO, int DICT>
static int launch_c(Engine *e, TiledState *S, const TiledArgs &A, const MassExtras &X) {
    if constexpr (!variant_built<P, D, AC, TPM, CS, HALO, DICT>()) return fail(e, TB_ERR_UNSUPPORTED, "this kernel variant is not part of a TB_DEV_BUILD library");
    else {
    // The dynamic shared-memory limit of a kernel is a per-device (per-context) attribute: remembered per template
    // instantiation AND device, under a lock (engines on several GPUs, or several facade GPU threads, in one process).
    static std::mutex mu;
    static size_t configured_smem[MAX_DEVICES] = {};
    static int blocks_per_sm_dev[MAX_DEVICES] = {};
    auto kern = tiled_step_kernel<P, D, AC, TPM, CS, HALO, DICT>;
    int blocks_per_sm = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        const int dv = e->device % MAX_DEVICES;
        if (configured_smem[dv] != S->smem_bytes || e->device >= MAX_DEVICES) {
            TB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S->smem_bytes));
            TB_CUDA(e, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm_dev[dv], kern, TM * TPM + 32, S->smem_bytes));
            if (blocks_per_sm_dev[dv] < 1) return fail(e, TB_ERR_CUDA, "tiled step kernel does not fit on an SM (%zu bytes of shared memory)", S->smem_bytes);
            configured_smem[dv] = S->smem_bytes;
        }
        blocks_per_sm = blocks_per_sm_dev[dv];
    }
    int64_t grid = std::min<int64_t>(S->n_active, (int64_t)e->sm_count * blocks_per_sm);
    if (const char *g = getenv("TB_GRID")) { int v = atoi(g); if (v > 0) grid = std::min<int64_t>(S->n_active, v); }
    S->grid = (int)grid;
    // programmatic dependent launch measured on B200 (C2): no gain (135.6 vs 137.0 G spring-steps/s) -> opt-in only
    static const bool use_pdl = getenv("TB_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(TM * TPM + 32); cfg.dynamicSmemBytes = S->smem_bytes; cfg.stream = e->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = use_pdl ? 1 : 0;
    TB_CUDA(e, cudaLaunchKernelEx(&cfg, kern, A, e->G, X));
    return TB_OK;
    }
}
template <class P, bool D, bool AC, int TPM>
static int launch_v(Engine *e, TiledState *S, const TiledArgs &A, const MassExtras &X) {
    static const bool force_halo = getenv("TB_FORCE_HALO") != nullptr;   // measurement aid: exchange code compiled in, inert
    constexpr bool CAN_DICT = !P::strict && TPM == 1;    // build_structure only chooses the dictionary mode then
    if (S->dict && !CAN_DICT) return fail(e, TB_ERR_INVALID, "internal: dictionary records with a kernel that cannot read them");
    constexpr int D1 = CAN_DICT ? 1 : 0, D2 = CAN_DICT ? 2 : 0;
    constexpr int CSS = P::strict ? CS_BIG : CS_SMALL, TPS = P::strict ? 1 : TPM;
    const bool big = P::strict || A.sp.cs == CS_BIG;
    if (A.n_peers > 0 || e->any_ghost || force_halo) {   // partitioned: one thread per mass (choose_tpm), exchange code compiled in
        if (TPM != 1) return fail(e, TB_ERR_INVALID, "internal: halo exchange needs one thread per mass");
        if (S->dict && S->dict_split) return big ? launch_c<P, D, AC, 1, CS_BIG, true, D2>(e, S, A, X) : launch_c<P, D, AC, 1, CSS, true, D2>(e, S, A, X);
        if (S->dict) return big ? launch_c<P, D, AC, 1, CS_BIG, true, D1>(e, S, A, X) : launch_c<P, D, AC, 1, CSS, true, D1>(e, S, A, X);
        return big ? launch_c<P, D, AC, 1, CS_BIG, true, 0>(e, S, A, X) : launch_c<P, D, AC, 1, CSS, true, 0>(e, S, A, X);
    }
    if (S->dict && S->dict_split) return big ? launch_c<P, D, AC, TPM, CS_BIG, false, D2>(e, S, A, X) : launch_c<P, D, AC, TPS, CSS, false, D2>(e, S, A, X);
    if (S->dict) return big ? launch_c<P, D, AC, TPM, CS_BIG, false, D1>(e, S, A, X) : launch_c<P, D, AC, TPS, CSS, false, D1>(e, S, A, X);
    return big ? launch_c<P, D, AC, TPM, CS_BIG, false, 0>(e, S, A, X) : launch_c<P, D, AC, TPS, CSS, false, 0>(e, S, A, X);
}
template <class P, int TPM> static int launch_f(Engine *e, TiledState *S, const TiledArgs &A, const MassExtras &X) {
    bool d = S->has_damping, a = S->has_active;
    if (d && a) return launch_v<P, true, true, TPM>(e, S, A, X);
    if (d) return launch_v<P, true, false, TPM>(e, S, A, X);
    if (a) return launch_v<P, false, true, TPM>(e, S, A, X);
    return launch_v<P, false, false, TPM>(e, S, A, X);
}

bool tiled_supported(const Engine *e) {
    if (e->nm == 0) return false;
    const TiledState *S = e->tiled;
    if (S && !(e->tiled_stale & TS_STRUCT) && !S->supported) return false;
    return true;
}

static void halo_release(Engine *e, TiledState *S);
static unsigned long long *g_prof_dev = nullptr;
static long long g_prof_launches = 0;
int tiled_prepare(Engine *e) {
    TimeScope ts("tiled_prepare");
    if (getenv("TB_PROFILE") && !g_prof_dev) { cudaMalloc((void **)&g_prof_dev, 16 * 8); cudaMemset(g_prof_dev, 0, 16 * 8); }
    if (!e->tiled) e->tiled = new TiledState();
    TiledState *S = e->tiled;
    if (e->any_actuated) TB_TRY(prepare_actuated(e));     // the tile builder records where their rest lengths live
    uint32_t stale = e->tiled_stale;
    if (S->n_tiles > 0 && (S->has_damping != e->any_damping || S->has_active != e->any_active)) stale |= TS_STRUCT;
    if (S->n_tiles == 0 && !(stale & TS_STRUCT) && S->supported) stale |= TS_STRUCT;
    if ((stale & TS_STRUCT) && S->halo_connected)
        return fail(e, TB_ERR_INVALID, "the device layout changed after tb_halo_connect: run tb_halo_prepare / connect / set_exports again");
    if ((stale & TS_PATCH) && !(stale & TS_STRUCT)) {
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
        if (e->canon_stale) TB_TRY(tiled_to_canonical(e));
        bool done = false;
        const auto w0 = std::chrono::steady_clock::now();
        TB_TRY(patch_structure(e, S, &done));
        if (getenv("TB_TIMING"))
            fprintf(stderr, "[tb timing] patch_structure (%zu masses, %zu springs) %s: %.2f ms\n", e->patch_masses.size(), e->patch_springs.size(),
                    done ? "done" : "needs a full build", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count());
        if (done) {
            e->graph_dirty = true;
            stale |= TS_PARAMS | TS_STATE | TS_FLAGS;       // the blobs were re-allocated; new tiles want their state
            if (S->api_to_slot) { release(e, S->api_to_slot, S->n_api); S->n_api = 0; }
        } else if (e->patch_needs_layout || !S->supported) stale |= TS_STRUCT;   // the layout cannot absorb the edit: build it afresh
        else e->removed_unpatched = true;                       // removals only: they stay in place with k = 0
    }
    e->patch_masses.clear(); e->patch_springs.clear(); e->patch_needs_layout = false;
    stale &= ~(uint32_t)TS_PATCH;
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
    if ((stale & TS_PARAMS) && !(stale & TS_STRUCT) && S->dict) {
        // the records are shared by parameter class: a parameter upload that splits a class needs new records
        TB_CUDA(e, cudaMemsetAsync(S->dict_bad, 0, sizeof(int), e->stream));
        dict_check_kernel<<<(unsigned)((e->ns + 255) / 256), 256, 0, e->stream>>>(e->ns, S->class_rep, S->dict_split ? nullptr : e->k, e->rest,
                                                                                  S->has_damping ? e->damping : nullptr,
                                                                                  S->has_active ? e->omega : nullptr,
                                                                                  S->has_active ? e->type : nullptr, e->removed_dev,
                                                                                  (!S->dict_split || S->has_damping) ? 1 : 0, S->dict_bad);
        int bad = 0;
        TB_CUDA(e, cudaMemcpyAsync(&bad, S->dict_bad, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
        // full dictionary: a tombstone still named by an incidence word shares its record (and its k) with its class
        bool forced = e->removed_unpatched && !S->dict_split;
        if (bad == 2) forced = true;          // only representatives were lost: the classes themselves still hold
        if (forced) bad |= 1;
        if (S->has_active) {      // the records of actuated springs index the sine table: it must still exist
            std::vector<double> tab;
            if (!omega_table(e, tab)) bad = 1;
        }
        if (bad) {
            if (!forced) { if (S->dict_split) S->split_failures++; else S->dict_failures++; }
            TB_TRY(build_structure(e, S, /*keep_slots=*/true));
            e->graph_dirty = true;
            if (!S->supported) return TB_OK;
            stale |= TS_PARAMS | TS_FLAGS;
        }
    }
    if (e->any_magnet && !S->api_to_slot) {           // the magnet pass looks masses up by API index
        TB_TRY(upload_vec(e, &S->api_to_slot, S->h_api_to_slot));
        S->n_api = S->h_api_to_slot.size();
    }
    unsigned gs = (unsigned)((S->n_slots + 255) / 256);
    if (stale & TS_STATE) {
        if (e->canon_stale && !(stale & TS_STRUCT)) TB_TRY(tiled_to_canonical(e));  // partial upload landed in canonical
        tiled_gather_state<<<gs, 256, 0, e->stream>>>(S->st[e->cur], e->pos[e->cur], e->vel[e->cur], S->slot_to_api, S->n_slots, e->nm, e->mflags,
                                                      S->halo_connected ? 1 : 0);
    }
    if (stale & (TS_FLAGS | TS_STATE))
        tiled_refresh_mass<<<(unsigned)S->n_tiles, TM, 0, e->stream>>>(S->blob, S->meta, S->slot_to_api, e->m, e->mflags);
    if ((stale & TS_PARAMS) && S->has_active) {
        // distinct omegas of the ACTIVE_* springs (bit patterns); more than MAX_OMEGA -> per-spring sin() in the kernel
        std::map<uint64_t, int> seen;
        std::vector<double> tab;
        bool table = e->h_omega.size() == (size_t)e->ns && e->h_type.size() == (size_t)e->ns;
        for (int64_t sp = 0; table && sp < e->ns; sp++) {
            if (e->h_type[(size_t)sp] != 2 && e->h_type[(size_t)sp] != 3) continue;
            uint64_t bits; memcpy(&bits, &e->h_omega[(size_t)sp], 8);
            if (seen.emplace(bits, (int)tab.size()).second) { tab.push_back(e->h_omega[(size_t)sp]); if (tab.size() > (size_t)MAX_OMEGA) table = false; }
        }
        if (!S->omega_tab) TB_CUDA(e, cudaMalloc((void **)&S->omega_tab, MAX_OMEGA * 8));
        S->n_omega = (table && !tab.empty()) ? (int)tab.size() : 0;
        if (S->n_omega) TB_CUDA(e, cudaMemcpyAsync(S->omega_tab, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice, e->stream));
        TB_CUDA(e, cudaStreamSynchronize(e->stream));   // `tab` is a local
        e->graph_dirty = true;
    }
    if (stale & TS_PARAMS)
        tiled_refresh_params<<<(unsigned)S->n_tiles, 128, 0, e->stream>>>(S->blob, S->meta, S->rec_spring, e->k, e->rest, e->damping,
                                                                          e->omega, S->has_damping ? 1 : 0, S->has_active ? 1 : 0, S->omega_tab, S->n_omega,
                                                                          S->dict ? S->rec_map : nullptr, S->dict_split ? S->krec_spring : nullptr,
                                                                          S->dict_split ? S->krec_off : nullptr, S->has_active ? e->type : nullptr);
    TB_CUDA(e, cudaGetLastError());
    e->tiled_stale = 0;
    e->removed_unpatched = false;
    return TB_OK;
}

int launch_tiled_step(Engine *e, bool use_ext, bool write_acc, int rk2_stage) {
    TiledState *S = e->tiled;
    if (!S || !S->supported) return fail(e, TB_ERR_INVALID, "tiled step requested without a tiled layout");
    TiledArgs A;
    A.n_tiles = (int32_t)S->n_active;
    A.write_acc = (e->flags & TB_INTEGRATOR_VERLET) ? 2 : (write_acc ? 1 : 0);
    A.rk2_stage = rk2_stage;
    A.prof = nullptr; A.prof_lite = 0;
    if (getenv("TB_PROFILE")) {
        A.prof = g_prof_dev; g_prof_launches++;
        A.prof_lite = atoi(getenv("TB_PROFILE")) == 2;
    }
    A.debug = 0;
    A.l2_hint = 1;
    if (const char *h = getenv("TB_L2_HINT")) A.l2_hint = atoi(h);
    if (const char *d = getenv("TB_DEBUG_SKIP")) A.debug = atoi(d);
    A.sp = plan_smem(S->cs, S->max_halo, S->max_blob, S->stage_vel, S->tpm, S->nst);
    A.st_in = S->st[e->cur]; A.st_out = S->st[e->cur ^ 1];
    A.blob = S->blob; A.slot_to_api = S->slot_to_api; A.halo = S->halo; A.ell = S->ell; A.meta = S->meta;
    A.acc = e->acc;
    A.t_in = e->t_dev + e->cur; A.t_out = e->t_dev + (e->cur ^ 1);
    A.omega_tab = S->omega_tab; A.n_omega = S->has_active ? S->n_omega : 0;
    A.act_tile_off = S->act_tile_off; A.act_rec = S->act_rec;
    A.rest_canon = e->rest; A.type = e->type; A.l_max = e->l_max; A.l_min = e->l_min; A.rate = e->rate;
    A.step_in = e->step_dev + e->cur; A.step_out = e->step_dev + (e->cur ^ 1);
    A.n_peers = 0; A.n_boundary_units = 0; A.n_boundary_tiles = 0; A.my_flags = nullptr; A.exp_tab = nullptr; A.boundary_done = nullptr; A.halo_err = nullptr;
    for (int p = 0; p < TB_MAX_PEERS; p++) { A.peer_flag[p] = nullptr; A.peer_out[p] = nullptr; }
    if (e->any_ghost || S->n_exp_tab) {
        if (!S->halo_connected) return fail(e, TB_ERR_INVALID, "ghost masses need tb_halo_prepare / tb_halo_connect / tb_halo_set_exports before tb_run_async");
        A.n_peers = S->n_peers;
        A.n_boundary_units = (int32_t)(S->n_boundary * (TM * S->tpm / 32));
        A.n_boundary_tiles = (int32_t)S->n_boundary;
        A.my_flags = S->flags_dev; A.exp_tab = S->exp_tab; A.boundary_done = S->done_dev; A.halo_err = S->err_dev;
        for (int p = 0; p < S->n_peers; p++) { A.peer_flag[p] = S->peer_flag[p]; A.peer_out[p] = S->peer_st[p][e->cur ^ 1]; }
    }
    MassExtras X;
    X.ext = e->any_magnet ? e->mag_force : ((use_ext && e->ext) ? e->ext : nullptr);   // magnet.cu folds the uploaded force in
    X.drag = e->drag;
    X.lc_ptr = e->any_local ? e->lc_ptr : nullptr;
    X.lc = e->lc;
    X.nm = e->nm;
    X.rk2_bvel = (rk2_stage && e->any_local) ? e->rk2_bvel : nullptr;
    int rc;
    if (e->flags & TB_MATH_STRICT) rc = launch_f<Strict, 1>(e, S, A, X);
    else if (S->tpm == 4) rc = launch_f<Fast, 4>(e, S, A, X);
    else if (S->tpm == 2) rc = launch_f<Fast, 2>(e, S, A, X);
    else rc = launch_f<Fast, 1>(e, S, A, X);
    if (rc != TB_OK) return rc;
    TB_CUDA(e, cudaGetLastError());
    return TB_OK;
}

int tiled_to_canonical(Engine *e) {
    TiledState *S = e->tiled;
    if (S && S->supported && S->n_slots > 0 && e->canon_stale) {
        tiled_scatter_state<<<(unsigned)((S->n_slots + 255) / 256), 256, 0, e->stream>>>(
            e->pos[e->cur], e->vel[e->cur], S->st[e->cur], S->slot_to_api, S->n_slots, e->nm);
        TB_CUDA(e, cudaGetLastError());
        TB_CUDA(e, cudaStreamSynchronize(e->stream));
    }
    e->canon_stale = false;
    return TB_OK;
}

void tiled_destroy(Engine *e) {
    if (g_prof_dev && e->tiled && e->tiled->n_tiles > 0) {   // TB_PROFILE=1: average clocks per tile and warp
        unsigned long long h[16];
        cudaMemcpy(h, g_prof_dev, sizeof h, cudaMemcpyDeviceToHost);
        const TiledState *S = e->tiled;
        double cw0 = (double)g_prof_launches * (double)S->n_tiles * (TM / 32), cwq = cw0 * (S->tpm - 1), pw = (double)g_prof_launches * (double)S->n_tiles;
        if (getenv("TB_PROFILE") && atoi(getenv("TB_PROFILE")) == 2)
            fprintf(stderr, "[tb profile lite] clocks/tile  consumer warps: wait_full %.0f of %.0f | producer: wait_empty %.0f of %.0f\n",
                    h[0] / cw0, h[1] / cw0, h[4] / pw, h[6] / pw);
        else
        fprintf(stderr, "[tb profile] clocks/tile  q0 warps: wait_full %.0f springs %.0f handoff %.0f mass_update %.0f ell_wait %.0f | q>0 warps: wait_full %.0f springs %.0f handoff %.0f tail %.0f ell_wait %.0f | producer: wait_empty %.0f wait_halo_list %.0f issue %.0f\n",
                h[0] / cw0, h[1] / cw0, h[2] / cw0, h[3] / cw0, h[7] / cw0, cwq > 0 ? h[8] / cwq : 0, cwq > 0 ? h[9] / cwq : 0, cwq > 0 ? h[10] / cwq : 0,
                cwq > 0 ? h[11] / cwq : 0, cwq > 0 ? h[15] / cwq : 0, h[4] / pw, h[5] / pw, h[6] / pw);
        cudaMemset(g_prof_dev, 0, sizeof h); g_prof_launches = 0;
    }
    if (!e->tiled) return;
    halo_release(e, e->tiled);
    if (e->tiled->omega_tab) cudaFree(e->tiled->omega_tab);
    if (e->tiled->flags_dev) { cudaFree(e->tiled->flags_dev); cudaFree(e->tiled->done_dev); cudaFree(e->tiled->err_dev); }
    free_structure(e, e->tiled);
    delete e->tiled;
    e->tiled = nullptr;
}

// ---------------------------------------------------------------- halo exchange: host side ------------------
struct HandleBlob {                 // the contents of a tb_halo_handle
    uint64_t magic;
    int64_t pid;
    int32_t device, cur;
    int64_t steps_done, state_doubles;
    cudaIpcMemHandle_t ipc[3];      // state[0], state[1], flags
    void *raw[3];                   // the same three, valid inside the owning process
};
static_assert(sizeof(HandleBlob) <= sizeof(tb_halo_handle), "tb_halo_handle too small");
constexpr uint64_t HANDLE_MAGIC = 0x7442323030484c4full;
constexpr size_t FLAGS_BYTES = 2u << 20;   // its own allocation granule, so that the IPC mapping starts at the flags

static void halo_release(Engine *e, TiledState *S) {
    (void)e;
    for (int p = 0; p < TB_MAX_PEERS; p++)
        for (int b = 0; b < 3; b++)
            if (S->ipc_opened[p][b]) { cudaIpcCloseMemHandle(S->ipc_opened[p][b]); S->ipc_opened[p][b] = nullptr; }
    if (S->exp_tab) { cudaFree(S->exp_tab); S->exp_tab = nullptr; }
    if (S->exp_flat) { cudaFree(S->exp_flat); S->exp_flat = nullptr; }
    S->n_exp_tab = S->n_exp_flat = 0;
    S->n_peers = 0;
    S->halo_connected = false;
}


// ---------------------------------------------------------------- resident small-system kernel -------------
// Systems of at most 8 tiles (<= 1,024 masses: the README example, a single soft robot) are bound by launch latency,
// not bandwidth: one step of the streaming kernel above costs ~6 us of which almost nothing is work.  This kernel
// runs ALL requested steps in ONE launch: one thread-block cluster, one CTA per tile, the state resident in shared
// memory for the whole run.  Per step every CTA copies the foreign masses its springs reach out of its neighbours'
// shared memory (distributed shared memory, ld.shared::cluster), evaluates its masses from local shared memory exactly
// like the streaming kernel (same ELL words, same records, same arithmetic, one thread per mass in ascending spring
// order -- also the oracle's order), writes the other state buffer and meets the cluster at ONE barrier.  The run loop
// of the reference (src/sim.cu:1800-1811: springs with t = T, masses, T += dt) becomes the loop of this kernel; HBM is
// touched at the start and at the end.
struct ResidentArgs {
    int32_t nsteps, cs, vs, stage_vel, dict_mode, verlet, write_acc, n_omega;
    int32_t off_halo, off_hptr, off_sintab, off_rs, off_pos, off_vel, total;     // shared-memory byte offsets (the blob sits at 0)
    const double *st_in; double *st_out;
    const uint8_t *blob; const int32_t *slot_to_api, *halo; const uint32_t *ell; const TileMeta *meta;
    double *acc; const double *t_in; double *t_out; const long long *step_in; long long *step_out; const double *omega_tab;
    // ACTUATED_* springs (tiles flagged TF_ACTUATED only), as in TiledArgs
    const uint32_t *act_tile_off; const int4 *act_rec;
    double *rest_canon; const int32_t *type; const double *l_max, *l_min, *rate;
};
constexpr int RES_ROWS = MAX_DEG;      // ELL words a thread keeps in registers
constexpr int RES_MAX_TILES = 8;       // portable thread-block cluster size

template <class P, bool DAMPING, bool ACTIVE>
__global__ void __launch_bounds__(TM, 1) resident_step_kernel(const __grid_constant__ ResidentArgs A, const __grid_constant__ Globals G,
                                                               const __grid_constant__ MassExtras X) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(128) unsigned char smem[];
    const int t = threadIdx.x, tile = blockIdx.x;
    const bool multi = cluster.num_blocks() > 1;      // several independent systems = several clusters in one launch
    const TileMeta M = A.meta[tile];
    const int cs = A.cs, vs = A.vs;
    // ---- one-time loads: blob, halo list, ELL words, own state ------------------------------------------------
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.blob + (size_t)M.blob_off16 * 16);
        uint4 *dst = reinterpret_cast<uint4 *>(smem);
        for (uint32_t q = t; q < M.blob_bytes / 16; q += TM) dst[q] = src[q];
        int32_t *hl = reinterpret_cast<int32_t *>(smem + A.off_halo);
        for (uint32_t h = t; h < M.n_halo; h += TM) hl[h] = A.halo[M.halo_off + h];
    }
    uint32_t ell[RES_ROWS];
    const int rows = (int)ell_chunks((int)M.deg) * ELL_CHUNK;
#pragma unroll
    for (int j = 0; j < RES_ROWS; j++)     // rows beyond the block: the padding word (dummy mass, record 0 with k = 0)
        ell[j] = (j < rows) ? __ldg(A.ell + M.ell_off + (size_t)j * TM + t) : ((uint32_t)(A.cs - 1) << ELL_NB_SHIFT);
    double *Pb = reinterpret_cast<double *>(smem + A.off_pos), *Vb = reinterpret_cast<double *>(smem + A.off_vel);
    auto Pc = [&](int buf, int c) { return Pb + (size_t)(buf * 3 + c) * cs; };
    auto Vc = [&](int buf, int c) { return Vb + (size_t)(buf * 3 + c) * vs; };
    {
        const double *sb = A.st_in + (size_t)tile * STATE_BLOCK + t;
        for (int c = 0; c < 3; c++) { Pc(0, c)[t] = sb[c * TM]; Vc(0, c)[t] = sb[(3 + c) * TM]; }
        if (t < N_DUMMY) for (int b = 0; b < 2; b++) for (int c = 0; c < 3; c++) {      // the dummy masses: far away, at rest
            Pc(b, c)[cs - 1 - t] = 1e30;
            if (A.stage_vel) Vc(b, c)[vs - 1 - t] = 0.0;
        }
    }
    __syncthreads();
    const int n_rec = reinterpret_cast<const int32_t *>(smem)[2], hdr3 = reinterpret_cast<const int32_t *>(smem)[3];
    const int np8 = ((n_rec + 1) & ~1) * 8;
    const int nk8 = A.dict_mode == 2 ? (((hdr3 >> 8) + 1) & ~1) * 8 : np8;
    const char *kb = reinterpret_cast<const char *>(smem + BLOB_REC), *rb = kb + nk8, *dmb = rb + np8;
    const char *omb = dmb + (DAMPING ? np8 : 0);
    const uint8_t *mapb = reinterpret_cast<const uint8_t *>(omb + (ACTIVE ? np8 : 0));
    const uint8_t fl = smem[BLOB_FLAGS + t];
    const double mass = reinterpret_cast<const double *>(smem + BLOB_M)[t];
    const bool live = (fl & F_VALID) && !(fl & (F_FIXED | F_GHOST));
    const int64_t i_api = A.slot_to_api[(size_t)tile * TM + t];
    const int32_t *hl = reinterpret_cast<const int32_t *>(smem + A.off_halo);
    double *sintab = reinterpret_cast<double *>(smem + A.off_sintab);
    // sine-table actuation: rest * (1 -/+ 0.2 sin(omega t)) once per record and step (the record carries its actuation
    // code), into a table of its own -- the blob stays as loaded
    double *rs = reinterpret_cast<double *>(smem + A.off_rs);
    const char *rsb = reinterpret_cast<const char *>(rs);
    const bool act_table = ACTIVE && A.n_omega > 0;
    // sin(omega t) is a long dependent chain: the LAST threads of the CTA (padding slots of a 5 x 5 x 5 brick) evaluate it
    // for the NEXT step at the top of a step, off the critical path of the barrier
    auto fill_sintab = [&](double t_for) {
        const int w = TM - 1 - t;
        if (w < A.n_omega) sintab[w] = sin(P::mul(A.omega_tab[w], t_for));
    };
    auto scale_records = [&]() {
        for (int r = t; r < n_rec; r += TM) {
            const int slot = *reinterpret_cast<const int *>(omb + 8 * r), ty = (slot >> 16) & 3;
            const double c = ty == 1 ? -0.2 : (ty == 2 ? 0.2 : 0.0);
            rs[r] = P::mul(*reinterpret_cast<const double *>(rb + 8 * r), P::add(1.0, P::mul(c, sintab[slot & 0xffff])));
        }
    };
    // ACTUATED_* records of this tile: advanced in the resident copy of the blob before every step's forces
    // (src/sim.cu:1174-1181), written back to HBM (blob and canonical rest array) at the end of the run
    const uint32_t act_lo = (M.tflags & TF_ACTUATED) ? A.act_tile_off[tile] : 0u, act_hi = (M.tflags & TF_ACTUATED) ? A.act_tile_off[tile + 1] : 0u;
    const bool has_act = act_hi > act_lo;             // uniform over the CTA
    auto advance_records = [&]() {
        for (uint32_t q = act_lo + (uint32_t)t; q < act_hi; q += TM) {
            const int4 r = A.act_rec[q];
            double *slot = reinterpret_cast<double *>(smem + r.x);
            *slot = advance_rest(A.type[r.y], *slot, A.l_max[r.y], A.l_min[r.y], A.rate[r.y], G.dt);
        }
    };
    V3 acc = {0.0, 0.0, 0.0};
    if (A.verlet && live) acc = V3{A.acc[i_api], A.acc[X.nm + i_api], A.acc[2 * X.nm + i_api]};
    double t_sim = *A.t_in;
    // Halo exchange by PUSH: the owner of a mass stores its new state straight into the halo slots of the CTAs whose
    // springs reach it (remote shared-memory stores do not stall the writer; a pull would pay a load round trip per
    // step).  Each CTA registers its halo slots with the owners once: exp_cnt[slot] destinations per own mass,
    // exp_dst[slot][k] = generic pointer to the destination's position buffer 0, x component.
    int *exp_cnt = reinterpret_cast<int *>(smem + A.off_hptr);
    double **exp_dst = reinterpret_cast<double **>(smem + A.off_hptr + TM * 4);
    const ptrdiff_t v_minus_p = (A.off_vel - A.off_pos) / 8;
    exp_cnt[t] = 0;
    cluster.sync();
    {
        double *my_p = cluster.map_shared_rank(Pb, cluster.block_rank());
        for (uint32_t h = t; h < M.n_halo; h += TM) {
            const int32_t hi = hl[h];
            if (hi < 0) continue;
            const int owner = hi / STATE_BLOCK - (tile - (int)cluster.block_rank()), slot = hi % STATE_BLOCK;
            const int k = atomicAdd(cluster.map_shared_rank(exp_cnt, owner) + slot, 1);
            cluster.map_shared_rank(exp_dst, owner)[slot * (RES_MAX_TILES - 1) + k] = my_p + TM + h;
        }
    }
    cluster.sync();
    const int n_exp = exp_cnt[t];
    auto push = [&](int buf, const V3 &p, const V3 &v) {
        for (int k = 0; k < n_exp; k++) {
            double *d = exp_dst[t * (RES_MAX_TILES - 1) + k];
            d[(size_t)(buf * 3 + 0) * cs] = p.x; d[(size_t)(buf * 3 + 1) * cs] = p.y; d[(size_t)(buf * 3 + 2) * cs] = p.z;
            if (A.stage_vel) {
                double *dv = d + v_minus_p;
                dv[(size_t)(buf * 3 + 0) * vs] = v.x; dv[(size_t)(buf * 3 + 1) * vs] = v.y; dv[(size_t)(buf * 3 + 2) * vs] = v.z;
            }
        }
    };
    push(0, V3{Pc(0, 0)[t], Pc(0, 1)[t], Pc(0, 2)[t]}, V3{Vc(0, 0)[t], Vc(0, 1)[t], Vc(0, 2)[t]});
    if (has_act && A.nsteps > 0) { advance_records(); __syncthreads(); }      // for the first step
    if (act_table) { fill_sintab(t_sim); __syncthreads(); scale_records(); }
    cluster.sync();                                   // every CTA's state buffer 0, halo included, is in place

    int cur = 0;
    for (int step = 0; step < A.nsteps; step++, cur ^= 1) {
        // the halo slots of buffer `cur` were filled by their owners before the barrier that ended the previous step
        if (step > 0) { if (multi) cluster.barrier_wait(); else __syncthreads(); }
        if (act_table) fill_sintab(__dadd_rn(t_sim, G.dt));      // for the next step; this step reads only `rs`
        const char *pbx = reinterpret_cast<const char *>(Pc(cur, 0)), *pby = reinterpret_cast<const char *>(Pc(cur, 1)),
                   *pbz = reinterpret_cast<const char *>(Pc(cur, 2));
        const char *vbx = reinterpret_cast<const char *>(Vc(cur, 0)), *vby = reinterpret_cast<const char *>(Vc(cur, 1)),
                   *vbz = reinterpret_cast<const char *>(Vc(cur, 2));
        V3 pos = {Pc(cur, 0)[t], Pc(cur, 1)[t], Pc(cur, 2)[t]}, vel = {Vc(cur, 0)[t], Vc(cur, 1)[t], Vc(cur, 2)[t]};
        if (live) {
            V3 force = {0.0, 0.0, 0.0};
            // one spring; padding words evaluate a dummy spring (far-away mass, k = 0) that contributes exactly zero, so
            // the fast form needs no branch and the springs of a chunk interleave
            auto spring_t = [&](const uint32_t en, auto per_spring_sine) {
                constexpr bool SINE = decltype(per_spring_sine)::value;   // more distinct omegas than the table holds: a call
                if (P::strict && !(en & ELL_VALID)) return;               // per spring, kept out of the common straight line
                const uint32_t nbo = en & 0xff8u;
                uint32_t rco = (en >> 13) & 0x7ff8u;
                const uint32_t kco = rco;
                if (A.dict_mode) rco = (uint32_t)mapb[(en >> 16) & 0xfffu] << 3;
                const double k = *(const double *)(kb + (A.dict_mode == 2 ? kco : rco));
                double rest = *(const double *)((act_table ? rsb : rb) + rco);    // table mode: scaled per record
                if (ACTIVE && SINE) {
                    const int ty = (en >> ELL_TYPE_SHIFT) & 3;
                    if (ty) rest = P::mul(rest, actuation_from_sin<P>(ty, sin_omega_t<P>(*(const double *)(omb + rco), t_sim)));
                }
                const V3 po = {*(const double *)(pbx + nbo), *(const double *)(pby + nbo), *(const double *)(pbz + nbo)};
                if (P::strict) {
                    V3 vo = {0.0, 0.0, 0.0};
                    if (A.stage_vel) vo = V3{*(const double *)(vbx + nbo), *(const double *)(vby + nbo), *(const double *)(vbz + nbo)};
                    SpringArgs spr;
                    spr.k = k; spr.rest_scaled = rest; spr.damping = DAMPING ? *(const double *)(dmb + rco) : 0.0;
                    if (en & ELL_SIDE) acc3<P>(force, spring_force<P, DAMPING>(po, pos, vo, vel, spr));
                    else { V3 f = spring_force<P, DAMPING>(pos, po, vel, vo, spr); acc3<P>(force, V3{-f.x, -f.y, -f.z}); }
                } else {   // the streaming kernel's fast form: force on this mass = (p_me - p_other) * coef
                    const double dx = pos.x - po.x, dy = pos.y - po.y, dz = pos.z - po.z;
                    const double L2 = fma(dz, dz, fma(dy, dy, dx * dx));
                    const double inv = rsqrt_fast(L2);
                    double coef = k * fma(rest, inv, -1.0);
                    if (DAMPING) {
                        const double dvx = *(const double *)(vbx + nbo) - vel.x, dvy = *(const double *)(vby + nbo) - vel.y,
                                     dvz = *(const double *)(vbz + nbo) - vel.z;
                        coef = fma(fma(dvz, dz, fma(dvy, dy, dvx * dx)) * (inv * inv), *(const double *)(dmb + rco), coef);
                    }
                    force.x = fma(dx, coef, force.x); force.y = fma(dy, coef, force.y); force.z = fma(dz, coef, force.z);
                }
            };
            // up to 28 rows (every 3-D lattice) as one straight line, so that independent springs overlap: with one warp
            // per scheduler there is nothing else to hide the fp64 / MUFU latencies behind
            constexpr int STRAIGHT = 28;
            if (ACTIVE && !act_table) {
#pragma unroll 1
                for (int g = 0; g < RES_ROWS / ELL_CHUNK; g++) {
                    if (g * ELL_CHUNK >= rows) break;
#pragma unroll
                    for (int u = 0; u < ELL_CHUNK; u++) {
                        uint32_t en = 0;
#pragma unroll
                        for (int j = 0; j < RES_ROWS; j++) if (j == g * ELL_CHUNK + u) en = ell[j];   // register array, dynamic g
                        spring_t(en, std::true_type{});
                    }
                }
            } else if (rows <= 8) {            // low-degree systems (the README cube: 7 springs per mass) skip the padding rows
#pragma unroll
                for (int j = 0; j < 8; j++) spring_t(ell[j], std::false_type{});
            } else if (rows <= 16) {
#pragma unroll
                for (int j = 0; j < 16; j++) spring_t(ell[j], std::false_type{});
            } else {
#pragma unroll
                for (int j = 0; j < STRAIGHT; j++) spring_t(ell[j], std::false_type{});
                if (rows > STRAIGHT) {
#pragma unroll
                    for (int j = STRAIGHT; j < RES_ROWS; j++) spring_t(ell[j], std::false_type{});
                }
            }
            force = mass_forces<P>(G, X, i_api, fl & F_LOCAL, mass, pos, vel, force);
            if (A.verlet) verlet_tail<P>(G.dt, mass, force, pos, vel, acc);
            else euler_tail<P>(G.dt, mass, force, pos, vel, acc);
        }
        const int nxt = cur ^ 1;
        Pc(nxt, 0)[t] = pos.x; Pc(nxt, 1)[t] = pos.y; Pc(nxt, 2)[t] = pos.z;
        Vc(nxt, 0)[t] = vel.x; Vc(nxt, 1)[t] = vel.y; Vc(nxt, 2)[t] = vel.z;
        push(nxt, pos, vel);
        t_sim = __dadd_rn(t_sim, G.dt);               // T += dt, src/sim.cu:1811
        if (has_act && step + 1 < A.nsteps) {         // the next step's rest lengths: after this step's readers, before the
            __syncthreads();                          // tables below and the barrier at the top of the next step
            advance_records();
        }
        if (act_table) {                              // the tables are read by this CTA's threads only: separate them from
            __syncthreads();                          // this step's readers, the barrier below publishes the new ones
            scale_records();
        }
        // all new states visible, nobody still reads the old ones: arrive here, wait at the top of the next step
        if (multi) cluster.barrier_arrive();
    }
    // ---- back to HBM ----------------------------------------------------------------------------------------
    if (multi) cluster.barrier_wait();   // also keeps every CTA's shared memory alive until all pushes have landed
    else __syncthreads();
    {
        double *ob = A.st_out + (size_t)tile * STATE_BLOCK + t;
        for (int c = 0; c < 3; c++) { ob[c * TM] = Pc(cur, c)[t]; ob[(3 + c) * TM] = Vc(cur, c)[t]; }
        if (live && (A.write_acc || A.verlet) && A.nsteps > 0) { A.acc[i_api] = acc.x; A.acc[X.nm + i_api] = acc.y; A.acc[2 * X.nm + i_api] = acc.z; }
        if (tile == 0 && t == 0) { *A.t_out = t_sim; *A.step_out = *A.step_in + A.nsteps; }
        if (has_act && A.nsteps > 0) {
            uint8_t *blob_tile = const_cast<uint8_t *>(A.blob) + (size_t)M.blob_off16 * 16;
            for (uint32_t q = act_lo + (uint32_t)t; q < act_hi; q += TM) {
                const int4 r = A.act_rec[q];
                const double x = *reinterpret_cast<const double *>(smem + r.x);
                *reinterpret_cast<double *>(blob_tile + r.x) = x;
                if (r.z) A.rest_canon[r.y] = x;
            }
        }
    }
}

static bool resident_env() { static const bool on = !(getenv("TB_RESIDENT") && atoi(getenv("TB_RESIDENT")) == 0); return on; }

// Can the whole system live in one cluster?  (All tiles processed by this engine, none waiting for a peer.)
bool tiled_resident_ok(const Engine *e) {
    const TiledState *S = e->tiled;
    if (!S || !S->supported || (e->flags & TB_NO_RESIDENT) || !resident_env()) return false;
    // one cluster per independent system; beyond a couple of waves of clusters the streaming kernel is the better machine
    // (measured on B200, batches of actuated 10^3 robots: 16 robots = 128 CTAs 6.4 vs 7.2 us per step, 37 robots = 296 CTAs
    // 12.3 vs 7.2 us): resident while every cluster gets an SM of its own
    const int64_t max_tiles = getenv("TB_RESIDENT_MAX_TILES") ? atoll(getenv("TB_RESIDENT_MAX_TILES")) : std::max(8, e->sm_count);
    if (S->n_tiles < 1 || S->res_cluster < 1 || S->n_tiles > max_tiles || S->n_active != S->n_tiles) return false;
    if (e->any_ghost || e->any_magnet || S->n_exp_tab || S->halo_connected) return false;   // the magnet pass runs between steps
    if (e->flags & TB_INTEGRATOR_RK2) return false;     // two launches per step (streaming or generic kernel)
    return true;
}

template <class P, bool D, bool AC>
static int launch_resident_t(Engine *e, TiledState *S, const ResidentArgs &A, const MassExtras &X) {
    auto kern = resident_step_kernel<P, D, AC>;
    {   // per device, see launch_c
        static std::mutex mu;
        static int configured[MAX_DEVICES];
        static bool init = false;
        std::lock_guard<std::mutex> lock(mu);
        if (!init) { for (int &c : configured) c = -1; init = true; }
        const int dv = e->device % MAX_DEVICES;
        if (configured[dv] != A.total || e->device >= MAX_DEVICES) {
            TB_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, A.total));
            configured[dv] = A.total;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)S->n_tiles); cfg.blockDim = dim3(TM); cfg.dynamicSmemBytes = (size_t)A.total; cfg.stream = e->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)S->res_cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    TB_CUDA(e, cudaLaunchKernelEx(&cfg, kern, A, e->G, X));
    return TB_OK;
}

// nsteps steps in one launch; the caller flips e->cur by nsteps & 1 afterwards
int launch_tiled_resident(Engine *e, int nsteps, bool write_acc) {
    TiledState *S = e->tiled;
    if (!tiled_resident_ok(e)) return fail(e, TB_ERR_INVALID, "resident kernel requested for a system it does not fit");
    ResidentArgs A;
    A.nsteps = nsteps;
    A.cs = S->cs; A.vs = S->stage_vel ? S->cs : TM; A.stage_vel = S->stage_vel ? 1 : 0;
    A.dict_mode = S->dict ? (S->dict_split ? 2 : 1) : 0;
    A.verlet = (e->flags & TB_INTEGRATOR_VERLET) ? 1 : 0;
    A.write_acc = write_acc ? 1 : 0;
    A.n_omega = S->has_active ? S->n_omega : 0;
    A.off_halo = (int32_t)up16((size_t)S->max_blob);
    A.off_hptr = A.off_halo + (int32_t)up16((size_t)(S->max_halo + 4) * 4);
    A.off_sintab = A.off_hptr + (int32_t)up16((size_t)TM * 4 + (size_t)TM * (RES_MAX_TILES - 1) * 8);
    A.off_rs = A.off_sintab + MAX_OMEGA * 8;
    A.off_pos = A.off_rs + (int32_t)up16((size_t)((S->dict ? S->max_dict : S->max_rec) + 2) * 8);
    A.off_vel = A.off_pos + 2 * 3 * A.cs * 8;
    A.total = A.off_vel + 2 * 3 * A.vs * 8;
    if ((size_t)A.total > SMEM_LIMIT) return fail(e, TB_ERR_INVALID, "resident kernel: working set exceeds shared memory");
    const int out = e->cur ^ (nsteps & 1);
    A.st_in = S->st[e->cur]; A.st_out = S->st[out];
    A.blob = S->blob; A.slot_to_api = S->slot_to_api; A.halo = S->halo; A.ell = S->ell; A.meta = S->meta;
    A.acc = e->acc;
    A.t_in = e->t_dev + e->cur; A.t_out = e->t_dev + out;
    A.step_in = e->step_dev + e->cur; A.step_out = e->step_dev + out;
    A.omega_tab = S->omega_tab;
    A.act_tile_off = S->act_tile_off; A.act_rec = S->act_rec;
    A.rest_canon = e->rest; A.type = e->type; A.l_max = e->l_max; A.l_min = e->l_min; A.rate = e->rate;
    MassExtras X;
    X.ext = nullptr;                                  // a pending one-shot force is consumed by a streaming step first
    X.drag = e->drag;
    X.lc_ptr = e->any_local ? e->lc_ptr : nullptr;
    X.lc = e->lc;
    X.nm = e->nm;
    const bool d = S->has_damping, a = S->has_active;
    int rc;
    if (e->flags & TB_MATH_STRICT) {
        rc = d ? (a ? launch_resident_t<Strict, true, true>(e, S, A, X) : launch_resident_t<Strict, true, false>(e, S, A, X))
               : (a ? launch_resident_t<Strict, false, true>(e, S, A, X) : launch_resident_t<Strict, false, false>(e, S, A, X));
    } else {
        rc = d ? (a ? launch_resident_t<Fast, true, true>(e, S, A, X) : launch_resident_t<Fast, true, false>(e, S, A, X))
               : (a ? launch_resident_t<Fast, false, true>(e, S, A, X) : launch_resident_t<Fast, false, false>(e, S, A, X));
    }
    if (rc != TB_OK) return rc;
    TB_CUDA(e, cudaGetLastError());
    return TB_OK;
}

static_assert(TM == 128 && STATE_BLOCK == 768, "PosView::ld (engine.h) spells these out");
void tiled_pos_view(const Engine *e, PosView *v) {
    const TiledState *S = e->tiled;
    v->base = S->st[e->cur]; v->api_to_slot = S->api_to_slot; v->nm = e->nm;
}

bool tiled_halo_connected(const Engine *e) { return e->tiled && e->tiled->halo_connected; }

int tiled_halo_check(Engine *e) {
    TiledState *S = e->tiled;
    if (!S || !S->halo_connected || !S->err_dev) return TB_OK;
    int err = 0;
    TB_CUDA(e, cudaMemcpy(&err, S->err_dev, sizeof err, cudaMemcpyDeviceToHost));
    if (err) {
        cudaMemset(S->err_dev, 0, sizeof err);
        return fail(e, TB_ERR_CUDA, "halo exchange: a peer engine did not signal its boundary state in time (are all engines of the partition advanced by the same tb_run_async calls?)");
    }
    return TB_OK;
}

// pushes the current state of every exported mass (tb_halo_push_now)
__global__ void halo_push_kernel(const double *st, const int4 *flat, int64_t n, double *p0, double *p1, double *p2, double *p3,
                                 double *p4, double *p5, double *p6, double *p7) {
    int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double *peers[TB_MAX_PEERS] = {p0, p1, p2, p3, p4, p5, p6, p7};
    const int4 x = flat[j];
    const double *src = st + x.x;
    double *dst = peers[x.y] + x.z;
    for (int c = 0; c < 6; c++) dst[c * TM] = src[c * TM];
    __threadfence_system();
}

}  // namespace tb

using namespace tb;

extern "C" {

int tb_halo_set_ghosts(tb_engine *e, int64_t n, const int32_t *mass_idx) {
    if (!e) return TB_ERR_INVALID;
    cudaSetDevice(e->device);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_halo_set_ghosts before tb_set_topology");
    if (n < 0 || (n > 0 && !mass_idx)) return fail(e, TB_ERR_INVALID, "tb_halo_set_ghosts: bad arguments");
    for (int64_t j = 0; j < n; j++) if (mass_idx[j] < 0 || mass_idx[j] >= e->nm) return fail(e, TB_ERR_INVALID, "tb_halo_set_ghosts: index out of range");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    for (uint8_t &f : e->h_flags) f = (uint8_t)(f & ~F_GHOST);
    for (int64_t j = 0; j < n; j++) e->h_flags[(size_t)mass_idx[j]] |= F_GHOST;
    e->any_ghost = n > 0;
    TB_CUDA(e, cudaMemcpy(e->mflags, e->h_flags.data(), (size_t)e->nm, cudaMemcpyHostToDevice));
    e->tiled_stale |= TS_STRUCT | TS_FLAGS;
    e->graph_dirty = true;
    if (e->tiled) halo_release(e, e->tiled);
    return TB_OK;
}

int tb_halo_prepare(tb_engine *e, tb_halo_handle *out) {
    if (!e || !out) return TB_ERR_INVALID;
    cudaSetDevice(e->device);
    if (e->nm == 0) return fail(e, TB_ERR_INVALID, "tb_halo_prepare before tb_set_topology");
    if (e->tiled) halo_release(e, e->tiled);
    // a layout whose tile table still carries the export fields of an earlier connection is built afresh; one that is current
    // and never exchanged (tb_run_async of zero steps built it just now) is taken as it is
    if (!e->tiled || e->tiled->exports_set || !e->tiled->supported || e->tiled->n_tiles == 0) e->tiled_stale |= TS_STRUCT;
    int rc = tiled_prepare(e);
    if (rc != TB_OK) return rc;
    TiledState *S = e->tiled;
    if (!S->supported) return fail(e, TB_ERR_UNSUPPORTED, "halo exchange needs the tiled layout: %s", S->why_not.c_str());
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (!S->flags_dev) {
        TB_CUDA(e, cudaMalloc((void **)&S->flags_dev, FLAGS_BYTES));
        TB_CUDA(e, cudaMalloc((void **)&S->done_dev, sizeof(unsigned)));
        TB_CUDA(e, cudaMalloc((void **)&S->err_dev, sizeof(int)));
    }
    // flags start at the current step count: "every export up to the present state has landed" (the caller uploads
    // consistent ghost state, or calls tb_halo_push_now)
    std::vector<unsigned long long> init(TB_MAX_PEERS, (unsigned long long)e->steps_done);
    TB_CUDA(e, cudaMemset(S->flags_dev, 0, FLAGS_BYTES));
    TB_CUDA(e, cudaMemcpy(S->flags_dev, init.data(), init.size() * 8, cudaMemcpyHostToDevice));
    TB_CUDA(e, cudaMemset(S->done_dev, 0, sizeof(unsigned)));
    TB_CUDA(e, cudaMemset(S->err_dev, 0, sizeof(int)));
    HandleBlob h;
    memset(&h, 0, sizeof h);
    h.magic = HANDLE_MAGIC; h.pid = (int64_t)getpid(); h.device = e->device; h.cur = e->cur;
    h.steps_done = e->steps_done; h.state_doubles = S->n_tiles * STATE_BLOCK;
    h.raw[0] = S->st[0]; h.raw[1] = S->st[1]; h.raw[2] = S->flags_dev;
    for (int b = 0; b < 3; b++) TB_CUDA(e, cudaIpcGetMemHandle(&h.ipc[b], h.raw[b]));
    memset(out, 0, sizeof *out);
    memcpy(out->bytes, &h, sizeof h);
    S->halo_prepared = true;
    return TB_OK;
}

int tb_halo_state_offsets(tb_engine *e, int64_t n, const int32_t *mass_idx, int64_t *offsets) {
    if (!e) return TB_ERR_INVALID;
    TiledState *S = e->tiled;
    if (!S || !S->halo_prepared) return fail(e, TB_ERR_INVALID, "tb_halo_state_offsets before tb_halo_prepare");
    for (int64_t j = 0; j < n; j++) {
        if (mass_idx[j] < 0 || mass_idx[j] >= e->nm) return fail(e, TB_ERR_INVALID, "tb_halo_state_offsets: index out of range");
        int32_t slot = S->h_api_to_slot[(size_t)mass_idx[j]];
        offsets[j] = (int64_t)(slot / TM) * STATE_BLOCK + slot % TM;
    }
    return TB_OK;
}

int tb_halo_connect(tb_engine *e, int32_t n_peers, const tb_halo_handle *peers, const int32_t *flag_slot_at_peer) {
    if (!e) return TB_ERR_INVALID;
    cudaSetDevice(e->device);
    TiledState *S = e->tiled;
    if (!S || !S->halo_prepared) return fail(e, TB_ERR_INVALID, "tb_halo_connect before tb_halo_prepare");
    if (n_peers < 0 || n_peers > TB_MAX_PEERS || (n_peers > 0 && (!peers || !flag_slot_at_peer)))
        return fail(e, TB_ERR_INVALID, "tb_halo_connect: between 0 and %d peers", TB_MAX_PEERS);
    for (int p = 0; p < n_peers; p++) {
        HandleBlob h;
        memcpy(&h, peers[p].bytes, sizeof h);
        if (h.magic != HANDLE_MAGIC) return fail(e, TB_ERR_INVALID, "tb_halo_connect: peer %d is not a tb_halo_handle", p);
        if (h.cur != e->cur || h.steps_done != e->steps_done)
            return fail(e, TB_ERR_INVALID, "tb_halo_connect: peer %d is at step %lld, this engine at %lld -- engines of a partition advance together", p, (long long)h.steps_done, (long long)e->steps_done);
        if (flag_slot_at_peer[p] < 0 || flag_slot_at_peer[p] >= TB_MAX_PEERS) return fail(e, TB_ERR_INVALID, "tb_halo_connect: bad flag slot");
        void *ptr[3];
        if (h.pid == (int64_t)getpid()) {
            for (int b = 0; b < 3; b++) ptr[b] = h.raw[b];
            if (h.device != e->device) {
                cudaError_t err = cudaDeviceEnablePeerAccess(h.device, 0);
                if (err != cudaSuccess && err != cudaErrorPeerAccessAlreadyEnabled) return fail(e, TB_ERR_CUDA, "no peer access from device %d to %d: %s", e->device, h.device, cudaGetErrorString(err));
                cudaGetLastError();
            }
        } else {
            for (int b = 0; b < 3; b++) {
                TB_CUDA(e, cudaIpcOpenMemHandle(&ptr[b], h.ipc[b], cudaIpcMemLazyEnablePeerAccess));
                S->ipc_opened[p][b] = ptr[b];
            }
        }
        S->peer_st[p][0] = (double *)ptr[0]; S->peer_st[p][1] = (double *)ptr[1];
        S->peer_flag[p] = (unsigned long long *)ptr[2] + flag_slot_at_peer[p];
    }
    S->n_peers = n_peers;
    S->halo_connected = true;
    e->graph_dirty = true;
    return TB_OK;
}

int tb_halo_set_exports(tb_engine *e, int64_t n, const int32_t *mass_idx, const int32_t *peer_idx, const int64_t *remote_offset) {
    if (!e) return TB_ERR_INVALID;
    cudaSetDevice(e->device);
    TiledState *S = e->tiled;
    if (!S || !S->halo_connected) return fail(e, TB_ERR_INVALID, "tb_halo_set_exports before tb_halo_connect");
    if (n < 0 || (n > 0 && (!mass_idx || !peer_idx || !remote_offset))) return fail(e, TB_ERR_INVALID, "tb_halo_set_exports: bad arguments");
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    if (S->exp_tab) { cudaFree(S->exp_tab); S->exp_tab = nullptr; }
    if (S->exp_flat) { cudaFree(S->exp_flat); S->exp_flat = nullptr; }
    std::vector<std::vector<std::pair<int32_t, int32_t>>> per_slot;   // only for slots of tiles that export
    std::vector<int> rows((size_t)S->n_tiles, 0);
    std::vector<int4> flat((size_t)n);
    S->exports_set = true;
    std::map<int64_t, std::vector<std::pair<int32_t, int32_t>>> by_slot;
    for (int64_t j = 0; j < n; j++) {
        if (mass_idx[j] < 0 || mass_idx[j] >= e->nm || peer_idx[j] < 0 || peer_idx[j] >= S->n_peers || remote_offset[j] < 0 ||
            remote_offset[j] >= ((int64_t)1 << 31))
            return fail(e, TB_ERR_INVALID, "tb_halo_set_exports: entry %lld out of range", (long long)j);
        if (e->h_flags[(size_t)mass_idx[j]] & F_GHOST) return fail(e, TB_ERR_INVALID, "tb_halo_set_exports: mass %d is a ghost here", mass_idx[j]);
        int64_t slot = S->h_api_to_slot[(size_t)mass_idx[j]];
        std::vector<std::pair<int32_t, int32_t>> &v = by_slot[slot];
        v.push_back({peer_idx[j], (int32_t)remote_offset[j]});
        rows[(size_t)(slot / TM)] = std::max(rows[(size_t)(slot / TM)], (int)v.size());
        flat[(size_t)j] = make_int4((int)((slot / TM) * STATE_BLOCK + slot % TM), peer_idx[j], (int)remote_offset[j], 0);
    }
    std::vector<int2> tab;
    for (int64_t T = 0; T < S->n_tiles; T++) {
        TileMeta &M = S->h_meta[(size_t)T];
        M.exp_off = (uint32_t)tab.size(); M.exp_rows = (uint32_t)rows[(size_t)T];
        if (rows[(size_t)T] && T >= S->n_boundary)
            return fail(e, TB_ERR_INVALID, "tb_halo_set_exports: an exported mass has no ghost neighbour on this engine (inconsistent partition)");
        tab.resize(tab.size() + (size_t)rows[(size_t)T] * TM, make_int2(-1, 0));
    }
    for (auto &kv : by_slot) {
        const TileMeta &M = S->h_meta[(size_t)(kv.first / TM)];
        for (size_t r = 0; r < kv.second.size(); r++)
            tab[(size_t)M.exp_off + r * TM + (size_t)(kv.first % TM)] = make_int2(kv.second[r].first, kv.second[r].second);
    }
    S->n_exp_tab = (int64_t)tab.size(); S->n_exp_flat = n;
    TB_CUDA(e, cudaMalloc((void **)&S->exp_tab, std::max<size_t>(1, tab.size()) * sizeof(int2)));
    TB_CUDA(e, cudaMalloc((void **)&S->exp_flat, std::max<size_t>(1, flat.size()) * sizeof(int4)));
    if (!tab.empty()) TB_CUDA(e, cudaMemcpy(S->exp_tab, tab.data(), tab.size() * sizeof(int2), cudaMemcpyHostToDevice));
    if (!flat.empty()) TB_CUDA(e, cudaMemcpy(S->exp_flat, flat.data(), flat.size() * sizeof(int4), cudaMemcpyHostToDevice));
    TB_CUDA(e, cudaMemcpy(S->meta, S->h_meta.data(), (size_t)S->n_tiles * sizeof(TileMeta), cudaMemcpyHostToDevice));
    e->graph_dirty = true;
    return TB_OK;
}

int tb_halo_push_now(tb_engine *e) {
    if (!e) return TB_ERR_INVALID;
    cudaSetDevice(e->device);
    TiledState *S = e->tiled;
    if (!S || !S->halo_connected) return fail(e, TB_ERR_INVALID, "tb_halo_push_now before tb_halo_connect");
    int rc = tiled_prepare(e);       // brings an upload that landed in the canonical arrays into the tiled state
    if (rc != TB_OK) return rc;
    if (S->n_exp_flat > 0) {
        double *p[TB_MAX_PEERS];
        for (int q = 0; q < TB_MAX_PEERS; q++) p[q] = S->peer_st[q][e->cur];
        halo_push_kernel<<<(unsigned)((S->n_exp_flat + 127) / 128), 128, 0, e->stream>>>(S->st[e->cur], S->exp_flat, S->n_exp_flat, p[0], p[1],
                                                                                       p[2], p[3], p[4], p[5], p[6], p[7]);
        TB_CUDA(e, cudaGetLastError());
    }
    TB_CUDA(e, cudaStreamSynchronize(e->stream));
    return TB_OK;
}

}  // extern "C"

namespace tb {

int tiled_info(const Engine *e, tb_info *info) {
    const TiledState *S = e->tiled;
    if (!S || !S->supported) return TB_OK;
    info->n_tiles = (int32_t)S->n_tiles;
    info->threads_per_mass = S->tpm;
    info->grid = S->grid;
    info->smem_bytes = (int32_t)S->smem_bytes;
    info->ell_unique_tiles = (int32_t)S->ell_unique_tiles;
    info->dict_records = S->dict ? S->max_dict : 0;
    info->dict_split_k = (S->dict && S->dict_split) ? 1 : 0;
    info->layout_uploads = S->n_uploads; info->layout_patches = S->n_patches;
    // read: blob + own state + halo gathers + halo lists + ELL words + meta; written: own state
    info->tiled_bytes_per_step = (int64_t)S->blob_bytes + 2 * S->n_tiles * STATE_BLOCK * 8 + S->n_halo * (4 + (S->stage_vel ? 48 : 24)) +
                                 S->n_tiles * (int64_t)sizeof(TileMeta) + S->n_ell * 4;
    return TB_OK;
}

}  // namespace tb
