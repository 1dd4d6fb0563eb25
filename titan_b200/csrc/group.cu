// group.cu -- several engines in ONE process behind one handle: a system partitioned over GPUs (boundary-mass halo exchange
// inside the step kernel, tiled.cu) or a batch of independent simulations sharded over GPUs (no communication).
// SURVEY.md 8(b) lists tb_create_partitioned / tb_batch_create among the entry points the C++ facade needs; the reference
// has nothing of the kind (one device, src/sim.cu:1704-1841).  Host code only: everything goes through the tb_* entry
// points of engine.cu / tiled.cu; peers in one process reach each other's state with cudaDeviceEnablePeerAccess
// (tb_halo_connect), no CUDA IPC.
//
// The caller describes the system in GLOBAL numbering plus an owner (engine index) per mass.  Engine r holds, in ascending
// global index, its own masses and ghost copies of the foreign endpoints of its cut springs, and every spring with at least
// one own endpoint, in global spring order -- so the per-mass summation order equals the single-engine one (bit-identical
// under TB_MATH_STRICT).  This is titan_b200/partition.py::GraphPartition + wire() in C++.
#include "engine.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

struct tb_group {
    std::vector<tb_engine *> eng;
    std::vector<int> device;
    uint32_t flags = 0;
    std::string err;
    int64_t nm = 0, ns = 0;
    std::vector<int32_t> owner;                 // [nm] engine index per mass
    std::vector<int32_t> left, right;           // global topology
    struct Part {
        std::vector<int64_t> mass_gid, spring_gid;      // ascending global ids of the local masses / springs
        std::vector<int32_t> ghost_local, own_local;    // local indices
        std::vector<int32_t> ghost_owner;               // per ghost_local entry
    };
    std::vector<Part> part;
    std::vector<int64_t> batch_sizes;           // tb_batch_create: masses per simulation (owner == NULL -> whole simulations per engine)
    bool connected = false, need_push = false, any_cut = false;
    double last_ms = 0.0;
};

namespace {

int gfail(tb_group *g, int status, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (g) g->err = buf;
    return status;
}
// an engine call failed: its message becomes the group's
int efail(tb_group *g, size_t r, int rc) {
    g->err = "engine " + std::to_string(r) + " (device " + std::to_string(g->device[r]) + "): " + tb_last_error(g->eng[r]);
    return rc;
}
#define G_TRY(g, r, call) do { int _rc = (call); if (_rc != TB_OK) return efail((g), (r), _rc); } while (0)

// one host thread per engine (each talks to its own device), or chunks of a loop over all host cores
template <class F> void per_engine(size_t W, F f) {
    if (W == 1) { f((size_t)0); return; }
    std::vector<std::thread> th;
    for (size_t r = 0; r < W; r++) th.emplace_back([&f, r] { f(r); });
    for (std::thread &t : th) t.join();
}
template <class F> void par_chunks(size_t n, F f) {
    const size_t nt = n < ((size_t)1 << 16) ? 1 : std::min<size_t>(16, std::max(1u, std::thread::hardware_concurrency()));
    if (nt == 1) { f((size_t)0, n); return; }
    std::vector<std::thread> th;
    for (size_t t = 0; t < nt; t++) th.emplace_back([&f, t, nt, n] { f(n * t / nt, n * (t + 1) / nt); });
    for (std::thread &t : th) t.join();
}

inline int32_t local_of(const tb_group::Part &p, int64_t gid) {
    auto it = std::lower_bound(p.mass_gid.begin(), p.mass_gid.end(), gid);
    return (it != p.mass_gid.end() && *it == gid) ? (int32_t)(it - p.mass_gid.begin()) : -1;
}
inline int32_t local_spring_of(const tb_group::Part &p, int64_t gid) {
    auto it = std::lower_bound(p.spring_gid.begin(), p.spring_gid.end(), gid);
    return (it != p.spring_gid.end() && *it == gid) ? (int32_t)(it - p.spring_gid.begin()) : -1;
}

// Global selection (range or index list) -> per engine: local indices + positions in the caller's arrays.
struct Route { std::vector<int32_t> local; std::vector<int64_t> at; };
void route_masses(const tb_group *g, int64_t first, int64_t count, const int32_t *idx, std::vector<Route> &out) {
    out.assign(g->eng.size(), Route());
    if (!idx) {
        for (size_t r = 0; r < g->eng.size(); r++) {
            const tb_group::Part &p = g->part[r];
            for (int32_t l : p.own_local) {
                const int64_t gid = p.mass_gid[(size_t)l];
                if (gid >= first && gid < first + count) { out[r].local.push_back(l); out[r].at.push_back(gid - first); }
            }
        }
    } else {
        for (int64_t j = 0; j < count; j++) {
            const size_t r = (size_t)g->owner[(size_t)idx[j]];
            out[r].local.push_back(local_of(g->part[r], idx[j]));
            out[r].at.push_back(j);
        }
    }
}
// springs live on every engine that owns one of their endpoints
void route_springs(const tb_group *g, int64_t first, int64_t count, const int32_t *idx, std::vector<Route> &out) {
    out.assign(g->eng.size(), Route());
    if (!idx) {
        for (size_t r = 0; r < g->eng.size(); r++) {
            const tb_group::Part &p = g->part[r];
            auto lo = std::lower_bound(p.spring_gid.begin(), p.spring_gid.end(), first);
            auto hi = std::lower_bound(p.spring_gid.begin(), p.spring_gid.end(), first + count);
            for (auto it = lo; it != hi; ++it) { out[r].local.push_back((int32_t)(it - p.spring_gid.begin())); out[r].at.push_back(*it - first); }
        }
    } else {
        for (int64_t j = 0; j < count; j++) {
            const int64_t s = idx[j];
            int ra = g->left[(size_t)s] >= 0 ? g->owner[(size_t)g->left[(size_t)s]] : -1;
            int rb = g->right[(size_t)s] >= 0 ? g->owner[(size_t)g->right[(size_t)s]] : -1;
            for (int r : {ra, rb == ra ? -1 : rb}) {
                if (r < 0) continue;
                const int32_t l = local_spring_of(g->part[(size_t)r], s);
                if (l >= 0) { out[(size_t)r].local.push_back(l); out[(size_t)r].at.push_back(j); }
            }
        }
    }
}
template <class T> const T *pick(const T *src, const Route &rt, int width, std::vector<T> &tmp) {
    if (!src) return nullptr;
    tmp.resize(rt.at.size() * (size_t)width);
    for (size_t q = 0; q < rt.at.size(); q++)
        for (int c = 0; c < width; c++) tmp[q * (size_t)width + (size_t)c] = src[(size_t)rt.at[q] * (size_t)width + (size_t)c];
    return tmp.data();
}
template <class T> void unpick(T *dst, const Route &rt, int width, const std::vector<T> &tmp) {
    if (!dst) return;
    for (size_t q = 0; q < rt.at.size(); q++)
        for (int c = 0; c < width; c++) dst[(size_t)rt.at[q] * (size_t)width + (size_t)c] = tmp[q * (size_t)width + (size_t)c];
}

}  // namespace

namespace {

// Halo handshake between the engines of the group (titan_b200/partition.py::connect, in-process form): prepare every engine,
// tell every owner where its peers keep the ghost copies, open the peers' buffers, then push the current boundary state.
int group_connect(tb_group *g) {
    const size_t W = g->eng.size();
    g->connected = false;
    if (!g->any_cut) { g->connected = true; return TB_OK; }      // independent pieces: nothing to exchange
    std::vector<tb_halo_handle> handle(W);
    std::vector<std::vector<int64_t>> ghost_off(W);
    for (size_t r = 0; r < W; r++) {
        G_TRY(g, r, tb_halo_prepare(g->eng[r], &handle[r]));
        const tb_group::Part &p = g->part[r];
        ghost_off[r].resize(p.ghost_local.size());
        if (!p.ghost_local.empty())
            G_TRY(g, r, tb_halo_state_offsets(g->eng[r], (int64_t)p.ghost_local.size(), p.ghost_local.data(), ghost_off[r].data()));
    }
    // peers of r: owners of its ghosts and engines holding ghosts it owns, ascending
    std::vector<std::vector<int32_t>> peers(W);
    for (size_t r = 0; r < W; r++) {
        std::vector<uint8_t> is(W, 0);
        for (int32_t o : g->part[r].ghost_owner) is[(size_t)o] = 1;
        for (size_t q = 0; q < W; q++)
            if (q != r) for (int32_t o : g->part[q].ghost_owner) if ((size_t)o == r) { is[q] = 1; break; }
        for (size_t q = 0; q < W; q++) if (is[q] && q != r) peers[r].push_back((int32_t)q);
        if (peers[r].size() > TB_MAX_PEERS) return gfail(g, TB_ERR_UNSUPPORTED, "engine %zu exchanges boundary state with %zu peers (at most %d)", r, peers[r].size(), TB_MAX_PEERS);
    }
    for (size_t r = 0; r < W; r++) {
        std::vector<tb_halo_handle> hs;
        std::vector<int32_t> flag_slot;
        for (int32_t q : peers[r]) {
            hs.push_back(handle[(size_t)q]);
            const std::vector<int32_t> &pq = peers[(size_t)q];
            flag_slot.push_back((int32_t)(std::find(pq.begin(), pq.end(), (int32_t)r) - pq.begin()));
        }
        G_TRY(g, r, tb_halo_connect(g->eng[r], (int32_t)hs.size(), hs.data(), flag_slot.data()));
        // exports: for every peer, its ghosts that this engine owns, in the peer's ghost order
        std::vector<int32_t> e_idx, e_peer;
        std::vector<int64_t> e_off;
        for (size_t j = 0; j < peers[r].size(); j++) {
            const size_t q = (size_t)peers[r][j];
            const tb_group::Part &pq = g->part[q];
            for (size_t h = 0; h < pq.ghost_local.size(); h++) {
                if ((size_t)pq.ghost_owner[h] != r) continue;
                const int32_t loc = local_of(g->part[r], pq.mass_gid[(size_t)pq.ghost_local[h]]);
                if (loc < 0) return gfail(g, TB_ERR_INVALID, "internal: engine %zu holds a ghost of a mass engine %zu does not have", q, r);
                e_idx.push_back(loc); e_peer.push_back((int32_t)j); e_off.push_back(ghost_off[q][h]);
            }
        }
        G_TRY(g, r, tb_halo_set_exports(g->eng[r], (int64_t)e_idx.size(), e_idx.data(), e_peer.data(), e_off.data()));
    }
    g->connected = true;
    g->need_push = true;
    return TB_OK;
}

int group_push(tb_group *g) {
    if (!g->any_cut) { g->need_push = false; return TB_OK; }
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_halo_push_now(g->eng[r]));     // returns when the stores are visible
    g->need_push = false;
    return TB_OK;
}

}  // namespace

extern "C" {

int tb_create_partitioned(int32_t ndev, const int32_t *devices, uint32_t flags, tb_group **out) {
    if (!out) return TB_ERR_INVALID;
    *out = nullptr;
    if (ndev < 1 || ndev > 64 || !devices) return TB_ERR_INVALID;
    tb_group *g = new (std::nothrow) tb_group();
    if (!g) return TB_ERR_OOM;
    g->flags = flags;
    for (int32_t r = 0; r < ndev; r++) {
        tb_engine *e = nullptr;
        int rc = tb_create(devices[r], flags, &e);
        if (rc != TB_OK) { for (tb_engine *x : g->eng) tb_destroy(x); delete g; return rc; }     // tb_last_error(NULL) has the reason
        g->eng.push_back(e); g->device.push_back(devices[r]);
    }
    g->part.resize((size_t)ndev);
    *out = g;
    return TB_OK;
}

int tb_batch_create(int32_t ndev, const int32_t *devices, uint32_t flags, int64_t nsims, const int64_t *masses_per_sim, tb_group **out) {
    if (nsims < 1 || !masses_per_sim) return TB_ERR_INVALID;
    int rc = tb_create_partitioned(ndev, devices, flags, out);
    if (rc != TB_OK) return rc;
    (*out)->batch_sizes.assign(masses_per_sim, masses_per_sim + nsims);
    return TB_OK;
}

void tb_group_destroy(tb_group *g) {
    if (!g) return;
    for (tb_engine *e : g->eng) tb_sync(e);      // no engine may still be storing into a peer that is about to go
    for (tb_engine *e : g->eng) tb_destroy(e);
    delete g;
}

const char *tb_group_last_error(const tb_group *g) { return g ? g->err.c_str() : tb_last_error(nullptr); }
int32_t tb_group_size(const tb_group *g) { return g ? (int32_t)g->eng.size() : 0; }
tb_engine *tb_group_engine(tb_group *g, int32_t r) { return (g && r >= 0 && (size_t)r < g->eng.size()) ? g->eng[(size_t)r] : nullptr; }

int tb_group_set_topology(tb_group *g, int64_t nm, int64_t ns, const int32_t *left, const int32_t *right, const int32_t *owner) {
    if (!g) return TB_ERR_INVALID;
    tb::TimeScope ts("tb_group_set_topology");
    const size_t W = g->eng.size();
    if (nm <= 0 || ns < 0 || (ns > 0 && (!left || !right))) return gfail(g, TB_ERR_INVALID, "tb_group_set_topology: bad sizes/pointers");
    g->nm = nm; g->ns = ns;
    g->left.assign(left, left + ns); g->right.assign(right, right + ns);
    g->owner.assign((size_t)nm, 0);
    if (owner) {
        for (int64_t i = 0; i < nm; i++) {
            if (owner[i] < 0 || (size_t)owner[i] >= W) return gfail(g, TB_ERR_INVALID, "tb_group_set_topology: owner of mass %lld out of range", (long long)i);
            g->owner[(size_t)i] = owner[i];
        }
    } else if (!g->batch_sizes.empty()) {
        // whole simulations per engine, contiguous blocks balanced by mass count
        int64_t total = 0;
        for (int64_t n : g->batch_sizes) total += n;
        if (total != nm) return gfail(g, TB_ERR_INVALID, "tb_group_set_topology: the batch holds %lld masses, the topology %lld", (long long)total, (long long)nm);
        int64_t at = 0, r = 0;
        for (int64_t n : g->batch_sizes) {
            while (r + 1 < (int64_t)W && at * (int64_t)W >= (r + 1) * total) r++;
            std::fill(g->owner.begin() + at, g->owner.begin() + at + n, (int32_t)r);
            at += n;
        }
    } else {
        for (int64_t i = 0; i < nm; i++) g->owner[(size_t)i] = (int32_t)std::min<int64_t>((int64_t)W - 1, i * (int64_t)W / nm);
    }
    // per engine: springs with an own endpoint (global order), masses touched by them plus own masses without springs
    g->any_cut = false;
    std::vector<std::vector<int32_t>> LL(W), RR(W);
    std::vector<uint8_t> cut(W, 0);
    per_engine(W, [&](size_t r) {            // host work only, one thread per engine
        tb_group::Part &p = g->part[r];
        p = tb_group::Part();
        std::vector<int32_t> g2l((size_t)nm, -1);            // -1 untouched, then the local index
        for (int64_t i = 0; i < nm; i++) if ((size_t)g->owner[(size_t)i] == r) g2l[(size_t)i] = 0;
        for (int64_t s = 0; s < ns; s++) {
            const int32_t a = left[s], b = right[s];
            const bool mine = (a >= 0 && (size_t)g->owner[(size_t)a] == r) || (b >= 0 && (size_t)g->owner[(size_t)b] == r);
            if (!mine) continue;
            p.spring_gid.push_back(s);
            if (a >= 0) g2l[(size_t)a] = 0;
            if (b >= 0) g2l[(size_t)b] = 0;
        }
        for (int64_t i = 0; i < nm; i++) if (g2l[(size_t)i] == 0) { g2l[(size_t)i] = (int32_t)p.mass_gid.size(); p.mass_gid.push_back(i); }
        std::vector<int32_t> &ll = LL[r], &rr = RR[r];
        ll.resize(p.spring_gid.size()); rr.resize(p.spring_gid.size());
        for (size_t q = 0; q < p.spring_gid.size(); q++) {
            const int64_t s = p.spring_gid[q];
            ll[q] = left[s] >= 0 ? g2l[(size_t)left[s]] : -1;
            rr[q] = right[s] >= 0 ? g2l[(size_t)right[s]] : -1;
        }
        for (size_t l = 0; l < p.mass_gid.size(); l++) {
            const int32_t o = g->owner[(size_t)p.mass_gid[l]];
            if ((size_t)o == r) p.own_local.push_back((int32_t)l);
            else { p.ghost_local.push_back((int32_t)l); p.ghost_owner.push_back(o); cut[r] = 1; }
        }
    });
    for (size_t r = 0; r < W; r++) {
        tb_group::Part &p = g->part[r];
        if (cut[r]) g->any_cut = true;
        if (p.mass_gid.empty()) return gfail(g, TB_ERR_INVALID, "tb_group_set_topology: engine %zu owns no mass (fewer masses than engines?)", r);
    }
    std::vector<int> rcs(W, TB_OK);
    per_engine(W, [&](size_t r) {
        tb_group::Part &p = g->part[r];
        rcs[r] = tb_set_topology(g->eng[r], (int64_t)p.mass_gid.size(), (int64_t)p.spring_gid.size(), LL[r].data(), RR[r].data());
        if (rcs[r] == TB_OK) rcs[r] = tb_halo_set_ghosts(g->eng[r], (int64_t)p.ghost_local.size(), p.ghost_local.empty() ? nullptr : p.ghost_local.data());
    });
    for (size_t r = 0; r < W; r++) if (rcs[r] != TB_OK) return efail(g, r, rcs[r]);
    g->connected = false; g->need_push = true;
    return TB_OK;
}

int tb_group_set_tile_hint(tb_group *g, const int32_t *tile_of_mass) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    for (size_t r = 0; r < g->eng.size(); r++) {
        if (!tile_of_mass) { G_TRY(g, r, tb_set_tile_hint(g->eng[r], nullptr)); continue; }
        const tb_group::Part &p = g->part[r];
        std::vector<int32_t> h(p.mass_gid.size());
        for (size_t l = 0; l < h.size(); l++) h[l] = tile_of_mass[(size_t)p.mass_gid[l]];
        G_TRY(g, r, tb_set_tile_hint(g->eng[r], h.data()));
    }
    g->connected = false;
    return TB_OK;
}

}  // extern "C"

extern "C" {

// Uploads go to the OWNER of a mass; the ghost copies on other engines are refreshed from the owners (tb_halo_push_now)
// before the next run.  Same field semantics as tb_set_mass_state.
int tb_group_set_mass_state(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *pos, const double *vel,
                            const double *m, const double *T, const double *extern_force, const uint8_t *fixed, const int32_t *drag,
                            const uint8_t *valid) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    if (count < 0 || (!idx && (first < 0 || first + count > g->nm))) return gfail(g, TB_ERR_INVALID, "tb_group_set_mass_state: range outside [0,%lld)", (long long)g->nm);
    if (idx) for (int64_t j = 0; j < count; j++) if (idx[j] < 0 || idx[j] >= g->nm) return gfail(g, TB_ERR_INVALID, "tb_group_set_mass_state: index %d out of range", idx[j]);
    std::vector<Route> rt;
    route_masses(g, first, count, idx, rt);
    std::vector<double> tp, tv, tm, tT, te;
    std::vector<uint8_t> tf, tval;
    std::vector<int32_t> td;
    for (size_t r = 0; r < g->eng.size(); r++) {
        if (rt[r].local.empty()) continue;
        G_TRY(g, r, tb_set_mass_state(g->eng[r], 0, (int64_t)rt[r].local.size(), rt[r].local.data(), pick(pos, rt[r], 3, tp), pick(vel, rt[r], 3, tv),
                                      pick(m, rt[r], 1, tm), pick(T, rt[r], 1, tT), pick(extern_force, rt[r], 3, te), pick(fixed, rt[r], 1, tf),
                                      pick(drag, rt[r], 1, td), pick(valid, rt[r], 1, tval)));
    }
    // a ghost copy carries the mass and the flags of its owner too (m scales nothing on a ghost, but fixed / valid decide which
    // springs the tile builder keeps): mirror those fields; the state itself arrives by the push
    if (m || fixed || valid) {
        std::vector<std::pair<int64_t, int64_t>> where;      // (global id, position in the caller's arrays), index-list uploads
        if (idx) {
            for (int64_t j = 0; j < count; j++) where.push_back({idx[j], j});
            std::sort(where.begin(), where.end());
        }
        for (size_t r = 0; r < g->eng.size(); r++) {
            const tb_group::Part &p = g->part[r];
            Route gr;
            for (int32_t l : p.ghost_local) {
                const int64_t gid = p.mass_gid[(size_t)l];
                int64_t at = -1;
                if (!idx) { if (gid >= first && gid < first + count) at = gid - first; }
                else {
                    auto it = std::lower_bound(where.begin(), where.end(), std::make_pair(gid, (int64_t)-1));
                    if (it != where.end() && it->first == gid) at = it->second;
                }
                if (at >= 0) { gr.local.push_back(l); gr.at.push_back(at); }
            }
            if (gr.local.empty()) continue;
            G_TRY(g, r, tb_set_mass_state(g->eng[r], 0, (int64_t)gr.local.size(), gr.local.data(), pick(pos, gr, 3, tp), pick(vel, gr, 3, tv),
                                          pick(m, gr, 1, tm), nullptr, nullptr, pick(fixed, gr, 1, tf), nullptr, pick(valid, gr, 1, tval)));
        }
        if (valid) g->connected = false;      // validity changes rebuild the tile structure of the engines involved
    }
    g->need_push = true;
    return TB_OK;
}

int tb_group_get_mass_state(tb_group *g, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel, double *acc,
                            double *T, double *extern_force) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    if (count < 0 || (!idx && (first < 0 || first + count > g->nm))) return gfail(g, TB_ERR_INVALID, "tb_group_get_mass_state: range outside [0,%lld)", (long long)g->nm);
    if (idx) for (int64_t j = 0; j < count; j++) if (idx[j] < 0 || idx[j] >= g->nm) return gfail(g, TB_ERR_INVALID, "tb_group_get_mass_state: index %d out of range", idx[j]);
    std::vector<Route> rt;
    route_masses(g, first, count, idx, rt);
    for (size_t r = 0; r < g->eng.size(); r++) {
        const size_t n = rt[r].local.size();
        if (!n) continue;
        std::vector<double> tp(pos ? 3 * n : 0), tv(vel ? 3 * n : 0), ta(acc ? 3 * n : 0), tT(T ? n : 0), te(extern_force ? 3 * n : 0);
        G_TRY(g, r, tb_get_mass_state(g->eng[r], 0, (int64_t)n, rt[r].local.data(), pos ? tp.data() : nullptr, vel ? tv.data() : nullptr,
                                      acc ? ta.data() : nullptr, T ? tT.data() : nullptr, extern_force ? te.data() : nullptr));
        unpick(pos, rt[r], 3, tp); unpick(vel, rt[r], 3, tv); unpick(acc, rt[r], 3, ta); unpick(T, rt[r], 1, tT); unpick(extern_force, rt[r], 3, te);
    }
    return TB_OK;
}

int tb_group_set_mass_magnet(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *rad, const double *stiffness,
                             const double *max_mag_force, const double *mag_scale) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    std::vector<Route> rt;
    route_masses(g, first, count, idx, rt);
    std::vector<double> a, b, c, d;
    for (size_t r = 0; r < g->eng.size(); r++) {
        if (rt[r].local.empty()) continue;
        G_TRY(g, r, tb_set_mass_magnet(g->eng[r], 0, (int64_t)rt[r].local.size(), rt[r].local.data(), pick(rad, rt[r], 1, a),
                                       pick(stiffness, rt[r], 1, b), pick(max_mag_force, rt[r], 1, c), pick(mag_scale, rt[r], 1, d)));
    }
    return TB_OK;
}

// every engine that holds a spring gets its parameters (a cut spring is evaluated by both owners)
int tb_group_set_spring_params(tb_group *g, int64_t first, int64_t count, const int32_t *idx, const double *k, const double *rest,
                               const double *damping, const int32_t *type, const double *omega, const double *l_max,
                               const double *l_min, const double *rate) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    tb::TimeScope ts("tb_group_set_spring_params");
    if (count < 0 || (!idx && (first < 0 || first + count > g->ns))) return gfail(g, TB_ERR_INVALID, "tb_group_set_spring_params: range outside [0,%lld)", (long long)g->ns);
    if (idx) for (int64_t j = 0; j < count; j++) if (idx[j] < 0 || idx[j] >= g->ns) return gfail(g, TB_ERR_INVALID, "tb_group_set_spring_params: index %d out of range", idx[j]);
    if (!idx) {
        // a range of global springs is a range of every engine's local springs (ascending global order on both sides): gather
        // the values on all host cores and upload them as a range -- no index lists (43 M springs: 3.2 s -> a fraction)
        std::vector<double> tmp[7];
        std::vector<int32_t> ty;
        for (size_t r = 0; r < g->eng.size(); r++) {
            const tb_group::Part &p = g->part[r];
            const size_t lo = (size_t)(std::lower_bound(p.spring_gid.begin(), p.spring_gid.end(), first) - p.spring_gid.begin());
            const size_t hi = (size_t)(std::lower_bound(p.spring_gid.begin(), p.spring_gid.end(), first + count) - p.spring_gid.begin());
            if (hi <= lo) continue;
            const double *src[7] = {k, rest, damping, omega, l_max, l_min, rate};
            const double *use[7];
            for (int a = 0; a < 7; a++) {
                use[a] = nullptr;
                if (!src[a]) continue;
                tmp[a].resize(hi - lo);
                double *dst = tmp[a].data(); const double *from = src[a];
                par_chunks(hi - lo, [&](size_t q0, size_t q1) { for (size_t q = q0; q < q1; q++) dst[q] = from[p.spring_gid[lo + q] - first]; });
                use[a] = dst;
            }
            if (type) {
                ty.resize(hi - lo);
                par_chunks(hi - lo, [&](size_t q0, size_t q1) { for (size_t q = q0; q < q1; q++) ty[q] = type[p.spring_gid[lo + q] - first]; });
            }
            G_TRY(g, r, tb_set_spring_params(g->eng[r], (int64_t)lo, (int64_t)(hi - lo), nullptr, use[0], use[1], use[2], type ? ty.data() : nullptr,
                                             use[3], use[4], use[5], use[6]));
        }
        if (type || damping) g->connected = false;
        return TB_OK;
    }
    std::vector<Route> rt;
    route_springs(g, first, count, idx, rt);
    std::vector<double> a, b, c, d, e2, f, h;
    std::vector<int32_t> ty;
    for (size_t r = 0; r < g->eng.size(); r++) {
        if (rt[r].local.empty()) continue;
        G_TRY(g, r, tb_set_spring_params(g->eng[r], 0, (int64_t)rt[r].local.size(), rt[r].local.data(), pick(k, rt[r], 1, a), pick(rest, rt[r], 1, b),
                                         pick(damping, rt[r], 1, c), pick(type, rt[r], 1, ty), pick(omega, rt[r], 1, d), pick(l_max, rt[r], 1, e2),
                                         pick(l_min, rt[r], 1, f), pick(rate, rt[r], 1, h)));
    }
    if (type || damping) g->connected = false;    // may change the tile structure (actuation classes, staged velocities): handshake again
    return TB_OK;
}

int tb_group_get_spring_rest(tb_group *g, int64_t first, int64_t count, const int32_t *idx, double *rest) {
    if (!g || g->nm == 0 || !rest) return TB_ERR_INVALID;
    std::vector<Route> rt;
    route_springs(g, first, count, idx, rt);
    for (size_t r = 0; r < g->eng.size(); r++) {       // both owners of a cut spring hold the same value
        const size_t n = rt[r].local.size();
        if (!n) continue;
        std::vector<double> t(n);
        G_TRY(g, r, tb_get_spring_rest(g->eng[r], 0, (int64_t)n, rt[r].local.data(), t.data()));
        unpick(rest, rt[r], 1, t);
    }
    return TB_OK;
}

int tb_group_remove_springs(tb_group *g, int64_t n, const int32_t *idx) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    std::vector<Route> rt;
    route_springs(g, 0, n, idx, rt);
    for (size_t r = 0; r < g->eng.size(); r++)
        if (!rt[r].local.empty()) G_TRY(g, r, tb_remove_springs(g->eng[r], (int64_t)rt[r].local.size(), rt[r].local.data()));
    return TB_OK;
}

int tb_group_set_planes(tb_group *g, int32_t n, const tb_plane *planes) {
    if (!g) return TB_ERR_INVALID;
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_set_planes(g->eng[r], n, planes));
    return TB_OK;
}
int tb_group_set_balls(tb_group *g, int32_t n, const tb_ball *balls) {
    if (!g) return TB_ERR_INVALID;
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_set_balls(g->eng[r], n, balls));
    return TB_OK;
}
int tb_group_set_globals(tb_group *g, const double gravity[3], double dt) {
    if (!g) return TB_ERR_INVALID;
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_set_globals(g->eng[r], gravity, dt));
    return TB_OK;
}

// All engines advance by the same number of steps; the launches are asynchronous, so one host thread feeds every device and
// the engines meet inside their step kernels (peer flags), not on the host.
int tb_group_run_async(tb_group *g, int64_t nsteps, double t0) {
    if (!g || g->nm == 0) return TB_ERR_INVALID;
    // layouts first (a tile build allocates, which waits for its whole device), then the handshake, then the boundary state
    {   tb::TimeScope ts("group layouts");
        // (each engine builds its layout on its own host thread: the builds are host work, a second or more per million masses)
        std::vector<int> rcs(g->eng.size(), TB_OK);
        per_engine(g->eng.size(), [&](size_t r) { rcs[r] = tb_run_async(g->eng[r], 0, t0); });
        for (size_t r = 0; r < g->eng.size(); r++) if (rcs[r] != TB_OK) return efail(g, r, rcs[r]);
    }
    if (!g->connected) { tb::TimeScope ts("group_connect"); int rc = group_connect(g); if (rc != TB_OK) return rc; }
    if (g->need_push) { tb::TimeScope ts("group_push"); int rc = group_push(g); if (rc != TB_OK) return rc; }
    if (nsteps == 0) return TB_OK;
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_run_async(g->eng[r], nsteps, t0));
    return TB_OK;
}

int tb_group_sync(tb_group *g) {
    if (!g) return TB_ERR_INVALID;
    g->last_ms = 0.0;
    for (size_t r = 0; r < g->eng.size(); r++) {
        G_TRY(g, r, tb_sync(g->eng[r]));
        double ms = 0.0;
        if (tb_last_run_ms(g->eng[r], &ms) == TB_OK) g->last_ms = std::max(g->last_ms, ms);
    }
    return TB_OK;
}

int tb_group_wait_steps(tb_group *g, int64_t steps) {
    if (!g) return TB_ERR_INVALID;
    for (size_t r = 0; r < g->eng.size(); r++) G_TRY(g, r, tb_wait_steps(g->eng[r], steps));
    return TB_OK;
}

int tb_group_last_run_ms(tb_group *g, double *ms) {
    if (!g || !ms) return TB_ERR_INVALID;
    *ms = g->last_ms;
    return TB_OK;
}

int tb_group_steps_done(tb_group *g, int64_t *steps) {
    if (!g || g->eng.empty()) return TB_ERR_INVALID;
    return tb_steps_done(g->eng[0], steps);
}

}  // extern "C"
