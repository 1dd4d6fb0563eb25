// sim.cpp -- titan::Simulation on top of the C ABI of libtitan_b200 (include/titan_b200.h).
//
// Layout of this file: (1) object bookkeeping (create/delete, builders, bulk setters -- host only),
// (2) the packer: host object graph <-> the engine's structure-of-arrays uploads/downloads,
// (3) lifecycle and the GPU thread.  Reference behaviour cited per function (paths relative to the reference).
#include <Titan/sim.h>

#include <titan_b200.h>

#include <algorithm>
#include <array>
#include <cassert>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <thread>
#include <unordered_set>

namespace titan {

// One GPU or several behind the same calls: every engine call of this file goes through a Dev, which forwards to the single
// engine or to the group (include/titan_b200.h, tb_group_*: the same signatures with global indices).
struct Dev { tb_engine *e; tb_group *g; };
namespace {
inline const char *dv_last_error(Dev d) { return d.g ? tb_group_last_error(d.g) : tb_last_error(d.e); }
inline int dv_set_topology(Dev d, int64_t nm, int64_t ns, const int32_t *l, const int32_t *r, const int32_t *owner) {
    return d.g ? tb_group_set_topology(d.g, nm, ns, l, r, owner) : tb_set_topology(d.e, nm, ns, l, r);
}
inline int dv_set_tile_hint(Dev d, const int32_t *h) { return d.g ? tb_group_set_tile_hint(d.g, h) : tb_set_tile_hint(d.e, h); }
inline int dv_set_mass_state(Dev d, int64_t first, int64_t count, const int32_t *idx, const double *pos, const double *vel, const double *m,
                             const double *T, const double *ext, const uint8_t *fixed, const int32_t *drag, const uint8_t *valid) {
    return d.g ? tb_group_set_mass_state(d.g, first, count, idx, pos, vel, m, T, ext, fixed, drag, valid)
               : tb_set_mass_state(d.e, first, count, idx, pos, vel, m, T, ext, fixed, drag, valid);
}
inline int dv_set_mass_magnet(Dev d, int64_t first, int64_t count, const int32_t *idx, const double *a, const double *b, const double *c, const double *e) {
    return d.g ? tb_group_set_mass_magnet(d.g, first, count, idx, a, b, c, e) : tb_set_mass_magnet(d.e, first, count, idx, a, b, c, e);
}
inline int dv_get_mass_state(Dev d, int64_t first, int64_t count, const int32_t *idx, double *pos, double *vel, double *acc, double *T, double *ext) {
    return d.g ? tb_group_get_mass_state(d.g, first, count, idx, pos, vel, acc, T, ext) : tb_get_mass_state(d.e, first, count, idx, pos, vel, acc, T, ext);
}
inline int dv_set_spring_params(Dev d, int64_t first, int64_t count, const int32_t *idx, const double *k, const double *rest, const double *damping,
                                const int32_t *type, const double *omega, const double *lmax, const double *lmin, const double *rate) {
    return d.g ? tb_group_set_spring_params(d.g, first, count, idx, k, rest, damping, type, omega, lmax, lmin, rate)
               : tb_set_spring_params(d.e, first, count, idx, k, rest, damping, type, omega, lmax, lmin, rate);
}
inline int dv_get_spring_rest(Dev d, int64_t first, int64_t count, const int32_t *idx, double *rest) {
    return d.g ? tb_group_get_spring_rest(d.g, first, count, idx, rest) : tb_get_spring_rest(d.e, first, count, idx, rest);
}
inline int dv_remove_springs(Dev d, int64_t n, const int32_t *idx) { return d.g ? tb_group_remove_springs(d.g, n, idx) : tb_remove_springs(d.e, n, idx); }
inline int dv_set_planes(Dev d, int32_t n, const tb_plane *p) { return d.g ? tb_group_set_planes(d.g, n, p) : tb_set_planes(d.e, n, p); }
inline int dv_set_balls(Dev d, int32_t n, const tb_ball *b) { return d.g ? tb_group_set_balls(d.g, n, b) : tb_set_balls(d.e, n, b); }
inline int dv_set_globals(Dev d, const double g[3], double dt) { return d.g ? tb_group_set_globals(d.g, g, dt) : tb_set_globals(d.e, g, dt); }
inline int dv_run_async(Dev d, int64_t n, double t0) { return d.g ? tb_group_run_async(d.g, n, t0) : tb_run_async(d.e, n, t0); }
inline int dv_sync(Dev d) { return d.g ? tb_group_sync(d.g) : tb_sync(d.e); }
inline int dv_wait_steps(Dev d, int64_t n) { return d.g ? tb_group_wait_steps(d.g, n) : tb_wait_steps(d.e, n); }
inline int dv_steps_done(Dev d, int64_t *n) { return d.g ? tb_group_steps_done(d.g, n) : tb_steps_done(d.e, n); }
inline int dv_alloc_host(Dev d, size_t bytes, void **out) { return tb_alloc_host(d.g ? tb_group_engine(d.g, 0) : d.e, bytes, out); }
}  // namespace

namespace {
const char *kEnded = "The simulation has ended. Cannot modify simulation after the end of the simulation.";
const char *kRunning = "The simulation is running. Stop the simulation to make changes.";
const char *kNotStarted = "The simulation has not started. Get and set commands cannot be called before sim.start()";
}  // namespace

// ================================================================ (1) bookkeeping ================================

Simulation::Simulation(int device_, int integrator_, Init)
    : device(device_), integrator(integrator_), dt(0.0001), T(0), _global_acc(0, 0, -9.81) {}   // src/sim.cu:76-86

Simulation::Simulation(const std::vector<int> &devices_, int integrator_, Init)
    : device(devices_.empty() ? 0 : devices_[0]), integrator(integrator_), devices(devices_), dt(0.0001), T(0), _global_acc(0, 0, -9.81) {}

Dev Simulation::dev() const { return Dev{engine, group}; }

Mass *Simulation::createMass(Mass *m) {   // src/sim.cu:222-253
    if (ENDED) throw std::runtime_error(kEnded);
    if (STARTED && RUNNING) throw std::runtime_error(kRunning);
    m->ref_count++;
    masses.push_back(m);
    if (STARTED) append_dirty = true;     // appended on the device at the next resume()/get()/set() (sync_topology)
    return m;
}

Mass *Simulation::createMass() {
    if (ENDED) throw std::runtime_error(kEnded);
    return createMass(new Mass());
}

Mass *Simulation::createMass(const Vec &pos) {
    if (ENDED) throw std::runtime_error(kEnded);
    return createMass(new Mass(pos));
}

Spring *Simulation::createSpring(Spring *s) {   // src/sim.cu:290-322 (exit(1) while running -> exception)
    if (ENDED) throw std::runtime_error(kEnded);
    if (STARTED && RUNNING) throw std::runtime_error(kRunning);
    if (s->_right) s->_right->ref_count++;
    if (s->_left) s->_left->ref_count++;
    springs.push_back(s);
    if (STARTED) append_dirty = true;
    return s;
}

Spring *Simulation::createSpring() {
    if (ENDED) throw std::runtime_error(kEnded);
    return createSpring(new Spring());
}

Spring *Simulation::createSpring(Mass *m1, Mass *m2) {
    if (ENDED) throw std::runtime_error(kEnded);
    return createSpring(new Spring(m1, m2));
}

// A deleted mass that springs still point at stays alive but invalid: those springs are skipped by the step
// (src/sim.cu:1163), exactly what the reference's `invalidate` kernel arranges (src/sim.cu:343-380).
void Simulation::deleteMass(Mass *m) {
    if (ENDED) throw std::runtime_error(kEnded);
    if (STARTED && RUNNING) throw std::runtime_error(kRunning);
    masses.erase(std::remove(masses.begin(), masses.end(), m), masses.end());
    m->valid = false;
    if (STARTED && m->slot >= 0 && !topology_dirty) {
        // on the device the mass freezes where it is and its springs exert no force from now on (tb_set_mass_state valid = 0:
        // flag byte + tombstones, no rebuild) -- what the reference's `invalidate` kernel arranges
        if (group) topology_dirty = true;
        else {
            const int32_t slot = m->slot;
            const uint8_t zero = 0;
            engine_check(tb_set_mass_state(engine, 0, 1, &slot, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, &zero), "tb_set_mass_state");
        }
    }
    m->slot = -1;
    if (m->ref_count > 1) zombies.push_back(m);
    m->decrementRefCount();
}

void Simulation::deleteSpring(Spring *s) {   // src/sim.cu:382-412
    if (ENDED) throw std::runtime_error(kEnded);
    if (STARTED && RUNNING) throw std::runtime_error(kRunning);
    springs.erase(std::remove(springs.begin(), springs.end(), s), springs.end());
    for (Mass *end : {s->_left, s->_right}) {
        if (!end) continue;
        if (!end->valid && end->ref_count == 1) zombies.erase(std::remove(zombies.begin(), zombies.end(), end), zombies.end());
        end->decrementRefCount();
    }
    if (STARTED) {
        // once started the spring becomes a tombstone on the device (no force from now on, indices of the others
        // unchanged): no topology rebuild; compaction happens at the next create*() re-upload
        if (s->slot >= 0 && !topology_dirty) {
            int32_t slot = s->slot;
            engine_check(dv_remove_springs(dev(), 1, &slot), "tb_remove_springs");
        }                                     // slot < 0: created while paused and never uploaded -- nothing to undo
        delete s;   // before start() the reference leaves the object to the caller
    }
}

void Simulation::deleteContainer(Container *c) {   // src/sim.cu:478-564
    if (ENDED) throw std::runtime_error(kEnded);
    if (RUNNING) throw std::runtime_error(kRunning);
    // springs first, so that no spring is left pointing at a freed mass
    std::unordered_set<Spring *> dead_springs(c->springs.begin(), c->springs.end());
    std::vector<Spring *> doomed;
    for (Spring *s : springs) if (dead_springs.count(s)) doomed.push_back(s);
    springs.erase(std::remove_if(springs.begin(), springs.end(), [&](Spring *s) { return dead_springs.count(s) != 0; }), springs.end());
    // Once started, on one GPU: the container's springs become tombstones and its masses freeze in place on the device, like
    // deleteSpring / deleteMass one by one (no re-upload; the device slots of the survivors do not move).  Objects that were
    // created while paused and never uploaded (slot < 0), a pending re-upload, or several GPUs: the full re-upload.
    bool in_place = STARTED && !topology_dirty && !group && engine;
    std::vector<int32_t> spring_slots, mass_slots;
    for (Spring *s : doomed) {
        if (s->slot >= 0) spring_slots.push_back(s->slot); else if (STARTED) in_place = false;
        for (Mass *end : {s->_left, s->_right}) {
            if (!end) continue;
            if (!end->valid && end->ref_count == 1) zombies.erase(std::remove(zombies.begin(), zombies.end(), end), zombies.end());
            end->decrementRefCount();
        }
        delete s;
    }
    std::unordered_set<Mass *> dead_masses(c->masses.begin(), c->masses.end());
    std::vector<Mass *> gone;
    for (Mass *m : masses) if (dead_masses.count(m)) gone.push_back(m);
    masses.erase(std::remove_if(masses.begin(), masses.end(), [&](Mass *m) { return dead_masses.count(m) != 0; }), masses.end());
    for (Mass *m : gone) {
        if (m->slot >= 0) mass_slots.push_back(m->slot); else if (STARTED) in_place = false;
        m->valid = false;
        m->slot = -1;
        if (m->ref_count > 1) zombies.push_back(m);
        m->decrementRefCount();
    }
    containers.erase(std::remove(containers.begin(), containers.end(), c), containers.end());
    delete c;
    if (!STARTED) return;
    if (!in_place) { topology_dirty = true; return; }
    if (!spring_slots.empty()) engine_check(tb_remove_springs(engine, (int64_t)spring_slots.size(), spring_slots.data()), "tb_remove_springs");
    if (!mass_slots.empty()) {
        const std::vector<uint8_t> zero(mass_slots.size(), 0);
        engine_check(tb_set_mass_state(engine, 0, (int64_t)mass_slots.size(), mass_slots.data(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                       nullptr, zero.data()), "tb_set_mass_state");
    }
}   // (the whole-array transfers check for themselves that the object lists still line up with the device slots)

Spring *Simulation::getSpringByIndex(int i) { assert(i >= 0 && (size_t)i < springs.size()); return springs[i]; }
Mass *Simulation::getMassByIndex(int i) { assert(i >= 0 && (size_t)i < masses.size()); return masses[i]; }
Container *Simulation::getContainerByIndex(int i) { assert(i >= 0 && (size_t)i < containers.size()); return containers[i]; }

Container *Simulation::createContainer() {
    Container *c = new Container();
    containers.push_back(c);
    return c;
}

void Simulation::adopt(Container *c) {   // the loops shared by createCube/Lattice/Beam/RobotLink, src/sim.cu:2062-2222
    masses.reserve(masses.size() + c->masses.size());
    springs.reserve(springs.size() + c->springs.size());
    for (Mass *m : c->masses) createMass(m);
    for (Spring *s : c->springs) createSpring(s);
    containers.push_back(c);
}

Cube *Simulation::createCube(const Vec &center, double side_length) {
    if (ENDED) throw std::runtime_error("The simulation has ended. Cannot create new objects");
    Cube *c = new Cube(center, side_length);
    adopt(c);
    return c;
}

Lattice *Simulation::createLattice(const Vec &center, const Vec &dims, int nx, int ny, int nz) {
    if (ENDED) throw std::runtime_error("The simulation has ended. New objects cannot be created.");
    Lattice *l = new Lattice(center, dims, nx, ny, nz);
    adopt(l);
    return l;
}

Beam *Simulation::createBeam(const Vec &center, const Vec &dims, int nx, int ny, int nz) {
    if (ENDED) throw std::runtime_error("The simulation has ended. New objects cannot be created.");
    Beam *b = new Beam(center, dims, nx, ny, nz);
    adopt(b);
    return b;
}

RobotLink *Simulation::createRobotLink(const Vec &pos1, const Vec &pos2, double mass, double max_exp_length,
                                       double min_exp_length, double expansion_rate, double k, double magnetic_force,
                                       double radius) {
    if (ENDED) throw std::runtime_error("The simulation has ended. New objects cannot be created.");
    RobotLink *l = new RobotLink(pos1, pos2, mass, max_exp_length, min_exp_length, expansion_rate, k, magnetic_force, radius);
    adopt(l);
    return l;
}

Container *Simulation::importFromSTL(const std::string &, double, int) {
    throw std::runtime_error("importFromSTL is not built in titan-b200 (host preprocessing outside the step path, SURVEY.md 8f)");
}

void Simulation::createPlane(const Vec &abc, double d) { createPlane(abc, d, 0.0, 0.0); }

void Simulation::createPlane(const Vec &abc, double d, double FRICTION_K, double FRICTION_S) {   // src/sim.cu:2251-2276
    if (ENDED) throw std::runtime_error("The simulation has ended. New objects cannot be created.");
    ContactPlane *p = new ContactPlane(abc, d);
    p->_FRICTION_K = FRICTION_K;
    p->_FRICTION_S = FRICTION_S;
    constraints.push_back(p);
    dev_planes.push_back(*p);
    constraints_dirty = true;
}

void Simulation::createBall(const Vec &center, double r) {   // src/sim.cu:2278-2288
    if (ENDED) throw std::runtime_error("The simulation has ended. New constraints cannot be added.");
    Ball *b = new Ball(center, r);
    constraints.push_back(b);
    dev_balls.push_back(*b);
    constraints_dirty = true;
}

// The reference clears only the host list; the device-side planes and balls stay active (src/sim.cu:2290-2293).
// Reproduced as is: dev_planes / dev_balls are untouched.
void Simulation::clearConstraints() {
    for (Constraint *c : constraints) delete c;
    constraints.clear();
    constraints_dirty = true;
}

void Simulation::setAllSpringConstantValues(double k) {
    if (ENDED) throw std::runtime_error(kEnded);
    for (Spring *s : springs) s->_k = k;
}

void Simulation::defaultRestLengths() {
    if (ENDED) throw std::runtime_error(kEnded);
    for (Spring *s : springs) s->defaultLength();
}

void Simulation::setAllMassValues(double m) {
    if (ENDED) throw std::runtime_error(kEnded);
    for (Mass *q : masses) q->m += m;
}

void Simulation::setTimeStep(double delta_t) {   // src/sim.cu:798-808
    if (ENDED) throw std::runtime_error(kEnded);
    if (delta_t <= 0) throw std::runtime_error("Cannot set time step to negative or zero value.");
    dt = delta_t;
    constraints_dirty = true;
}

double Simulation::getTimeStep() { return dt; }

void Simulation::setGlobalAcceleration(const Vec &global_acc) {   // src/sim.cu:2334-2340
    if (RUNNING) throw std::runtime_error("The simulation is running. The global force parameter cannot be changed during runtime");
    _global_acc = global_acc;
    constraints_dirty = true;
}

void Simulation::printPositions() {
    if (ENDED) throw std::runtime_error("The simulation has ended. You cannot view parameters of the simulation after it has been stopped.");
    std::cout << "\nHOST MASSES: " << std::endl;
    int count = 0;
    for (Mass *m : masses) std::cout << count++ << ": " << m->pos << std::endl;
    std::cout << std::endl;
}

// ================================================================ (2) the packer =================================
void Simulation::engine_check(int rc, const char *what) {
    if (rc == TB_OK) return;
    std::string msg = std::string(what) + ": " + (engine ? dv_last_error(dev()) : tb_last_error(nullptr));
    throw std::runtime_error(msg);
}

namespace {

// Modelled bytes one tile of a bx x by x bz brick streams per step: `tile` mass slots, one 16-byte record per spring
// touching the brick (cut springs are stored by both bricks), 24 bytes per halo mass.
double brick_cost(int bx, int by, int bz, int tile) {
    double a = bx - 1, b = by - 1, c = bz - 1;
    double inner = (a * by * bz + bx * b * bz + bx * by * c) + 2 * (a * b * bz + a * by * c + bx * b * c) + 4 * a * b * c;
    double recs = 26.0 * bx * by * bz - inner;
    double halo = (double)(bx + 2) * (by + 2) * (bz + 2) - (double)bx * by * bz;
    return tile * 104.0 + 16.0 * recs + 24.0 * halo;
}
int cdiv(int a, int b) { return (a + b - 1) / b; }

// Brick edge lengths (bx,by,bz), bx*by*bz <= tile, minimising the modelled bytes of a step over an nx x ny x nz lattice.
double brick_dims(int nx, int ny, int nz, int tile, int out[3]) {
    double best = std::numeric_limits<double>::max();
    out[0] = out[1] = out[2] = 1;
    for (int bx = 1; bx <= std::min(nx, tile); bx++)
        for (int by = 1; by <= std::min(ny, tile / bx); by++) {
            int bz = std::min(nz, tile / (bx * by));
            if (bz < 1) continue;
            double cost = (double)cdiv(nx, bx) * cdiv(ny, by) * cdiv(nz, bz) * brick_cost(bx, by, bz, tile);
            if (cost < best) { best = cost; out[0] = bx; out[1] = by; out[2] = bz; }
        }
    return best;
}

// Layers of bricks along i (thickness, by, bz), thickest first: when nx is not a multiple of the best brick's
// thickness one shape wastes its last layer, so the thicknesses are chosen by dynamic programming, each layer with its
// cheapest cross-section (same model; titan_b200/lattice.py::brick_layers is the array-side twin).
struct BrickLayer { int t, by, bz; bool operator==(const BrickLayer &o) const { return t == o.t && by == o.by && bz == o.bz; } };
std::vector<BrickLayer> brick_layers(int nx, int ny, int nz, int tile) {
    int uni[3];
    double uniform_cost = brick_dims(nx, ny, nz, tile, uni);
    std::vector<BrickLayer> uniform((size_t)cdiv(nx, uni[0]), BrickLayer{uni[0], uni[1], uni[2]});
    const int tmax = std::min(nx, tile);
    std::vector<double> layer_cost((size_t)tmax + 1, std::numeric_limits<double>::max());
    std::vector<BrickLayer> layer((size_t)tmax + 1);
    for (int t = 1; t <= tmax; t++)
        for (int by = 1; by <= std::min(ny, tile / t); by++) {
            int bz = std::min(nz, tile / (t * by));
            if (bz < 1) continue;
            double c = (double)cdiv(ny, by) * cdiv(nz, bz) * brick_cost(t, by, bz, tile);
            if (c < layer_cost[(size_t)t]) { layer_cost[(size_t)t] = c; layer[(size_t)t] = BrickLayer{t, by, bz}; }
        }
    std::vector<double> cost((size_t)nx + 1, std::numeric_limits<double>::max());
    std::vector<int> pick((size_t)nx + 1, 0);
    cost[0] = 0;
    for (int n = 1; n <= nx; n++)
        for (int t = 1; t <= std::min(n, tmax); t++)
            if (cost[(size_t)(n - t)] + layer_cost[(size_t)t] < cost[(size_t)n]) { cost[(size_t)n] = cost[(size_t)(n - t)] + layer_cost[(size_t)t]; pick[(size_t)n] = t; }
    if (uniform_cost <= cost[(size_t)nx]) return uniform;
    std::vector<BrickLayer> out;
    for (int n = nx; n > 0; n -= pick[(size_t)n]) out.push_back(layer[(size_t)pick[(size_t)n]]);
    std::stable_sort(out.begin(), out.end(), [](const BrickLayer &a, const BrickLayer &b) { return a.t > b.t; });
    return out;
}

// Tile id of every cell of an nx x ny x nz grid (i outermost), starting at `base`; returns the number of tiles.  Bricks
// are numbered super-block by super-block (8^3 bricks each) rather than plane by plane, so that the persistent CTAs
// work on a compact neighbourhood whose halo state is still in L2.
int32_t brick_ids(int nx, int ny, int nz, int32_t base, std::vector<int32_t> &ids) {
    const int B = 8;
    ids.assign((size_t)nx * ny * nz, 0);
    std::vector<BrickLayer> layers = brick_layers(nx, ny, nz, 128);
    int32_t next = base;
    int i0 = 0;
    for (size_t j = 0; j < layers.size();) {
        size_t run = 1;
        while (j + run < layers.size() && layers[j + run] == layers[j]) run++;
        const BrickLayer L = layers[j];
        const int planes = std::min(nx - i0, L.t * (int)run);
        const int nti = cdiv(planes, L.t), ntj = cdiv(ny, L.by), ntk = cdiv(nz, L.bz);
        std::vector<int32_t> rank((size_t)nti * ntj * ntk);
        if (std::max(nti, std::max(ntj, ntk)) > B) {
            std::vector<int32_t> order(rank.size());
            for (size_t q = 0; q < order.size(); q++) order[q] = (int32_t)q;
            auto key = [&](int32_t q) {
                int bi = q / (ntj * ntk), bj = (q / ntk) % ntj, bk = q % ntk;
                return std::array<int, 6>{bi / B, bj / B, bk / B, bi % B, bj % B, bk % B};
            };
            std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return key(a) < key(b); });
            for (size_t r = 0; r < order.size(); r++) rank[(size_t)order[r]] = (int32_t)r;
        } else {
            for (size_t q = 0; q < rank.size(); q++) rank[q] = (int32_t)q;
        }
        for (int i = 0; i < planes; i++)
            for (int jj = 0; jj < ny; jj++)
                for (int k = 0; k < nz; k++)
                    ids[((size_t)(i0 + i) * ny + jj) * nz + k] = next + rank[((size_t)(i / L.t) * ntj + jj / L.by) * ntk + k / L.bz];
        next += (int32_t)rank.size();
        i0 += planes;
        j += run;
    }
    return next - base;
}

}  // namespace

// test hook (tests/test_facade.py): the tile hint start() hands the engine for an nx x ny x nz lattice
extern "C" __attribute__((visibility("default"))) int titan_b200_brick_hint(int nx, int ny, int nz, int32_t *out) {
    std::vector<int32_t> ids;
    int32_t n = brick_ids(nx, ny, nz, 0, ids);
    std::copy(ids.begin(), ids.end(), out);
    return n;
}

namespace {

tb_plane to_plane(const ContactPlane &p) {
    tb_plane r;
    for (int c = 0; c < 3; c++) r.normal[c] = p._normal[c];
    r.offset = p._offset; r.friction_k = p._FRICTION_K; r.friction_s = p._FRICTION_S;
    return r;
}

struct MassPack {   // contiguous arrays for one tb_set_mass_state / tb_set_mass_magnet call
    std::vector<int32_t> idx, drag;
    std::vector<double> pos, vel, m, T, ext, rad, stiff, maxmag, magscale;
    std::vector<uint8_t> fixed, valid;
    void reserve(size_t n) {
        idx.reserve(n); drag.reserve(n); pos.reserve(3 * n); vel.reserve(3 * n); m.reserve(n); T.reserve(n); ext.reserve(3 * n);
        rad.reserve(n); stiff.reserve(n); maxmag.reserve(n); magscale.reserve(n); fixed.reserve(n); valid.reserve(n);
    }
};

}  // namespace

// ---- whole-simulation transfers (setAll / getAll / start) --------------------------------------------------------
// The reference's setAll()/getAll() rebuild a packed array of 240-byte CUDA_MASS / 80-byte CUDA_SPRING structs on one host
// thread and copy it through pageable memory (src/sim.cu:720-765,1094-1116).  Here the host objects are packed by all
// host cores straight into page-locked staging arrays (pos, vel, T every time: the dynamic state), and the parameter
// arrays (m, fixed, drag, magnet fields; k, rest, damping, type, omega, l_max, l_min, rate) are compared with a shadow of
// what the device already holds and uploaded only when they differ -- same device contents afterwards, a fraction of the
// bytes and none of the tile-structure rebuilds a spring-type upload would otherwise trigger.
template <class F> static void parallel_for(size_t n, F f) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = n < 32768 ? 1u : std::min(32u, hw);
    if (nt == 1) { f((size_t)0, n); return; }
    std::vector<std::thread> th;
    const size_t per = (n + nt - 1) / nt;
    for (unsigned t = 0; t < nt; t++) {
        const size_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
        if (lo < hi) th.emplace_back([=, &f] { f(lo, hi); });
    }
    for (std::thread &t : th) t.join();
}

// TITAN_B200_TIMING=1: wall-clock phases of the whole-simulation transfers on stderr
struct PhaseTimer {
    const char *what; bool on; std::chrono::steady_clock::time_point t0;
    explicit PhaseTimer(const char *w) : what(w), on(getenv("TITAN_B200_TIMING") != nullptr), t0(std::chrono::steady_clock::now()) {}
    void lap(const char *phase) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        std::cerr << "[titan_b200] " << what << ": " << phase << " " << std::chrono::duration<double, std::milli>(t1 - t0).count() << " ms" << std::endl;
        t0 = t1;
    }
};

struct Simulation::HostStage {
    size_t cap = 0;
    double *pos = nullptr, *vel = nullptr, *acc = nullptr, *T = nullptr, *ext = nullptr;   // pinned (owned by the engine)
    bool mass_valid = false, spring_valid = false, ext_uploaded_nonzero = false;
    std::vector<double> m, rad, stiff, maxmag, magscale, k, rest, damping, omega, lmax, lmin, rate;
    std::vector<uint8_t> fixed;
    std::vector<int32_t> drag, type;
};

bool Simulation::upload_masses_whole() {
    const size_t n = masses.size();
    if (n == 0 || !engine) return false;
    if (!hs) hs = new HostStage();
    PhaseTimer pt("setAll masses");
    if (hs->cap < n) {
        const size_t cap = n + n / 4 + 1024;
        void *p = nullptr;
        engine_check(dv_alloc_host(dev(), cap * 13 * sizeof(double), &p), "tb_alloc_host");
        double *d = (double *)p;
        hs->pos = d; hs->vel = d + 3 * cap; hs->acc = d + 6 * cap; hs->ext = d + 9 * cap; hs->T = d + 12 * cap;
        hs->cap = cap;
    }
    const bool first = !hs->mass_valid || hs->m.size() != n;
    if (first) {
        hs->m.resize(n); hs->rad.resize(n); hs->stiff.resize(n); hs->maxmag.resize(n); hs->magscale.resize(n);
        hs->fixed.resize(n); hs->drag.resize(n);
    }
    std::atomic<bool> in_order{true}, ext_nonzero{false}, lc_changed{false};
    std::atomic<unsigned> changed{0};
    auto differs = [](double a, double b) { return std::memcmp(&a, &b, 8) != 0; };
    parallel_for(n, [&](size_t lo, size_t hi) {
        bool nz = false, lc = false;
        unsigned ch = 0;
        for (size_t j = lo; j < hi; j++) {
            Mass *q = masses[j];
            if (q->slot != (int32_t)j) { in_order = false; return; }
            for (int c = 0; c < 3; c++) {
                hs->pos[3 * j + c] = q->pos[c]; hs->vel[3 * j + c] = q->vel[c]; hs->ext[3 * j + c] = q->extern_force[c];
                nz |= q->extern_force[c] != 0.0;
            }
            hs->T[j] = q->T;
            if (differs(hs->m[j], q->m)) { hs->m[j] = q->m; ch |= 1u; }
            const uint8_t fx = q->fixed ? 1 : 0;
            if (hs->fixed[j] != fx) { hs->fixed[j] = fx; ch |= 2u; }
            const int32_t dr = (int32_t)q->drag_coefficient;   // int truncation, include/Titan/object.h:213
            if (hs->drag[j] != dr) { hs->drag[j] = dr; ch |= 4u; }
            if (differs(hs->rad[j], q->rad)) { hs->rad[j] = q->rad; ch |= 8u; }
            if (differs(hs->stiff[j], q->stiffness)) { hs->stiff[j] = q->stiffness; ch |= 8u; }
            if (differs(hs->maxmag[j], q->max_mag_force)) { hs->maxmag[j] = q->max_mag_force; ch |= 8u; }
            if (differs(hs->magscale[j], q->mag_scale_factor)) { hs->magscale[j] = q->mag_scale_factor; ch |= 8u; }
            std::vector<LocalConstraintRecord> &held = dev_local[j];
            bool same = held.size() == q->local.size();
            for (size_t r = 0; same && r < held.size(); r++)
                same = held[r].kind == q->local[r].kind && held[r].vec == q->local[r].vec && held[r].num == q->local[r].num;
            if (!same) { held = q->local; lc = true; }
        }
        if (nz) ext_nonzero = true;
        if (lc) lc_changed = true;
        if (ch) changed |= ch;
    });
    if (!in_order) { hs->mass_valid = false; return false; }
    pt.lap("pack");
    const unsigned ch = first ? 0xfu : changed.load();
    // the device consumes extern_force (one shot, src/sim.cu:1365) but fixed masses keep theirs: upload whenever this or the
    // previous upload carried a non-zero force
    const bool up_ext = ext_nonzero || hs->ext_uploaded_nonzero || first;
    std::vector<uint8_t> valid;
    if (first) valid.assign(n, 1);
    engine_check(dv_set_mass_state(dev(), 0, (int64_t)n, nullptr, hs->pos, hs->vel, (ch & 1u) ? hs->m.data() : nullptr, hs->T,
                                   up_ext ? hs->ext : nullptr, (ch & 2u) ? hs->fixed.data() : nullptr, (ch & 4u) ? hs->drag.data() : nullptr,
                                   first ? valid.data() : nullptr), "tb_set_mass_state");
    hs->ext_uploaded_nonzero = ext_nonzero;
    if (ch & 8u)
        engine_check(dv_set_mass_magnet(dev(), 0, (int64_t)n, nullptr, hs->rad.data(), hs->stiff.data(), hs->maxmag.data(), hs->magscale.data()), "tb_set_mass_magnet");
    hs->mass_valid = true;
    if (lc_changed) local_dirty = true;
    if (local_dirty) upload_local_constraints();
    pt.lap("upload");
    return true;
}

// CUDA_MASS(Mass&), src/mass.cu:74-100: m, T, pos, vel, extern_force, magnet fields, fixed, drag, valid = true.
void Simulation::upload_masses(Mass *const *list, size_t count) {
    if (count == 0) return;
    if (list == masses.data() && count == masses.size() && upload_masses_whole()) return;
    if (hs) hs->mass_valid = false;        // a partial upload: the shadow no longer mirrors the device
    MassPack P;
    P.reserve(count);
    bool lc_changed = false;
    for (size_t j = 0; j < count; j++) {
        Mass *q = list[j];
        if (q->slot < 0) continue;   // not on the device (deleted, or foreign)
        P.idx.push_back(q->slot);
        for (int c = 0; c < 3; c++) { P.pos.push_back(q->pos[c]); P.vel.push_back(q->vel[c]); P.ext.push_back(q->extern_force[c]); }
        P.m.push_back(q->m); P.T.push_back(q->T);
        P.fixed.push_back(q->fixed ? 1 : 0); P.valid.push_back(1);
        P.drag.push_back((int32_t)q->drag_coefficient);   // int truncation, include/Titan/object.h:213
        P.rad.push_back(q->rad); P.stiff.push_back(q->stiffness); P.maxmag.push_back(q->max_mag_force); P.magscale.push_back(q->mag_scale_factor);
        std::vector<LocalConstraintRecord> &held = dev_local[(size_t)q->slot];
        bool same = held.size() == q->local.size();
        for (size_t r = 0; same && r < held.size(); r++)
            same = held[r].kind == q->local[r].kind && held[r].vec == q->local[r].vec && held[r].num == q->local[r].num;
        if (!same) { held = q->local; lc_changed = true; }
    }
    const int64_t n = (int64_t)P.idx.size();
    if (n == 0) return;
    const bool whole = n == (int64_t)masses.size() && P.idx.front() == 0 && P.idx.back() == n - 1 && std::is_sorted(P.idx.begin(), P.idx.end());
    const int32_t *idx = whole ? nullptr : P.idx.data();
    engine_check(dv_set_mass_state(dev(), 0, n, idx, P.pos.data(), P.vel.data(), P.m.data(), P.T.data(), P.ext.data(),
                                   P.fixed.data(), P.drag.data(), P.valid.data()), "tb_set_mass_state");
    engine_check(dv_set_mass_magnet(dev(), 0, n, idx, P.rad.data(), P.stiff.data(), P.maxmag.data(), P.magscale.data()), "tb_set_mass_magnet");
    if (lc_changed) local_dirty = true;
    if (local_dirty) upload_local_constraints();
}

bool Simulation::upload_springs_whole() {
    const size_t n = springs.size();
    if (n == 0 || !engine) return false;
    if (!hs) hs = new HostStage();
    PhaseTimer pt("setAll springs");
    const bool first = !hs->spring_valid || hs->k.size() != n;
    if (first) {
        hs->k.resize(n); hs->rest.resize(n); hs->damping.resize(n); hs->omega.resize(n); hs->lmax.resize(n); hs->lmin.resize(n);
        hs->rate.resize(n); hs->type.resize(n);
    }
    // one pass over the host objects by all cores: every field is compared with the shadow IN PLACE (bit pattern) and the
    // shadow updated, so afterwards the shadow holds the new values and the flags say which arrays the device lacks
    std::atomic<bool> in_order{true}, actuated{false};
    std::atomic<unsigned> changed{0};
    auto differs = [](double a, double b) { return std::memcmp(&a, &b, 8) != 0; };
    parallel_for(n, [&](size_t lo, size_t hi) {
        bool act = false;
        unsigned ch = 0;
        for (size_t j = lo; j < hi; j++) {
            const Spring *sp = springs[j];
            if (sp->slot != (int32_t)j) { in_order = false; return; }
            if (differs(hs->k[j], sp->_k)) { hs->k[j] = sp->_k; ch |= 1u; }
            if (differs(hs->rest[j], sp->_rest)) { hs->rest[j] = sp->_rest; ch |= 2u; }
            if (differs(hs->damping[j], sp->_damping)) { hs->damping[j] = sp->_damping; ch |= 4u; }
            if (differs(hs->omega[j], sp->_omega)) { hs->omega[j] = sp->_omega; ch |= 8u; }
            if (differs(hs->lmax[j], sp->_l_max)) { hs->lmax[j] = sp->_l_max; ch |= 16u; }
            if (differs(hs->lmin[j], sp->_l_min)) { hs->lmin[j] = sp->_l_min; ch |= 32u; }
            if (differs(hs->rate[j], sp->_rate)) { hs->rate[j] = sp->_rate; ch |= 64u; }
            if (hs->type[j] != (int32_t)sp->_type) { hs->type[j] = (int32_t)sp->_type; ch |= 128u; }
            act |= sp->_type == ACTUATED_EXPAND || sp->_type == ACTUATED_CONTRACT;
        }
        if (act) actuated = true;
        if (ch) changed |= ch;
    });
    if (!in_order) { hs->spring_valid = false; return false; }
    pt.lap("scan");
    unsigned ch = first ? 0xffu : changed.load();
    if (actuated) ch |= 2u;                  // the device advances _rest of ACTUATED_* springs: the host value always goes down
    if (ch)
        engine_check(dv_set_spring_params(dev(), 0, (int64_t)n, nullptr, (ch & 1u) ? hs->k.data() : nullptr, (ch & 2u) ? hs->rest.data() : nullptr,
                                          (ch & 4u) ? hs->damping.data() : nullptr, (ch & 128u) ? hs->type.data() : nullptr,
                                          (ch & 8u) ? hs->omega.data() : nullptr, (ch & 16u) ? hs->lmax.data() : nullptr,
                                          (ch & 32u) ? hs->lmin.data() : nullptr, (ch & 64u) ? hs->rate.data() : nullptr), "tb_set_spring_params");
    hs->spring_valid = true;
    pt.lap("upload");
    return true;
}

bool Simulation::download_masses_whole() {
    const size_t n = masses.size();
    if (n == 0 || !engine || !hs || hs->cap < n) return false;
    std::atomic<bool> in_order{true};
    parallel_for(n, [&](size_t lo, size_t hi) {
        for (size_t j = lo; j < hi; j++) if (masses[j]->slot != (int32_t)j) { in_order = false; return; }
    });
    if (!in_order) return false;
    PhaseTimer pt("getAll");
    engine_check(dv_get_mass_state(dev(), 0, (int64_t)n, nullptr, hs->pos, hs->vel, hs->acc, hs->T, hs->ext), "tb_get_mass_state");
    pt.lap("download");
    parallel_for(n, [&](size_t lo, size_t hi) {
        for (size_t j = lo; j < hi; j++) {
            Mass *q = masses[j];
            q->pos = Vec(hs->pos[3 * j], hs->pos[3 * j + 1], hs->pos[3 * j + 2]);
            q->vel = Vec(hs->vel[3 * j], hs->vel[3 * j + 1], hs->vel[3 * j + 2]);
            q->acc = Vec(hs->acc[3 * j], hs->acc[3 * j + 1], hs->acc[3 * j + 2]);
            q->extern_force = Vec(hs->ext[3 * j], hs->ext[3 * j + 1], hs->ext[3 * j + 2]);
            q->T = hs->T[j];
        }
    });
    pt.lap("unpack");
    return true;
}

// CUDA_SPRING(const Spring&), src/spring.cu:37-63.  Endpoints are topology and only change through sync_topology().
void Simulation::upload_springs(Spring *const *list, size_t count) {
    if (count == 0) return;
    if (list == springs.data() && count == springs.size() && upload_springs_whole()) return;
    if (hs) hs->spring_valid = false;
    std::vector<int32_t> idx, type;
    std::vector<double> k, rest, damping, omega, lmax, lmin, rate;
    for (size_t j = 0; j < count; j++) {
        const Spring *s = list[j];
        if (s->slot < 0) continue;
        idx.push_back(s->slot); type.push_back((int32_t)s->_type);
        k.push_back(s->_k); rest.push_back(s->_rest); damping.push_back(s->_damping); omega.push_back(s->_omega);
        lmax.push_back(s->_l_max); lmin.push_back(s->_l_min); rate.push_back(s->_rate);
    }
    const int64_t n = (int64_t)idx.size();
    if (n == 0) return;
    const bool whole = n == (int64_t)springs.size() && idx.front() == 0 && idx.back() == n - 1 && std::is_sorted(idx.begin(), idx.end());
    engine_check(dv_set_spring_params(dev(), 0, n, whole ? nullptr : idx.data(), k.data(), rest.data(), damping.data(), type.data(),
                                      omega.data(), lmax.data(), lmin.data(), rate.data()), "tb_set_spring_params");
}

// Mass::operator=(CUDA_MASS&), src/mass.cu:21-49: the device changes pos, vel, acc, T and consumes extern_force.
void Simulation::download_masses(Mass *const *list, size_t count) {
    if (list == masses.data() && count == masses.size() && download_masses_whole()) return;
    std::vector<int32_t> idx;
    std::vector<Mass *> who;
    for (size_t j = 0; j < count; j++) if (list[j]->slot >= 0) { idx.push_back(list[j]->slot); who.push_back(list[j]); }
    const int64_t n = (int64_t)idx.size();
    if (n == 0) return;
    std::vector<double> pos(3 * n), vel(3 * n), acc(3 * n), tt(n), ext(3 * n);
    const bool whole = n == (int64_t)masses.size() && idx.front() == 0 && idx.back() == n - 1 && std::is_sorted(idx.begin(), idx.end());
    engine_check(dv_get_mass_state(dev(), 0, n, whole ? nullptr : idx.data(), pos.data(), vel.data(), acc.data(), tt.data(), ext.data()),
                 "tb_get_mass_state");
    for (int64_t j = 0; j < n; j++) {
        Mass *q = who[(size_t)j];
        q->pos = Vec(pos[3 * j], pos[3 * j + 1], pos[3 * j + 2]);
        q->vel = Vec(vel[3 * j], vel[3 * j + 1], vel[3 * j + 2]);
        q->acc = Vec(acc[3 * j], acc[3 * j + 1], acc[3 * j + 2]);
        q->extern_force = Vec(ext[3 * j], ext[3 * j + 1], ext[3 * j + 2]);
        q->T = tt[j];
    }
}

void Simulation::upload_constraints() {
    std::vector<tb_plane> planes;
    for (const ContactPlane &p : dev_planes) planes.push_back(to_plane(p));
    std::vector<tb_ball> balls;
    for (const Ball &b : dev_balls) {
        tb_ball r;
        for (int c = 0; c < 3; c++) r.center[c] = b._center[c];
        r.radius = b._radius;
        balls.push_back(r);
    }
    engine_check(dv_set_planes(dev(), (int32_t)planes.size(), planes.data()), "tb_set_planes");
    engine_check(dv_set_balls(dev(), (int32_t)balls.size(), balls.data()), "tb_set_balls");
    const double g[3] = {_global_acc[0], _global_acc[1], _global_acc[2]};
    engine_check(dv_set_globals(dev(), g, dt), "tb_set_globals");
    dt_engine = dt;        // what execute() accumulates T with: setTimeStep() while running takes effect at the next resume()
    constraints_dirty = false;
}

// The per-mass constraint lists the device holds (dev_local) as one CSR table.
void Simulation::upload_local_constraints() {
    const size_t nm = masses.size();
    std::vector<int64_t> ptr(nm + 1, 0);
    std::vector<tb_local> rec;
    for (size_t i = 0; i < nm; i++) {
        for (const LocalConstraintRecord &r : dev_local[i]) {
            tb_local t;
            t.kind = r.kind; t.pad = 0; t.num = r.num;
            for (int c = 0; c < 3; c++) t.vec[c] = r.vec[c];
            rec.push_back(t);
        }
        ptr[i + 1] = (int64_t)rec.size();
    }
    if (group) {
        if (!rec.empty()) throw std::runtime_error("per-mass constraints (Mass::addConstraint) are not available on a multi-GPU Simulation");
        local_dirty = false;
        return;
    }
    engine_check(tb_set_local_constraints(engine, rec.empty() ? nullptr : ptr.data(), rec.data()), "tb_set_local_constraints");
    local_dirty = false;
}

// Topology + every field, from the host objects (toArray, src/sim.cu:940-1041).
void Simulation::upload_everything() {
    const int64_t nm = (int64_t)masses.size(), ns = (int64_t)springs.size();
    PhaseTimer pt_topo("start");
    parallel_for((size_t)nm, [&](size_t lo, size_t hi) { for (size_t i = lo; i < hi; i++) masses[i]->slot = (int32_t)i; });
    std::vector<int32_t> left((size_t)ns), right((size_t)ns);
    parallel_for((size_t)ns, [&](size_t lo, size_t hi) {          // pointer chasing over every Spring and its two Masses: all cores
        for (size_t s = lo; s < hi; s++) {
            Spring *sp = springs[s];
            sp->slot = (int32_t)s;
            // nullptr or deleted endpoint -> -1: the spring is skipped by the step (src/sim.cu:1163)
            left[s] = (sp->_left && sp->_left->valid) ? sp->_left->slot : -1;
            right[s] = (sp->_right && sp->_right->valid) ? sp->_right->slot : -1;
        }
    });
    pt_topo.lap("topology from the host objects");
    // ---- several GPUs: who owns which mass -----------------------------------------------------------------------
    // (a) one grid container holding every mass (createLattice): i-slabs, whole planes per GPU, each slab with its own bricks;
    // (b) bodies no spring connects (every mass in exactly one container): whole containers to the GPUs in turn, balanced by
    //     mass count -- nothing is exchanged; (c) anything else: equal contiguous index ranges, tiles by coordinate bisection.
    std::vector<int32_t> owner;
    std::vector<int32_t> hint((size_t)nm, -1);
    bool covered = true;
    const size_t W = group ? devices.size() : 1;
    const Container *grid = nullptr;
    if (containers.size() == 1 && containers[0]->grid_nx && containers[0]->masses.size() == (size_t)nm &&
        (size_t)containers[0]->grid_nx * containers[0]->grid_ny * containers[0]->grid_nz == (size_t)nm)
        grid = containers[0];
    if (grid) for (int64_t i = 0; i < nm && grid; i++) if (grid->masses[(size_t)i] != masses[(size_t)i]) grid = nullptr;   // same order
    if (group && grid && (size_t)grid->grid_nx >= W) {
        const int nx = grid->grid_nx, plane = grid->grid_ny * grid->grid_nz;
        owner.resize((size_t)nm);
        int32_t base = 0;
        for (size_t r = 0; r < W; r++) {
            const int lo = (int)(r * (size_t)(nx / (int)W) + std::min<size_t>(r, (size_t)(nx % (int)W)));
            const int hi = lo + nx / (int)W + (r < (size_t)(nx % (int)W) ? 1 : 0);
            std::vector<int32_t> ids;
            const int32_t n_tiles = brick_ids(hi - lo, grid->grid_ny, grid->grid_nz, base, ids);
            for (int64_t i = (int64_t)lo * plane; i < (int64_t)hi * plane; i++) { owner[(size_t)i] = (int32_t)r; hint[(size_t)i] = ids[(size_t)(i - (int64_t)lo * plane)]; }
            base += n_tiles;
        }
    } else {
        // tile hint: bricks of lattice cells when every mass belongs to exactly one regular grid container
        std::vector<int32_t> cid((size_t)nm, -1);
        bool one_container_each = true;
        int32_t base = 0, c_index = 0;
        for (Container *c : containers) {
            for (const Mass *m : c->masses) {
                if (m->slot < 0 || masses[(size_t)m->slot] != m) continue;
                if (cid[(size_t)m->slot] >= 0) one_container_each = false;
                cid[(size_t)m->slot] = c_index;
            }
            c_index++;
            if (!c->grid_nx || c->masses.size() != (size_t)c->grid_nx * c->grid_ny * c->grid_nz) continue;
            std::vector<int32_t> ids;
            const int32_t n_tiles = brick_ids(c->grid_nx, c->grid_ny, c->grid_nz, base, ids);
            for (size_t q = 0; q < ids.size(); q++) {
                const Mass *m = c->masses[q];
                if (m->slot < 0 || masses[(size_t)m->slot] != m) { covered = false; continue; }
                if (hint[(size_t)m->slot] >= 0) covered = false;   // shared between two grids
                hint[(size_t)m->slot] = ids[q];
            }
            base += n_tiles;
        }
        for (int64_t i = 0; covered && i < nm; i++) covered = hint[(size_t)i] >= 0;
        if (group) {
            for (int64_t i = 0; one_container_each && i < nm; i++) one_container_each = cid[(size_t)i] >= 0;
            for (int64_t sidx = 0; one_container_each && sidx < ns; sidx++)
                if (left[(size_t)sidx] >= 0 && right[(size_t)sidx] >= 0 && cid[(size_t)left[(size_t)sidx]] != cid[(size_t)right[(size_t)sidx]]) one_container_each = false;
            if (one_container_each && containers.size() >= W) {
                std::vector<int64_t> count(containers.size(), 0);
                for (int64_t i = 0; i < nm; i++) count[(size_t)cid[(size_t)i]]++;
                std::vector<int32_t> where(containers.size(), 0);
                int64_t at = 0;
                size_t r = 0;
                for (size_t c = 0; c < containers.size(); c++) {
                    while (r + 1 < W && at * (int64_t)W >= (int64_t)(r + 1) * nm) r++;
                    where[c] = (int32_t)r;
                    at += count[c];
                }
                owner.resize((size_t)nm);
                for (int64_t i = 0; i < nm; i++) owner[(size_t)i] = where[(size_t)cid[(size_t)i]];
            }
        }
    }
    pt_topo.lap("owners and tile hints");
    engine_check(dv_set_topology(dev(), nm, ns, left.data(), right.data(), owner.empty() ? nullptr : owner.data()), "tb_set_topology");
    pt_topo.lap("tb_set_topology");
    dev_nm = nm; dev_ns = ns;
    if (hs) { hs->mass_valid = hs->spring_valid = hs->ext_uploaded_nonzero = false; }   // the device is back at its defaults
    if (covered && nm > 0) engine_check(dv_set_tile_hint(dev(), hint.data()), "tb_set_tile_hint");

    dev_local.assign((size_t)nm, {});
    local_dirty = false;
    upload_masses(masses.data(), masses.size());
    upload_springs(springs.data(), springs.size());
    if (local_dirty) upload_local_constraints();
    upload_constraints();
    topology_dirty = false;
    append_dirty = false;
}

// createMass/createSpring/delete* while paused (src/sim.cu:231-253,300-322,343-564): the engine's arrays are
// rebuilt for the new object lists.  Surviving objects keep the DEVICE's dynamic state (pos, vel, T, pending
// extern_force, _rest); parameters (m, fixed, k, ...) are taken from the host objects.
// Masses and springs created while paused sit at the end of the lists with slot < 0: they are appended on the device
// (tb_add_masses / tb_add_springs: only the tiles they touch are rebuilt).  false = this edit needs the full re-upload.
bool Simulation::append_topology() {
    if (group) return false;
    std::vector<Mass *> new_m;
    for (Mass *q : masses) if (q->slot < 0) new_m.push_back(q);
    std::vector<Spring *> new_s;
    for (Spring *sp : springs) if (sp->slot < 0) new_s.push_back(sp);
    if (!new_m.empty()) {
        int rc = tb_add_masses(engine, (int64_t)new_m.size());
        if (rc == TB_ERR_UNSUPPORTED) return false;
        engine_check(rc, "tb_add_masses");
        for (Mass *q : new_m) q->slot = (int32_t)dev_nm++;
        dev_local.resize((size_t)dev_nm);
        if (hs) hs->mass_valid = false;
        upload_masses(new_m.data(), new_m.size());
    }
    if (!new_s.empty()) {
        std::vector<int32_t> left, right;
        for (Spring *sp : new_s) {
            left.push_back((sp->_left && sp->_left->valid) ? sp->_left->slot : -1);
            right.push_back((sp->_right && sp->_right->valid) ? sp->_right->slot : -1);
        }
        int rc = tb_add_springs(engine, (int64_t)new_s.size(), left.data(), right.data());
        if (rc == TB_ERR_UNSUPPORTED) {
            if (!new_m.empty()) throw std::runtime_error(std::string("tb_add_springs: ") + tb_last_error(engine));
            return false;
        }
        engine_check(rc, "tb_add_springs");
        for (Spring *sp : new_s) sp->slot = (int32_t)dev_ns++;
        if (hs) hs->spring_valid = false;
        upload_springs(new_s.data(), new_s.size());
    }
    return true;
}

void Simulation::sync_topology() {
    if (!topology_dirty && append_dirty) {
        if (append_topology()) { append_dirty = false; return; }
        topology_dirty = true;
    }
    if (!topology_dirty) return;
    std::vector<Mass *> old_m;
    for (Mass *q : masses) if (q->slot >= 0) old_m.push_back(q);
    std::vector<Spring *> old_s;
    for (Spring *s : springs) if (s->slot >= 0) old_s.push_back(s);
    // remember what the host objects hold, fetch the device state into them, upload, then restore the host view
    struct Saved { Vec pos, vel, acc, ext; double T; };
    std::vector<Saved> host_m(old_m.size());
    for (size_t j = 0; j < old_m.size(); j++) host_m[j] = {old_m[j]->pos, old_m[j]->vel, old_m[j]->acc, old_m[j]->extern_force, old_m[j]->T};
    download_masses(old_m.data(), old_m.size());
    std::vector<double> host_rest(old_s.size());
    if (!old_s.empty()) {
        std::vector<int32_t> idx;
        std::vector<double> rest(old_s.size());
        for (Spring *s : old_s) idx.push_back(s->slot);
        engine_check(dv_get_spring_rest(dev(), 0, (int64_t)idx.size(), idx.data(), rest.data()), "tb_get_spring_rest");
        for (size_t j = 0; j < old_s.size(); j++) { host_rest[j] = old_s[j]->_rest; old_s[j]->_rest = rest[j]; }
    }
    upload_everything();
    for (size_t j = 0; j < old_m.size(); j++) {
        Mass *q = old_m[j];
        q->pos = host_m[j].pos; q->vel = host_m[j].vel; q->acc = host_m[j].acc; q->extern_force = host_m[j].ext; q->T = host_m[j].T;
    }
    for (size_t j = 0; j < old_s.size(); j++) old_s[j]->_rest = host_rest[j];
}

void Simulation::get(Mass *m) {   // src/sim.cu:589-602
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    if (ENDED) throw std::runtime_error(kEnded);
    sync_topology();
    download_masses(&m, 1);
}

void Simulation::set(Mass *m) {   // src/sim.cu:604-616
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    if (ENDED) throw std::runtime_error(kEnded);
    sync_topology();
    upload_masses(&m, 1);
}

void Simulation::get(Spring *s) {   // src/sim.cu:618-631; Spring::update copies _rest only (src/spring.cu:12-14)
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    if (ENDED) throw std::runtime_error(kEnded);
    sync_topology();
    if (s->slot < 0) return;
    int32_t idx = s->slot;
    engine_check(dv_get_spring_rest(dev(), 0, 1, &idx, &s->_rest), "tb_get_spring_rest");
}

void Simulation::set(Spring *s) {   // src/sim.cu:633-641
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    sync_topology();
    upload_springs(&s, 1);
}

void Simulation::getAll() {   // src/sim.cu:643-654: masses only
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    if (ENDED) throw std::runtime_error(kEnded);
    sync_topology();
    download_masses(masses.data(), masses.size());
}

void Simulation::get(Container *c) {   // src/sim.cu:1051-1092 (without its whole-simulation over-read)
    if (ENDED) throw std::runtime_error("The simulation has ended. Cannot get updates from the GPU after the end of the simulation.");
    if (!STARTED) { std::cerr << "sim.get() does nothing if called before the simulation has started." << std::endl; return; }
    sync_topology();
    download_masses(c->masses.data(), c->masses.size());
}

void Simulation::set(Container *c) {   // src/sim.cu:656-718 (the container's own springs, not the simulation's)
    if (ENDED) throw std::runtime_error(kEnded);
    if (!STARTED) { std::cerr << kNotStarted << std::endl; return; }
    sync_topology();
    upload_masses(c->masses.data(), c->masses.size());
    upload_springs(c->springs.data(), c->springs.size());
}

void Simulation::setAll() {   // src/sim.cu:720-765
    if (ENDED) throw std::runtime_error(kEnded);
    if (!STARTED) return;
    sync_topology();
    upload_masses(masses.data(), masses.size());
    upload_springs(springs.data(), springs.size());
}

// ================================================================ (3) lifecycle ==================================
void Simulation::throw_if_gpu_error() {
    std::string msg;
    {
        std::lock_guard<std::mutex> lk(mtx);
        msg.swap(gpu_error);
    }
    if (!msg.empty()) throw std::runtime_error(msg);
}

void Simulation::start() {   // src/sim.cu:1547-1591
    if (ENDED) throw std::runtime_error("The simulation has ended. Cannot call sim.start() after the end of the simulation.");
    if (masses.size() == 0) throw std::runtime_error("No masses have been added. Please add masses before starting the simulation.");
    if (STARTED) throw std::runtime_error("The simulation has already been started.");
    if (dt <= 0) throw std::runtime_error("Simultation timestep Simulation::dt is invalid. Please choose a positive non-zero value.");
    std::cout << "Starting simulation with " << masses.size() << " masses and " << springs.size() << " springs." << std::endl;

    // no CUDA device -> TB_ERR_NO_DEVICE -> exception: there is no CPU path
    const uint32_t eflags = TB_MATH_FAST | (integrator == 1 ? TB_INTEGRATOR_VERLET : 0u) | (integrator == 2 ? TB_INTEGRATOR_RK2 : 0u);
    if (devices.size() > 1) {
        std::vector<int32_t> dv(devices.begin(), devices.end());
        engine_check(tb_create_partitioned((int32_t)dv.size(), dv.data(), eflags, &group), "tb_create_partitioned");
        engine = tb_group_engine(group, 0);      // non-null marks "on the device"; every call goes through dev()
    } else {
        engine_check(tb_create(device, eflags, &engine), "tb_create");
    }
    try {
        upload_everything();
        engine_check(dv_run_async(dev(), 0, 0.0), "tb_run_async");   // validates the scene and brings the device layout up to date
    } catch (...) {
        if (group) tb_group_destroy(group); else tb_destroy(engine);
        engine = nullptr; group = nullptr;
        throw;
    }
    T = 0;
    RUNNING = true;
    STARTED = true;
    gpu_thread = std::thread(&Simulation::_run, this);
}

void Simulation::_run() {
    execute();
    GPU_DONE = true;
    cv.notify_all();
}

// The run loop of src/sim.cu:1704-1841 (Euler branch :1800-1811), batched and pipelined: instead of one launch group
// and four device synchronisations per step, whole runs of steps are enqueued (CUDA graphs inside the engine), and the
// next batch is enqueued BEFORE the loop waits for the previous one, so the device never drains between batches (a
// small system on the resident kernel finishes 500 steps in under a millisecond).  How many steps fit before the
// next breakpoint is decided by replaying the reference's own test `*bpts.begin() <= T` over the same fp64
// accumulation `T += dt`.  The breakpoint set is re-read before every batch, and batches are sized to about a
// millisecond of device time, so a breakpoint set while running (pause(time()), stop()) is honoured within two batches.
void Simulation::execute() {
    int64_t chunk = 50;                       // steps per enqueue, adapted to ~1 ms of device time (50 .. 1000)
    double T_enq = T;                         // simulation time after every step enqueued so far
    int64_t enq = 0;                          // steps enqueued by this loop since the last drain
    int64_t base = 0;                         // engine step count at the last drain
    struct Batch { int64_t upto; double T_after; std::chrono::steady_clock::time_point t_enq; int64_t n; };
    std::deque<Batch> inflight;
    dv_steps_done(dev(), &base);
    auto fail_and_park = [&](const char *what) {
        std::unique_lock<std::mutex> lk(mtx);
        gpu_error = std::string(what) + dv_last_error(dev());
        RUNNING = false;
        cv.notify_all();
        cv.wait(lk, [&] { return ENDED.load(); });
    };
    while (true) {
        double next_bp = std::numeric_limits<double>::infinity();
        {
            std::unique_lock<std::mutex> lk(mtx);
            if (!bpts.empty() && *bpts.begin() <= T_enq) {
                int rc = dv_sync(dev());   // cudaDeviceSynchronize at the breakpoint, :1708
                if (rc != TB_OK) gpu_error = std::string("tb_sync: ") + dv_last_error(dev());
                T = T_enq;
                inflight.clear();
                bpts.erase(bpts.begin());
                RUNNING = false;
                cv.notify_all();
                cv.wait(lk, [&] { return RUNNING.load() || ENDED.load(); });
                if (ENDED) return;
                // the user thread may have stepped / edited while paused: pick up its time and the engine's step count
                T_enq = T; enq = 0;
                dv_steps_done(dev(), &base);
                continue;
            }
            if (ENDED) return;
            if (!bpts.empty()) next_bp = *bpts.begin();
        }
        const double step_dt = dt_engine;     // the dt the engine was given at the last resume()/start()
        double t = T_enq;
        int64_t n = 0;
        if (integrator == 2) { const double h = 0.5 * step_dt; while (n < chunk && !(next_bp <= t)) { t += h; t += h; n++; } }   // T += 0.5 * dt twice, src/sim.cu:1789,1798
        else while (n < chunk && !(next_bp <= t)) { t += step_dt; n++; }
        if (dv_run_async(dev(), n, T_enq) != TB_OK) { fail_and_park("step failed: "); return; }
        enq += n;
        T_enq = t;
        inflight.push_back({base + enq, t, std::chrono::steady_clock::now(), n});
        while (inflight.size() >= 2) {        // wait for the older batch; the newer one keeps the device busy
            const Batch b = inflight.front();
            inflight.pop_front();
            if (dv_wait_steps(dev(), b.upto) != TB_OK) { fail_and_park("step failed: "); return; }
            T = b.T_after;
            cv.notify_all();
            // about one millisecond of device time per batch (the wait above returns once per batch in steady state)
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b.t_enq).count() / 2.0;
            if (ms > 0 && b.n == chunk) {
                const double want = (double)chunk / std::max(ms, 1e-3);
                chunk = (int64_t)std::min(1000.0, std::max(50.0, 0.5 * (double)chunk + 0.5 * want));
            }
        }
    }
}

void Simulation::resume() {   // src/sim.cu:1684-1702
    if (ENDED) throw std::runtime_error("The simulation has ended. Cannot resume the simulation.");
    if (!STARTED) throw std::runtime_error("The simulation has not started. You cannot resume a simulation before calling sim.start().");
    if (masses.size() == 0) throw std::runtime_error("No masses have been added. Please add masses before starting the simulation.");
    throw_if_gpu_error();
    if (RUNNING) return;
    sync_topology();
    if (constraints_dirty) upload_constraints();   // planes/balls created while paused, new dt or g (update_constraints, :1756-1771)
    engine_check(dv_run_async(dev(), 0, T), "tb_run_async");
    {
        std::lock_guard<std::mutex> lk(mtx);
        RUNNING = true;
    }
    cv.notify_all();
}

void Simulation::step(int n) {
    if (ENDED) throw std::runtime_error("The simulation has ended. Control functions cannot be called.");
    if (!STARTED) throw std::runtime_error("The simulation has not started. Call sim.start() and pause before sim.step().");
    if (RUNNING) throw std::runtime_error(kRunning);
    if (n <= 0) return;
    sync_topology();
    if (constraints_dirty) upload_constraints();
    engine_check(dv_run_async(dev(), n, T), "tb_run_async");
    double t = T;
    if (integrator == 2) { const double h = 0.5 * dt_engine; for (int j = 0; j < n; j++) { t += h; t += h; } }
    else for (int j = 0; j < n; j++) t += dt_engine;
    T = t;
    engine_check(dv_sync(dev()), "tb_sync");
}

void Simulation::setBreakpoint(double time) {   // src/sim.cu:814-820
    if (ENDED) throw std::runtime_error("The simulation has ended. Cannot set breakpoints after the end of the simulation run.");
    std::lock_guard<std::mutex> lk(mtx);
    bpts.insert(time);
}

void Simulation::pause(double t) {   // src/sim.cu:1843-1849
    if (ENDED) throw std::runtime_error("The simulation has ended. Control functions cannot be called.");
    setBreakpoint(t);
    waitForEvent();
}

void Simulation::wait(double t) {   // src/sim.cu:1851-1860
    if (ENDED) throw std::runtime_error("The simulation has ended. Control functions cannot be called.");
    const double until = time() + t;
    std::unique_lock<std::mutex> lk(mtx);
    while (RUNNING && T <= until) cv.wait_for(lk, std::chrono::microseconds(200));
}

void Simulation::waitUntil(double t) {   // src/sim.cu:1862-1870
    if (ENDED && !FREED) throw std::runtime_error("The simulation has ended. Control functions cannot be called.");
    std::unique_lock<std::mutex> lk(mtx);
    while (RUNNING && T <= t) cv.wait_for(lk, std::chrono::microseconds(200));
}

void Simulation::waitForEvent() {   // src/sim.cu:1872-1880
    if (ENDED && !FREED) throw std::runtime_error("The simulation has ended. Control functions cannot be called.");
    {
        std::unique_lock<std::mutex> lk(mtx);
        cv.wait(lk, [&] { return !RUNNING; });
    }
    throw_if_gpu_error();
}

double Simulation::time() { return T; }
bool Simulation::running() { return RUNNING; }

void Simulation::stop() { stop(time()); }   // src/sim.cu:1517-1531

void Simulation::stop(double t) {   // src/sim.cu:1533-1545
    if (RUNNING) {
        setBreakpoint(t);
        waitForEvent();
    }
    {
        std::lock_guard<std::mutex> lk(mtx);
        ENDED = true;
    }
    cv.notify_all();
    if (gpu_thread.joinable()) gpu_thread.join();
    freeGPU();
}

// Everything the simulation created is released; user-held pointers dangle afterwards, as in the reference
// (src/sim.cu:131-182).
void Simulation::freeGPU() {
    if (FREED) return;
    if (group) { tb_group_destroy(group); group = nullptr; engine = nullptr; }
    if (engine) { tb_destroy(engine); engine = nullptr; }
    delete hs; hs = nullptr;               // its pinned arrays went with the engine
    for (Spring *s : springs) delete s;
    for (Mass *m : masses) delete m;
    for (Mass *m : zombies) delete m;
    for (Container *c : containers) delete c;
    for (Constraint *c : constraints) delete c;
    springs.clear(); masses.clear(); zombies.clear(); containers.clear(); constraints.clear();
    dev_planes.clear(); dev_balls.clear(); dev_local.clear();
    FREED = true;
    ENDED = true;
}

void Simulation::reset() {   // src/sim.cu:102-129 (which leaks; this one releases)
    if (STARTED && !FREED) {
        if (RUNNING) { setBreakpoint(time()); waitForEvent(); }
        { std::lock_guard<std::mutex> lk(mtx); ENDED = true; }
        cv.notify_all();
        if (gpu_thread.joinable()) gpu_thread.join();
    }
    FREED = false;
    freeGPU();
    bpts.clear();
    RUNNING = false; STARTED = false; ENDED = false; FREED = false; GPU_DONE = false;
    topology_dirty = false; append_dirty = false; constraints_dirty = true; local_dirty = false;
    dev_nm = dev_ns = 0;
    T = 0;
    _global_acc = Vec(0, 0, -9.81);
}

Simulation::~Simulation() {   // src/sim.cu:184-220
    std::cerr << "Simulation destructor called." << std::endl;
    if (STARTED && !FREED) {
        if (RUNNING) {
            try { setBreakpoint(time()); waitForEvent(); } catch (...) {}
        }
        { std::lock_guard<std::mutex> lk(mtx); ENDED = true; }
        cv.notify_all();
    }
    if (gpu_thread.joinable()) gpu_thread.join();
    freeGPU();
}

}  // namespace titan
