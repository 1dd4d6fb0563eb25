/*
 * titan_api_driver.cu -- a USER PROGRAM of the public Titan API (include/Titan/sim.h); it contains no step
 * arithmetic.  TEST / BENCH INFRASTRUCTURE ONLY.  One source, two engines:
 *   oracle/_ref/titan_ref_cuda      linked (oracle/build_ref.sh) against the REFERENCE's own, headless-patched
 *                                   vec/mass/spring/object/sim.cu: the golden-fixture generator and the
 *                                   "Titan's own CUDA build on the same B200" baseline of bench.py
 *   tests/cpp/bin/titan_facade_dump compiled unchanged (tests/cpp/Makefile) against this repo's drop-in facade
 *                                   (include/Titan + libTitan.so); tests/test_facade.py compares the two
 * Modes:
 *   dump      <scenario> <out.bin>          initial scene + snapshots after pause(t)
 *   init      <scenario> <out.bin>          initial scene only (host builders; needs no GPU)
 *   bench     <config> <steps>              wall clock around pause() for `steps` steps
 *   bench_e2e <config> <steps> <batches>    end to end through host Mass objects: every batch is
 *                                           setAll() + resume() + pause(t + steps dt) + getAll(), all timed
 *
 * Dump format (little endian): repeated records  u32 name_len, name, u32 dtype(0=f64,1=i32,2=u8), u64 count, data.
 */
#include <Titan/sim.h>

#ifndef DUMP_IMPL                 // the facade build of this same file passes its own name
#define DUMP_IMPL "titan_cuda_reference"
#endif

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

using namespace titan;

static FILE *g_out = nullptr;

static void put(const std::string &name, uint32_t dtype, const void *data, uint64_t count) {
    static const size_t width[3] = {8, 4, 1};
    uint32_t n = (uint32_t)name.size();
    fwrite(&n, 4, 1, g_out);
    fwrite(name.data(), 1, n, g_out);
    fwrite(&dtype, 4, 1, g_out);
    fwrite(&count, 8, 1, g_out);
    fwrite(data, width[dtype], count, g_out);
}
static void put_f64(const std::string &name, const std::vector<double> &v) { put(name, 0, v.data(), v.size()); }
static void put_i32(const std::string &name, const std::vector<int> &v) { put(name, 1, v.data(), v.size()); }

static void dump_masses(Simulation &sim, const std::string &prefix) {
    std::vector<double> pos, vel, acc, m, T;
    for (Mass *q : sim.masses) {
        for (int c = 0; c < 3; c++) { pos.push_back(q->pos[c]); vel.push_back(q->vel[c]); acc.push_back(q->acceleration()[c]); }
        m.push_back(q->m);
        T.push_back(q->T);
    }
    put_f64(prefix + "pos", pos); put_f64(prefix + "vel", vel); put_f64(prefix + "acc", acc);
    put_f64(prefix + "m", m); put_f64(prefix + "T", T);
}

/* magnet fields of the initial scene; written only by scenes that have any, so older dumps keep their record set */
static void dump_magnet_fields(Simulation &sim) {
    std::vector<double> rad, stiff, maxmag, scale;
    bool any = false;
    for (Mass *q : sim.masses) {
        rad.push_back(q->rad); stiff.push_back(q->stiffness); maxmag.push_back(q->max_mag_force); scale.push_back(q->mag_scale_factor);
        any = any || q->rad != 0 || q->max_mag_force != 0 || q->mag_scale_factor != 0;
    }
    if (!any) return;
    put_f64("init.rad", rad); put_f64("init.stiffness", stiff); put_f64("init.max_mag", maxmag); put_f64("init.mag_scale", scale);
}

static void dump_springs(Simulation &sim, const std::string &prefix, bool topology) {
    std::unordered_map<Mass *, int> index;
    int n = 0;
    for (Mass *q : sim.masses) index[q] = n++;
    std::vector<int> left, right, type;
    std::vector<double> k, rest, damping, omega, lmax, lmin, rate;
    for (Spring *s : sim.springs) {
        left.push_back(index.at(s->_left)); right.push_back(index.at(s->_right)); type.push_back((int)s->_type);
        k.push_back(s->_k); rest.push_back(s->_rest); damping.push_back(s->_damping); omega.push_back(s->_omega);
        lmax.push_back(s->_l_max); lmin.push_back(s->_l_min); rate.push_back(s->_rate);
    }
    put_f64(prefix + "rest", rest);
    if (!topology) return;
    put_i32(prefix + "left", left); put_i32(prefix + "right", right); put_i32(prefix + "type", type);
    put_f64(prefix + "k", k); put_f64(prefix + "damping", damping); put_f64(prefix + "omega", omega);
    put_f64(prefix + "l_max", lmax); put_f64(prefix + "l_min", lmin); put_f64(prefix + "rate", rate);
}

/* every spring must carry defined actuator fields before they are dumped / uploaded */
static void define_actuator_fields(Simulation &sim) {
    for (Spring *s : sim.springs) {
        if (s->_type != ACTUATED_EXPAND && s->_type != ACTUATED_CONTRACT) { s->_l_max = 0; s->_l_min = 0; s->_rate = 0; }
    }
}

struct Scenario {
    std::vector<double> pauses;  /* absolute simulation times handed to pause() */
};

/* Scene descriptions.  tests/scenarios.py holds the same descriptions for the oracle / engine side; the
 * dump also carries the full initial state so a mismatch in the two descriptions is caught by the test. */
static Scenario build(Simulation &sim, const std::string &name) {
    Scenario sc;
    if (name == "c1_readme") { /* README.md:13-18 */
        sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 5, 5, 5);
        sim.createPlane(Vec(0, 0, 1), 0);
        sc.pauses = {0.0, 0.001, 0.01, 0.1};
    } else if (name == "contact_friction") {
        Lattice *l = sim.createLattice(Vec(0, 0, 0.499), Vec(1, 1, 1), 6, 6, 6);
        int n = 0;
        for (Mass *q : l->masses) { q->vel = Vec(0.3 + 0.001 * (n % 7), -0.2 + 0.002 * (n % 5), -0.1); n++; }
        n = 0;
        for (Spring *s : l->springs) { if (n % 3 == 0) s->addDamping(0.05 + 0.01 * (n % 4)); n++; }
        sim.createPlane(Vec(0, 0, 2), 0, 0.5, 0.6);
        sim.createPlane(Vec(1, 0, 0.2), -0.6, 0.0, 0.3);
        sc.pauses = {0.0, 0.001, 0.01, 0.05};
    } else if (name == "actuated") {
        Lattice *l = sim.createLattice(Vec(0, 0, 1.0), Vec(0.9, 0.9, 0.9), 4, 4, 4);
        int n = 0;
        for (Spring *s : l->springs) {
            s->_type = (SpringType)(n % 6);
            s->_omega = 4 * 3.14159265358979323846 * (1 + (n % 3));
            s->_damping = (n % 2) ? 0.1 : 0.0;
            s->_l_max = s->_rest * 1.001; s->_l_min = s->_rest * 0.999; s->_rate = 0.05 + 0.01 * (n % 5);
            n++;
        }
        sim.createBall(Vec(0, 0, -0.2), 0.8);
        sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
        sc.pauses = {0.0, 0.001, 0.01, 0.05};
    } else if (name == "local_constraints") {
        Lattice *l = sim.createLattice(Vec(0, 0, 2), Vec(1, 1, 1), 4, 4, 4);
        int n = 0;
        for (Mass *q : l->masses) {
            q->vel = Vec(0.05 * ((n % 3) - 1), 0.04 * ((n % 5) - 2), 0.03 * ((n % 7) - 3));
            if (n % 11 == 0) q->fix();
            if (n % 5 == 1) q->setDrag(2.0 + (n % 3));
            if (n % 7 == 2) q->addConstraint(CONSTRAINT_PLANE, Vec(0.2, 0.1, 1), 0.3);
            if (n % 7 == 3) q->addConstraint(DIRECTION, Vec(1, 0.5, 0.25), 0.2);
            if (n % 7 == 4) q->addConstraint(CONTACT_PLANE, Vec(0, 0.1, 1), 1.7);
            if (n % 7 == 5) q->addConstraint(BALL, Vec(0.1, 0.1, 1.9), 0.6);
            if (n % 13 == 6) { q->addConstraint(CONTACT_PLANE, Vec(0, 0, 1), 1.6); q->addConstraint(DIRECTION, Vec(0, 1, 1), 0.0); }
            if (n % 4 == 0) q->setExternalForce(Vec(0.5, -0.25, 1.0 + 0.1 * n));
            q->m = 0.1 + 0.01 * (n % 4);
            n++;
        }
        sim.setGlobalAcceleration(Vec(0.1, -0.2, -9.8));
        sim.setTimeStep(0.0002);
        sc.pauses = {0.0, 0.002, 0.01};
    } else if (name == "cube_beam") {
        sim.createCube(Vec(0, 0, 3), 1.0);
        Beam *b = sim.createBeam(Vec(0, 2, 3), Vec(2, 0.5, 0.5), 6, 3, 3);
        for (Spring *s : b->springs) s->_k = 5e4;
        sim.createPlane(Vec(0, 0, 1), 2.45, 0.2, 0.3);
        sc.pauses = {0.0, 0.001, 0.02};
    } else if (name == "magnets") {
        /* magnetic connector masses (RobotLink) + the occupancy-grid pass.  Every interaction inside the 0.14 m cut-off
         * is between a FREE and a FIXED mass: the reference reads neighbour positions while their threads may already
         * have moved them (src/sim.cu:1223-1281 inside massForcesAndUpdate), so only such scenes are reproducible. */
        for (int g = 0; g < 5; g++) {
            double x0 = -0.988 + 0.5 * g, len = 0.09;
            RobotLink *l = sim.createRobotLink(Vec(x0, 0.01, 0.5), Vec(x0 + 0.03, 0.02, 0.5 + len), 0.05, 0.12, 0.08, 0.01,
                                               2000.0, 0.004 * (g + 1), 0.015);
            l->masses[0]->fix();
            Mass *p = sim.createMass(Vec(x0 + 0.035 + 0.002 * g, 0.015, 0.5 + len + 0.02 + 0.004 * g));   /* fixed magnet */
            p->rad = 0.015; p->stiffness = 4000.0; p->max_mag_force = 0.003; p->mag_scale_factor = 0.8;
            p->fix();
            sim.createMass(Vec(x0 + 0.03, 0.02, 0.5 + len + 0.35));   /* free, plain, same cell column, beyond the cut-off */
            Mass *r = sim.createMass(Vec(x0 + 0.03, 0.02 - 0.03, 0.5 + len - 0.03));   /* fixed, susceptible only */
            r->mag_scale_factor = 0.5;
            r->fix();
        }
        sc.pauses = {0.0, 0.001, 0.005};
    } else {
        fprintf(stderr, "unknown scenario %s\n", name.c_str());
        exit(2);
    }
    define_actuator_fields(sim);
    return sc;
}

static int do_dump(const std::string &name, const char *path, bool init_only = false) {
    g_out = fopen(path, "wb");
    if (!g_out) { perror(path); return 1; }
    Simulation sim;
    Scenario sc = build(sim, name);
    dump_masses(sim, "init.");
    dump_magnet_fields(sim);
    dump_springs(sim, "init.", true);
    if (init_only) { fclose(g_out); return 0; }
    sim.start();
    int snap = 0;
    for (double t : sc.pauses) {
        if (snap > 0) sim.resume();
        sim.pause(t);
        sim.getAll();
        for (Spring *s : sim.springs) sim.get(s);
        std::string p = "snap" + std::to_string(snap) + ".";
        std::vector<double> tt = {t, sim.time()};
        put_f64(p + "time", tt);
        dump_masses(sim, p);
        dump_springs(sim, p, false);
        snap++;
    }
    fclose(g_out);
    sim.stop();
    return 0;
}

/* bench: wall clock around pause() for `steps` steps, excluding start()'s upload */
static int do_bench(const std::string &cfg, long steps) {
    Simulation sim;
    if (cfg == "c1") {
        sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 5, 5, 5);
        sim.createPlane(Vec(0, 0, 1), 0);
    } else if (cfg == "c2" || cfg.rfind("lattice", 0) == 0) {
        int n = (cfg == "c2") ? 100 : atoi(cfg.c_str() + 7);
        double side = 0.1 * (n - 1);
        sim.createLattice(Vec(0, 0, side / 2 - 0.001), Vec(side, side, side), n, n, n);
        sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
    } else if (cfg == "c3_64") {
        /* SURVEY.md 8(d) C3 restricted to what one Titan Simulation can hold: 64 actuated 10^3 robots on an 8 x 8 grid
         * (1.2 m pitch, inside the +-5.12 m occupancy grid), spring s of each robot passive / contract-then-expand /
         * expand-then-contract by s % 3, omega = 4 pi, one frictional plane */
        for (int r = 0; r < 64; r++) {
            Lattice *l = sim.createLattice(Vec(-4.2 + 1.2 * (r % 8), -4.2 + 1.2 * (r / 8), 0.55), Vec(0.9, 0.9, 0.9), 10, 10, 10);
            int n = 0;
            for (Spring *s : l->springs) {
                s->_type = (n % 3 == 1) ? ACTIVE_CONTRACT_THEN_EXPAND : (n % 3 == 2) ? ACTIVE_EXPAND_THEN_CONTRACT : PASSIVE_SOFT;
                s->_omega = 4 * 3.14159265358979323846;
                n++;
            }
        }
        sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
    } else {
        fprintf(stderr, "unknown bench config %s\n", cfg.c_str());
        return 2;
    }
    size_t ns = sim.springs.size(), nm = sim.masses.size();
    double dt = sim.getTimeStep();
    auto t0 = std::chrono::steady_clock::now();
    sim.start();
    sim.pause(0.0);
    auto t1 = std::chrono::steady_clock::now();
    /* warm-up of 10 steps, then the timed run */
    double T = sim.time();
    sim.resume(); sim.pause(T + 9.5 * dt);
    double Ta = sim.time();
    auto t2 = std::chrono::steady_clock::now();
    sim.resume(); sim.pause(Ta + (steps - 0.5) * dt);
    auto t3 = std::chrono::steady_clock::now();
    double Tb = sim.time();
    long done = lround((Tb - Ta) / dt);
    double sec = std::chrono::duration<double>(t3 - t2).count();
    printf("{\"impl\": \"" DUMP_IMPL "\", \"config\": \"%s\", \"masses\": %zu, \"springs\": %zu, \"steps\": %ld, "
           "\"seconds\": %.6f, \"spring_steps_per_s\": %.6e, \"upload_seconds\": %.3f}\n",
           cfg.c_str(), nm, ns, done, sec, (double)ns * done / sec, std::chrono::duration<double>(t1 - t0).count());
    fflush(stdout);
    sim.stop();
    return 0;
}

/* bench_e2e: what a user of the reference's integration loop pays per batch of `steps` steps when the state crosses the
 * boundary both ways: setAll() (host Mass objects -> device), resume(), pause(), getAll() (device -> host Mass objects).
 * bytes = 48 B (pos, vel) per mass each way is the payload counted; the call moves what its implementation moves. */
static int do_bench_e2e(const std::string &cfg, long steps, long batches) {
    Simulation sim;
    if (cfg == "c1") {
        sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 5, 5, 5);
        sim.createPlane(Vec(0, 0, 1), 0);
    } else if (cfg == "c2" || cfg.rfind("lattice", 0) == 0) {
        int n = (cfg == "c2") ? 100 : atoi(cfg.c_str() + 7);
        double side = 0.1 * (n - 1);
        sim.createLattice(Vec(0, 0, side / 2 - 0.001), Vec(side, side, side), n, n, n);
        sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
    } else {
        fprintf(stderr, "unknown bench_e2e config %s\n", cfg.c_str());
        return 2;
    }
    size_t ns = sim.springs.size(), nm = sim.masses.size();
    double dt = sim.getTimeStep();
    auto t0 = std::chrono::steady_clock::now();
    sim.start();
    sim.pause(0.0);
    auto t1 = std::chrono::steady_clock::now();
    auto batch = [&]() {
        double T = sim.time();
        sim.setAll();
        sim.resume();
        sim.pause(T + (steps - 0.5) * dt);
        sim.getAll();
        return lround((sim.time() - T) / dt);
    };
    for (int w = 0; w < 2; w++) batch();                       /* warm-up */
    long done = 0;
    auto t2 = std::chrono::steady_clock::now();
    for (long b = 0; b < batches; b++) done += batch();
    auto t3 = std::chrono::steady_clock::now();
    double sec = std::chrono::duration<double>(t3 - t2).count(), check = 0;
    for (Mass *q : sim.masses) check += q->pos[2];
    printf("{\"impl\": \"" DUMP_IMPL "\", \"mode\": \"e2e\", \"config\": \"%s\", \"masses\": %zu, \"springs\": %zu, "
           "\"steps\": %ld, \"batches\": %ld, \"seconds\": %.6f, \"spring_steps_per_s\": %.6e, \"upload_seconds\": %.3f, "
           "\"h2d_bytes_per_batch\": %zu, \"d2h_bytes_per_batch\": %zu, \"mean_z\": %.17g, "
           "\"call\": \"Simulation::setAll + resume + pause + getAll\"}\n",
           cfg.c_str(), nm, ns, done, batches, sec, (double)ns * done / sec, std::chrono::duration<double>(t1 - t0).count(),
           nm * 48, nm * 48, check / (double)nm);
    fflush(stdout);
    sim.stop();
    return 0;
}

#ifdef TITAN_B200_FACADE
/* bench_multi: ONE lattice over ndev GPUs through titan::Simulation(std::vector<int>) -- the facade cuts it into i-slabs
 * (tb_create_partitioned), the boundary masses are exchanged inside the step kernel.  Wall clock around pause(). */
static int do_bench_multi(const std::string &cfg, long steps, int ndev) {
    std::vector<int> devices;
    for (int d = 0; d < ndev; d++) devices.push_back(d);
    Simulation sim(devices);
    int n = (cfg == "c2") ? 100 : atoi(cfg.c_str() + 7);
    if (n < 2) { fprintf(stderr, "bench_multi: config latticeN\n"); return 2; }
    double side = 0.1 * (n - 1);
    auto t0 = std::chrono::steady_clock::now();
    sim.createLattice(Vec(0, 0, side / 2 - 0.001), Vec(side, side, side), n, n, n);
    sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
    size_t ns = sim.springs.size(), nm = sim.masses.size();
    double dt = sim.getTimeStep();
    auto t1 = std::chrono::steady_clock::now();
    sim.start();
    sim.pause(0.0);
    auto t2 = std::chrono::steady_clock::now();
    double T = sim.time();
    sim.resume(); sim.pause(T + 99.5 * dt);                    /* warm-up: 100 steps */
    double Ta = sim.time();
    auto t3 = std::chrono::steady_clock::now();
    sim.resume(); sim.pause(Ta + (steps - 0.5) * dt);
    auto t4 = std::chrono::steady_clock::now();
    long done = lround((sim.time() - Ta) / dt);
    double sec = std::chrono::duration<double>(t4 - t3).count();
    sim.getAll();
    double check = 0;
    for (Mass *q : sim.masses) check += q->pos[2];
    printf("{\"impl\": \"" DUMP_IMPL "\", \"mode\": \"multi\", \"config\": \"%s\", \"n_gpus\": %d, \"masses\": %zu, \"springs\": %zu, "
           "\"steps\": %ld, \"seconds\": %.6f, \"spring_steps_per_s\": %.6e, \"build_seconds\": %.3f, \"upload_seconds\": %.3f, \"mean_z\": %.17g}\n",
           cfg.c_str(), ndev, nm, ns, done, sec, (double)ns * done / sec, std::chrono::duration<double>(t1 - t0).count(),
           std::chrono::duration<double>(t2 - t1).count(), check / (double)nm);
    fflush(stdout);
    sim.stop();
    return 0;
}
#endif

int main(int argc, char **argv) {
#ifdef TITAN_B200_FACADE
    if (argc >= 5 && !strcmp(argv[1], "bench_multi")) return do_bench_multi(argv[2], atol(argv[3]), atoi(argv[4]));
#endif
    if (argc >= 5 && !strcmp(argv[1], "bench_e2e")) return do_bench_e2e(argv[2], atol(argv[3]), atol(argv[4]));
    if (argc >= 4 && !strcmp(argv[1], "dump")) return do_dump(argv[2], argv[3]);
    if (argc >= 4 && !strcmp(argv[1], "init")) return do_dump(argv[2], argv[3], true);
    if (argc >= 4 && !strcmp(argv[1], "bench")) return do_bench(argv[2], atol(argv[3]));
    fprintf(stderr, "usage: %s dump <scenario> <out.bin> | bench <c1|c2|c3_64|latticeN> <steps> | bench_e2e <c1|c2|latticeN> <steps> <batches>\n", argv[0]);
    return 2;
}
