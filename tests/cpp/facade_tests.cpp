// facade_tests.cpp -- behaviour tests of the titan::Simulation drop-in, written against <Titan/sim.h> only, in the
// shape of the reference's own tests (test/physics/*_unittest.cpp use gtest; this file carries its own tiny
// harness because gtest is not in the image).
//
//   titan_facade_tests host     lifecycle / bookkeeping rules that need no GPU (run by `pytest -m "not gpu"`)
//   titan_facade_tests gpu      everything that starts a simulation (run by `pytest -m gpu`)
// Exit code = number of failed checks.
#include <Titan/sim.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>

using namespace titan;

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                                          \
    do {                                                                                     \
        g_checks++;                                                                          \
        if (!(cond)) { g_failed++; std::printf("  FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); } \
    } while (0)

static bool throws(const std::function<void()> &f, const char *needle = nullptr) {
    try { f(); } catch (const std::runtime_error &e) {
        if (needle && !std::strstr(e.what(), needle)) { std::printf("  unexpected message: %s\n", e.what()); return false; }
        return true;
    }
    return false;
}
static void section(const char *name) { std::printf("[%s]\n", name); }

// ------------------------------------------------------------------------------------------------ host ----------
static void host_tests() {
    section("defaults and bulk setters");
    {
        Simulation sim;
        CHECK(sim.getTimeStep() == 0.0001);                               // src/sim.cu:78
        CHECK(throws([&] { sim.setTimeStep(0.0); }, "negative or zero")); // src/sim.cu:803-805
        CHECK(throws([&] { sim.setTimeStep(-1.0); }));
        sim.setTimeStep(0.001);
        CHECK(sim.getTimeStep() == 0.001);
        CHECK(throws([&] { sim.start(); }, "No masses have been added"));  // src/sim.cu:1552-1554
        CHECK(throws([&] { sim.resume(); }, "has not started"));           // src/sim.cu:1689-1691
        Mass *a = sim.createMass(Vec(0, 0, 1)), *b = sim.createMass(Vec(0, 3, 5));
        CHECK(a->m == 0.1 && a->T == 0 && a->pos == Vec(0, 0, 1) && a->vel == Vec(0, 0, 0));   // mass.h:18
        Mass *c = sim.createMass();
        CHECK(c->m == 1.0);                                               // src/mass.cu:9
        Spring *s = sim.createSpring(a, b);
        CHECK(s->_k == 10000.0 && s->_rest == 5.0 && s->_type == PASSIVE_SOFT && s->_damping == 0.0 && s->_omega == 0.0);
        Spring *free_spring = sim.createSpring();
        CHECK(free_spring->_left == nullptr && free_spring->_rest == 1.0);
        sim.setAllMassValues(0.5);                                        // ADDS, src/sim.cu:795
        CHECK(a->m == 0.6 && c->m == 1.5);
        sim.setAllSpringConstantValues(123.0);
        CHECK(s->_k == 123.0);
        sim.deleteSpring(free_spring);
        CHECK(sim.springs.size() == 1);
        b->pos = Vec(0, 0, 3);
        sim.defaultRestLengths();
        CHECK(s->_rest == 2.0);
        // get/set before start() only complain on stderr (src/sim.cu:590-593,605-608)
        sim.get(a); sim.set(a); sim.getAll(); sim.setAll();
        CHECK(!sim.running() && sim.time() == 0.0);
        CHECK(sim.getMassByIndex(1) == b && sim.getSpringByIndex(0) == s);
    }
    section("builders and containers");
    {
        Simulation sim;
        Lattice *l = sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 5, 5, 5);   // README.md:13-18
        CHECK(l->masses.size() == 125 && l->springs.size() == 1036);            // Ns(5) = 1036
        CHECK(sim.masses.size() == 125 && sim.springs.size() == 1036 && sim.containers.size() == 1);
        CHECK(l->nx == 5 && l->_dims == Vec(5, 5, 5));
        CHECK(l->masses[0]->pos == Vec(-2.5, -2.5, 7.5) && l->masses[124]->pos == Vec(2.5, 2.5, 12.5));
        CHECK(l->masses[1]->pos == Vec(-2.5, -2.5, 8.75));                      // k is the fastest index
        CHECK(l->springs[0]->_left == l->masses[0] && l->springs[0]->_right == l->masses[1]);
        Cube *c = sim.createCube(Vec(0, 0, 3), 1.0);
        CHECK(c->masses.size() == 8 && c->springs.size() == 28);
        CHECK(c->masses[5]->pos == Vec(0.5, -0.5, 3.5));                        // corner (i&1, i>>1&1, i>>2&1)
        Beam *b = sim.createBeam(Vec(0, 2, 3), Vec(2, 0.5, 0.5), 6, 3, 3);
        CHECK(b->masses.size() == 54);
        Container *g = sim.createContainer();
        g->add(c); g->add(l->masses[0]); g->add(l->springs[0]);
        CHECK(g->masses.size() == 9 && g->springs.size() == 29);
        c->translate(Vec(1, 2, 3));
        CHECK(c->masses[5]->pos == Vec(1.5, 1.5, 6.5));
        c->setMassValues(0.1);                                                  // ADDS, src/object.cu:148
        CHECK(c->masses[0]->m == 0.2);
        c->setSpringConstants(7.0); c->setRestLengths(0.25);
        CHECK(c->springs[27]->_k == 7.0 && c->springs[27]->_rest == 0.25);
        // rotation by pi about z through the centre of mass maps (x,y) -> (-x,-y) within rounding
        Vec before = l->masses[0]->pos;
        l->rotate(Vec(0, 0, 1), 3.14159265358979323846);
        CHECK(std::fabs(l->masses[0]->pos[0] + before[0]) < 1e-12 && std::fabs(l->masses[0]->pos[1] + before[1]) < 1e-12);
        CHECK(l->masses[0]->pos[2] == before[2]);
        size_t nm = sim.masses.size(), ns = sim.springs.size();
        sim.deleteContainer(c);
        CHECK(sim.masses.size() == nm - 8 && sim.springs.size() == ns - 28 && sim.containers.size() == 3);
        RobotLink *r = sim.createRobotLink(Vec(0, 0, 0), Vec(0.1, 0, 0), 0.2, 0.15, 0.1, 0.009, 5000, 10.0);
        CHECK(r->s->_rest == 0.1 && r->s->_l_max == 0.15 && r->ml->isMagnetic());
        CHECK(r->expand() && r->s->_type == ACTUATED_EXPAND);
        CHECK(!r->contract() && r->s->_type == PASSIVE_SOFT);                   // already at minimum length
        CHECK(throws([&] { sim.importFromSTL("x.stl"); }, "not built"));
    }
    section("reference counting");
    {
        Simulation sim;
        Mass *a = sim.createMass(Vec(0, 0, 0)), *b = sim.createMass(Vec(1, 0, 0)), *c = sim.createMass(Vec(2, 0, 0));
        Spring *ab = sim.createSpring(a, b);
        sim.deleteMass(c);                       // unreferenced: gone
        sim.deleteMass(a);                       // still held by ab: stays alive (invalid) until ab goes
        CHECK(sim.masses.size() == 1 && ab->_left == a && a->pos == Vec(0, 0, 0));
        sim.deleteSpring(ab);
        CHECK(sim.springs.empty());
        delete ab;                               // before start() the caller keeps the spring object
    }
    section("no CPU fallback");
    {
        // On a machine without a CUDA device start() must fail loudly (TB_ERR_NO_DEVICE), never step on the CPU.
        Simulation sim;
        sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 3, 3, 3);
        bool threw = throws([&] { sim.start(); });
        if (threw) {
            std::printf("  start() threw (no CUDA device here) -- ok\n");
            CHECK(!sim.running());
            CHECK(throws([&] { sim.resume(); }, "has not started"));
        } else {
            sim.stop();
            CHECK(throws([&] { sim.createMass(); }, "has ended"));
        }
    }
}

// ------------------------------------------------------------------------------------------------ gpu -----------
static double rel_err(double a, double b) { return std::fabs(a - b) / std::fmax(1.0, std::fabs(b)); }

static void gpu_tests() {
    section("README example: step count of pause(t) and free fall");   // README.md:13-18, src/sim.cu:1704-1714
    {
        Simulation sim;
        sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 5, 5, 5);
        sim.createPlane(Vec(0, 0, 1), 0);
        sim.start();
        CHECK(sim.running() || sim.time() >= 0);
        sim.pause(1.0);
        // T is accumulated by repeated `T += dt`: 10,000 additions of 1e-4 give 0.99999999999990..., so the
        // breakpoint at 1.0 is reached after 10,001 steps
        double t = 0; int n = 0;
        while (!(1.0 <= t)) { t += 0.0001; n++; }
        CHECK(n == 10001);
        CHECK(sim.time() == t);
        CHECK(!sim.running());
        sim.getAll();
        // the speed clamp (|v| <= 1 m/s, src/sim.cu:1359-1360) limits the fall: after the clamp engages the lattice
        // descends at exactly 1 m/s
        Mass *m0 = sim.masses[0];
        CHECK(std::fabs(m0->vel.norm() - 1.0) < 1e-9);
        CHECK(m0->pos[2] < 7.5 && m0->pos[2] > 6.4);
        CHECK(m0->T == t);                                        // per-mass clock advanced once per step
        CHECK(rel_err(m0->acceleration()[2], -9.81) < 1e-6);      // rigid free fall: spring forces cancel
        sim.stop();
        CHECK(throws([&] { sim.pause(2.0); }, "has ended"));
        CHECK(throws([&] { sim.createMass(); }, "has ended"));
    }
    section("single mass on a plane");   // test/physics/simple_unittest.cpp:16-22
    {
        Simulation sim;
        Mass *m = sim.createMass(Vec(0, 0, 0.05));
        sim.createPlane(Vec(0, 0, 1), 0);
        sim.start();
        sim.pause(0.5);
        sim.get(m);
        CHECK(m->pos[0] == 0.0 && m->pos[1] == 0.0);
        CHECK(m->pos[2] > -0.01 && m->pos[2] < 0.06);             // bouncing on the penalty plane, never tunnelling
        sim.stop();
    }
    section("external force acts for exactly one step");   // src/sim.cu:1300,1365; test/physics/external_unittest.cpp
    {
        Simulation sim;
        sim.setGlobalAcceleration(Vec(0, 0, 0));
        Mass *m = sim.createMass(Vec(0, 0, 1));
        m->setExternalForce(Vec(0.1, 0, 0));                      // a = F/m = 1 m/s^2 for one step of 1e-4 s
        sim.start();
        sim.pause(0.0);
        sim.step(1);
        sim.get(m);
        CHECK(rel_err(m->vel[0], 1e-4) < 1e-12 && rel_err(m->acceleration()[0], 1.0) < 1e-12);
        sim.step(9);
        sim.get(m);
        CHECK(rel_err(m->vel[0], 1e-4) < 1e-12);                  // no further push
        CHECK(rel_err(m->pos[0], 10 * 1e-4 * 1e-4) < 1e-12);
        CHECK(rel_err(sim.time(), 10 * 1e-4) < 1e-12);
        m->setExternalForce(Vec(0, 0.2, 0));
        sim.set(m);                                               // re-arms the one-shot force
        sim.step(2);
        sim.get(m);
        CHECK(rel_err(m->vel[1], 2e-4) < 1e-12);
        sim.stop();
    }
    section("get / modify / set cycle on a container");   // test/physics/rotate_unittest.cpp:12-39
    {
        Simulation sim;
        Lattice *l = sim.createLattice(Vec(0, 0, 10), Vec(5, 5, 5), 10, 10, 10);
        sim.setAllSpringConstantValues(1E5);
        l->rotate(Vec(0, 0, 1), -0.78);
        sim.createPlane(Vec(0, 0, 1), 0, 10, 10);
        sim.setGlobalAcceleration(Vec(0, 0, -9.8));
        sim.start();
        int cycles = 0;
        while (true) {
            sim.pause(sim.time() + 0.1);
            sim.get(l);
            Vec before = l->masses[999]->pos;
            l->rotate(Vec(0, 0, 1), 0.5);
            sim.set(l);
            Mass probe = *l->masses[999];
            sim.get(l->masses[999]);                              // read back what set() uploaded
            CHECK(l->masses[999]->pos == probe.pos && !(probe.pos == before));
            cycles++;
            if (sim.time() > 0.35) break;
            sim.resume();
        }
        CHECK(cycles == 4);
        sim.stop();
    }
    section("two simulations side by side; breakpoints; wait");   // test/physics/multiagent_unittest.cpp (per-instance state)
    {
        Simulation a, b;
        a.createLattice(Vec(0, 0, 2), Vec(1, 1, 1), 4, 4, 4);
        b.createLattice(Vec(0, 0, 2), Vec(1, 1, 1), 4, 4, 4);
        b.setTimeStep(0.0002);
        a.setBreakpoint(0.02); a.setBreakpoint(0.01);
        a.start(); b.start();
        a.waitForEvent();
        CHECK(std::fabs(a.time() - 0.01) < 2e-4);                 // the earliest breakpoint fires first
        a.resume(); a.waitForEvent();
        CHECK(std::fabs(a.time() - 0.02) < 2e-4);
        b.waitUntil(0.05);
        CHECK(b.time() > 0.05);
        double tb0 = b.time();
        b.wait(0.05);
        CHECK(b.time() > tb0 + 0.05);
        b.pause(b.time());                                        // stops at the next chunk boundary
        double tb1 = b.time();
        CHECK(!b.running());
        a.getAll(); b.getAll();
        CHECK(a.masses[0]->T == a.time() && b.masses[0]->T == tb1);
        a.stop(); b.stop();
    }
    section("topology edits while paused");   // src/sim.cu:231-253,300-322,343-412
    {
        Simulation sim;
        sim.setGlobalAcceleration(Vec(0, 0, 0));
        Mass *a = sim.createMass(Vec(0, 0, 0)), *b = sim.createMass(Vec(1, 0, 0));
        Spring *s = sim.createSpring(a, b);
        s->_rest = 0.5;                                           // stretched: the masses attract
        s->_k = 1.0;                                              // soft, so the 1 m/s speed clamp stays out of the way
        sim.start();
        sim.pause(0.001);
        sim.getAll();
        double va = a->vel[0];
        CHECK(va > 0 && b->vel[0] < 0);
        Mass *c = sim.createMass(Vec(0, 1, 0));                   // new mass + spring while paused
        Spring *t = sim.createSpring(a, c);
        t->_rest = 0.5;
        t->_k = 1.0;
        sim.resume();
        sim.pause(0.002);
        sim.getAll();
        CHECK(a->vel[0] > va && a->vel[1] > 0 && c->vel[1] < 0);  // old state kept, new spring active
        sim.deleteSpring(s);
        double vb = b->vel[0];
        sim.resume();
        sim.pause(0.003);
        sim.getAll();
        CHECK(b->vel[0] == vb);                                   // b is free now: constant velocity
        sim.deleteMass(c);                                        // t still references c: t is skipped from now on
        double vay = a->vel[1];
        sim.resume();
        sim.pause(0.004);
        sim.getAll();
        CHECK(a->vel[1] == vay && sim.masses.size() == 2);
        sim.stop();
    }
    section("deleteContainer while paused: tombstones and frozen masses on the device, the other body is undisturbed");
    {
        // reference: Simulation::deleteContainer src/sim.cu:478-564
        auto run = [&](bool with_b, std::vector<double> &out) {
            Simulation sim;
            Lattice *a = sim.createLattice(Vec(0, 0, 0.3), Vec(0.4, 0.4, 0.4), 5, 5, 5);
            Lattice *b = with_b ? sim.createLattice(Vec(2, 0, 0.3), Vec(0.4, 0.4, 0.4), 5, 5, 5) : nullptr;
            sim.createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
            sim.start();
            sim.pause(0.01);
            if (b) sim.deleteContainer(b);
            CHECK(sim.masses.size() == 125 && sim.springs.size() == 1036 && sim.containers.size() == 1);
            sim.resume();
            sim.pause(0.03);
            sim.getAll();
            for (Mass *q : a->masses) for (int c = 0; c < 3; c++) { out.push_back(q->pos[c]); out.push_back(q->vel[c]); }
            Mass *extra = sim.createMass(Vec(0, 0, 2));           // the object lists and the device slots no longer line up: still consistent
            sim.resume();
            sim.pause(0.031);
            sim.get(extra);
            CHECK(extra->pos[2] < 2.0 && extra->pos[2] > 1.99);
            sim.stop();
        };
        std::vector<double> alone, after_delete;
        run(false, alone);
        run(true, after_delete);
        double worst = 0;
        for (size_t q = 0; q < alone.size() && q < after_delete.size(); q++) worst = std::max(worst, std::fabs(alone[q] - after_delete[q]));
        std::printf("  max |never there - deleted| = %.3e\n", worst);
        CHECK(alone.size() == after_delete.size() && worst <= 1e-12);
    }
    section("RobotLink: magnetic connectors pull each other (occupancy-grid + magnet pass)");
    {
        // reference: createRobotLink src/sim.cu:2201-2222, RobotLink src/object.cu:368-384, magnet pass :1223-1281
        auto gap_after = [](bool attached) {
            Simulation sim;
            sim.setGlobalAcceleration(Vec(0, 0, 0));
            RobotLink *a = sim.createRobotLink(Vec(0, 0, 1), Vec(0.1, 0, 1), 0.2, 0.15, 0.1, 0.009, 5000, 10.0);
            RobotLink *b = sim.createRobotLink(Vec(0.15, 0, 1), Vec(0.25, 0, 1), 0.2, 0.15, 0.1, 0.009, 5000, 10.0);
            if (!attached) { a->detach(); b->detach(); }
            sim.start();
            sim.pause(0.002);
            sim.getAll();
            double gap = b->masses[0]->pos[0] - a->masses[1]->pos[0];
            sim.stop();
            return gap;
        };
        double pulled = gap_after(true), loose = gap_after(false);
        CHECK(pulled < 0.05 - 1e-4);            // the facing connectors (5 cm apart, neighbouring grid cells) approach
        CHECK(std::fabs(loose - 0.05) < 1e-9);  // detached (max_mag_force = 0) and shells apart: nothing acts
    }
}

// Simulation(std::vector<int> devices): one simulation over several GPUs through the same public API.  `second` = 0 puts both
// engines on device 0 (single-GPU box; run with TB_GRID=16 so that they are co-resident), 1 uses a second GPU.  With
// TB_TPM=1 both sides sum a mass's springs in the same order, so the states are identical to the last bit.
static void multi_tests(int second) {
    section("one lattice over two GPUs (i-slabs, halo exchange in the step kernel) equals the single-GPU run");
    {
        auto run = [&](bool multi, std::vector<double> &out) {
            std::unique_ptr<Simulation> sim(multi ? new Simulation(std::vector<int>{0, second}) : new Simulation());
            Lattice *l = sim->createLattice(Vec(0, 0, 0.29), Vec(1.1, 0.5, 0.5), 12, 6, 6);
            int n = 0;
            for (Mass *q : l->masses) { q->vel = Vec(0.2 + 0.001 * (n % 7), 0, -0.3 + 0.002 * (n % 5)); n++; }
            sim->createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
            sim->start();
            sim->pause(0.0);                                                   // a breakpoint before any step was timed
            sim->resume();
            sim->pause(0.01);
            sim->getAll();
            l->masses[100]->pos = l->masses[100]->pos + Vec(0.01, 0, 0);      // edit across the API, both sides alike
            sim->setAll();
            sim->resume();
            sim->pause(0.02);
            sim->getAll();
            for (Mass *q : sim->masses) for (int c = 0; c < 3; c++) { out.push_back(q->pos[c]); out.push_back(q->vel[c]); }
            CHECK(std::fabs(sim->time() - 0.02) < 2e-4);
            sim->stop();
        };
        std::vector<double> one, two;
        run(false, one);
        run(true, two);
        CHECK(one.size() == two.size() && one.size() == 12 * 6 * 6 * 6);
        double worst = 0;
        for (size_t q = 0; q < one.size() && q < two.size(); q++) worst = std::max(worst, std::fabs(one[q] - two[q]));
        std::printf("  max |single - two engines| = %.3e\n", worst);
        CHECK(worst <= 1e-12);
    }
    section("independent bodies go whole to the GPUs (no exchange)");
    {
        auto run = [&](bool multi, std::vector<double> &out) {
            std::unique_ptr<Simulation> sim(multi ? new Simulation(std::vector<int>{0, second}) : new Simulation());
            for (int r = 0; r < 5; r++) {
                Lattice *l = sim->createLattice(Vec(1.0 * r, 0, 0.3 + 0.01 * r), Vec(0.4, 0.4, 0.4), 5, 5, 5);
                int n = 0;
                for (Spring *s : l->springs) { s->_type = (n % 3 == 1) ? ACTIVE_CONTRACT_THEN_EXPAND : (n % 3 == 2) ? ACTIVE_EXPAND_THEN_CONTRACT : PASSIVE_SOFT; s->_omega = 12.566370614359172; n++; }
            }
            sim->createPlane(Vec(0, 0, 1), 0, 0.5, 0.6);
            sim->start();
            sim->pause(0.015);
            sim->getAll();
            for (Mass *q : sim->masses) for (int c = 0; c < 3; c++) out.push_back(q->pos[c]);
            sim->stop();
        };
        std::vector<double> one, two;
        run(false, one);
        run(true, two);
        double worst = 0;
        for (size_t q = 0; q < one.size() && q < two.size(); q++) worst = std::max(worst, std::fabs(one[q] - two[q]));
        std::printf("  max |single - two engines| = %.3e\n", worst);
        CHECK(one.size() == two.size() && worst <= 1e-12);
    }
    section("per-mass constraints are refused on several GPUs");
    {
        Simulation sim(std::vector<int>{0, second});
        Lattice *l = sim.createLattice(Vec(0, 0, 1), Vec(1, 1, 1), 4, 4, 4);
        l->masses[3]->addConstraint(CONSTRAINT_PLANE, Vec(0, 0, 1), 0.1);
        CHECK(throws([&] { sim.start(); }, "multi-GPU"));
    }
}

int main(int argc, char **argv) {
    const std::string what = argc > 1 ? argv[1] : "host";
    if (what == "host") host_tests();
    else if (what == "gpu") gpu_tests();
    else if (what == "multi") multi_tests(argc > 2 ? std::atoi(argv[2]) : 0);
    else { std::fprintf(stderr, "usage: %s host|gpu|multi [second device]\n", argv[0]); return 2; }
    std::printf("%d checks, %d failed\n", g_checks, g_failed);
    return g_failed;
}
