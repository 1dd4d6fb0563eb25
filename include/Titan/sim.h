// Titan/sim.h -- titan::Simulation, the public entry point (drop-in for the reference's include/Titan/sim.h:38-129).
//
// Same surface, same two-thread model (user thread + one GPU thread per started simulation), same lifecycle rules
// and exception messages -- but underneath, start() packs the host object graph into the B200 engine behind the C
// ABI of include/titan_b200.h (fp64 structure-of-arrays state, CSR incidence, tiled step kernel under CUDA graphs),
// and the run loop enqueues whole batches of steps instead of 4 launches + 4 device syncs per step
// (reference src/sim.cu:1704-1841).  The number of steps between breakpoints is decided on the host by the very
// same repeated fp64 accumulation `T += dt` the reference performs, so pause(1.0) at dt = 1e-4 runs 10,001 steps
// here as it does there.
//
// Differences a user can observe (all documented in INTEGRATION.md):
//   * lifecycle flags are per instance (the reference's are static, sim.h:150-154): several simulations can coexist;
//   * get(Container*) / set(Container*) touch exactly the container's masses and springs (the reference over-reads,
//     src/sim.cu:1080-1081,696-701);
//   * misuse that made the reference exit(1) throws std::runtime_error instead;
//   * magnetic masses (RobotLink): the occupancy-grid pass sums pair forces in a fixed order from start-of-step
//     positions (the reference's order and read/update interleaving are scheduling-dependent);
//   * step(n) is an extension (BASELINE.json names it; the reference has no such call);
//   * graphics members are absent (headless).
#ifndef TITAN_B200_SIM_H
#define TITAN_B200_SIM_H

#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "mass.h"
#include "object.h"
#include "spring.h"
#include "vec.h"

#define TITAN_B200_FACADE 1   // lets a user program tell this drop-in from the reference headers (multi-GPU constructor, step())
struct tb_engine;  // include/titan_b200.h
struct tb_group;

namespace titan {

class Simulation {
public:
    // Create
    Mass *createMass();
    Mass *createMass(const Vec &pos);
    Spring *createSpring();
    Spring *createSpring(Mass *m1, Mass *m2);

    // Delete
    void deleteMass(Mass *m);
    void deleteSpring(Spring *s);

    // device -> host (valid while paused)
    void get(Mass *m);
    void get(Spring *s);       // brings back _rest only, like the reference
    void get(Container *c);
    // host -> device
    void set(Mass *m);
    void set(Spring *s);
    void set(Container *c);

    void getAll();             // masses only, like the reference
    void setAll();

    // Global constraints
    void createPlane(const Vec &abc, double d);
    void createPlane(const Vec &abc, double d, double FRICTION_K, double FRICTION_S);
    void createBall(const Vec &center, double r);
    void clearConstraints();

    // Containers
    Container *createContainer();
    void deleteContainer(Container *c);
    Cube *createCube(const Vec &center, double side_length);
    Lattice *createLattice(const Vec &center, const Vec &dims, int nx = 10, int ny = 10, int nz = 10);
    RobotLink *createRobotLink(const Vec &pos1, const Vec &pos2, double mass, double max_exp_length,
                               double min_exp_length, double expansion_rate, double k, double magnetic_force,
                               double radius = 0.015);
    Beam *createBeam(const Vec &center, const Vec &dims, int nx = 10, int ny = 10, int nz = 10);
    Container *importFromSTL(const std::string &path, double density = 10.0, int num_rays = 5);  // not built: throws

    // Bulk modifications (host side only; follow with setAll() once started)
    void setAllSpringConstantValues(double k);
    void setAllMassValues(double m);   // ADDS m, like the reference (src/sim.cu:795)
    void setTimeStep(double delta_t);
    double getTimeStep();
    void setGlobalAcceleration(const Vec &global_acc);
    void defaultRestLengths();

    // Control
    void start();
    void stop();
    void stop(double time);
    void pause(double t);
    void resume();
    void reset();
    void setBreakpoint(double time);
    void wait(double t);
    void waitUntil(double t);
    void waitForEvent();
    double time();
    bool running();
    void step(int n);   // extension: advance exactly n steps from a paused state and return paused

    void printPositions();

    // The reference picks its integrator when the LIBRARY is compiled (CMake option VERLET -> -DVERLET,
    // CMakeLists.txt:13,90-93; src/sim.cu:1350-1354).  libTitan.so holds both; a user translation unit compiled with
    // -DVERLET gets the Verlet one through these inline constructors, everything else the default Euler branch.
#if defined(RK2)            /* "overrides Verlet and other methods", CMakeLists.txt:95-100 */
#define TITAN_B200_INTEGRATOR 2
#elif defined(VERLET)
#define TITAN_B200_INTEGRATOR 1
#else
#define TITAN_B200_INTEGRATOR 0
#endif
    struct Init {};
    Simulation() : Simulation(0, TITAN_B200_INTEGRATOR, Init()) {}
    explicit Simulation(int device) : Simulation(device, TITAN_B200_INTEGRATOR, Init()) {}   // extension: choose the CUDA device (the reference always uses device 0)
    Simulation(int device, int integrator, Init);   // integrator: 0 = Euler (default build), 1 = VERLET build, 2 = RK2 build (call sites repaired)
    // extension: ONE simulation over several GPUs of this machine (tb_create_partitioned, include/titan_b200.h).  A single
    // lattice is cut into i-slabs, independent bodies (containers no spring connects) go whole to the GPUs in turn; the
    // boundary masses are exchanged inside the step kernel over NVLink peer memory.  Same API, same results (the per-mass
    // summation order is kept); per-mass constraints (Mass::addConstraint) are single-GPU only.
    explicit Simulation(const std::vector<int> &devices) : Simulation(devices, TITAN_B200_INTEGRATOR, Init()) {}
    Simulation(const std::vector<int> &devices, int integrator, Init);
    ~Simulation();
    Simulation(const Simulation &) = delete;
    Simulation &operator=(const Simulation &) = delete;

    Spring *getSpringByIndex(int i);
    Mass *getMassByIndex(int i);
    Container *getContainerByIndex(int i);

    std::vector<Mass *> masses;
    std::vector<Spring *> springs;
    std::vector<Container *> containers;

private:
    Mass *createMass(Mass *m);
    Spring *createSpring(Spring *s);
    void adopt(Container *c);

    // ---- engine side -------------------------------------------------------------------------------------------
    tb_engine *engine = nullptr;
    tb_group *group = nullptr;             // several GPUs: the engines live in the group, `engine` is its first (non-null = started)
    int device = 0;
    int integrator = 0;
    std::vector<int> devices;
    struct Dev dev() const;
    void engine_check(int rc, const char *what);
    void upload_everything();              // topology + all state (start, and after topology edits)
    void upload_masses(Mass *const *list, size_t count);
    void upload_springs(Spring *const *list, size_t count);
    void download_masses(Mass *const *list, size_t count);
    void upload_constraints();             // global planes/balls + globals (g, dt)
    void upload_local_constraints();
    void sync_topology();                  // create*/delete* while paused reach the device: appended / invalidated in place, or re-uploaded
    bool append_topology();                // new masses / springs at the end of the lists -> tb_add_masses / tb_add_springs
    bool append_dirty = false;
    int64_t dev_nm = 0, dev_ns = 0;        // masses / springs the device holds (deleted ones keep their index)
    struct HostStage;                      // pinned staging arrays + shadow of the parameter arrays the device holds (sim.cpp)
    HostStage *hs = nullptr;
    bool upload_masses_whole();            // setAll / start fast paths: packed by all host cores into pinned memory, parameter
    bool upload_springs_whole();           //   arrays uploaded only when they differ from what the device already holds
    bool download_masses_whole();
    std::vector<std::vector<LocalConstraintRecord>> dev_local;   // per engine slot: what the device currently holds
    bool topology_dirty = false, constraints_dirty = true, local_dirty = false;
    std::vector<Mass *> zombies;           // deleted masses still referenced by springs (freed with the simulation)

    // ---- lifecycle ---------------------------------------------------------------------------------------------
    double dt;
    double dt_engine = 0.0001;   // the time step last uploaded to the engine (upload_constraints): what the run loop advances T by
    std::atomic<double> T;   // advanced by the GPU thread as batches of steps COMPLETE; equals device time whenever paused
    Vec _global_acc;
    std::vector<Constraint *> constraints;
    std::vector<ContactPlane> dev_planes;  // what createPlane pushed to the device list (survives clearConstraints,
    std::vector<Ball> dev_balls;           //  like the reference's d_planes/d_balls, src/sim.cu:2290-2293)
    std::set<double> bpts;

    std::atomic<bool> RUNNING{false}, STARTED{false}, ENDED{false}, FREED{false}, GPU_DONE{false};
    std::mutex mtx;                        // guards bpts and the RUNNING transitions
    std::condition_variable cv;            // signalled on every RUNNING / ENDED change
    std::thread gpu_thread;
    std::string gpu_error;                 // set by the GPU thread if an engine call failed
    void freeGPU();
    void _run();
    void execute();
    void throw_if_gpu_error();
};

}  // namespace titan

#endif
