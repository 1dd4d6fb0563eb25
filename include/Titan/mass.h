// Titan/mass.h -- titan::Mass, the user-visible mass handle (drop-in for the reference's include/Titan/mass.h:16-87).
//
// A Mass is a HOST object; the device never sees it.  At Simulation::start()/set() its fields are packed into the
// engine's fp64 structure-of-arrays state (include/titan_b200.h: tb_set_mass_state / tb_set_mass_magnet /
// tb_set_local_constraints) and at get()/getAll() unpacked again -- the field mapping is the reference's
// CUDA_MASS(Mass&) / Mass::operator=(CUDA_MASS&) (src/mass.cu:21-100): m, T, pos, vel, extern_force, magnet fields,
// fixed and constraints go down; pos, vel, acc, T, extern_force come back.
#ifndef TITAN_B200_MASS_H
#define TITAN_B200_MASS_H

#include <cmath>
#include <cstdint>
#include <vector>

#include "object.h"
#include "vec.h"

namespace titan {

class Mass {
public:
    Mass(const Vec &position, double mass = 0.1, bool fixed = false, double radius = 0.0, double stiffness = 1000,
         double maximum_magnet_force = 0.0, double magnet_scale_factor = 0.0);

    double m;    // kg
    double T;    // local time: advanced by dt per step unless the mass is fixed (src/sim.cu:1292-1295,1362)
    Vec pos;     // m
    Vec vel;     // m/s

    // magnet-link properties (include/Titan/mass.h:27-31): once any mass has rad > 0 or max_mag_force != 0 the engine runs
    // the occupancy-grid + magnet pass every step (titan_b200/csrc/magnet.cu)
    double rad;
    double stiffness;
    double max_mag_force;
    double mag_scale_factor;
    bool isMagnetic() { return (bool)std::round(max_mag_force); }
    void setExternalForce(const Vec &v) { extern_force = v; }  // acts for ONE step after each upload (src/sim.cu:1365)
    Vec acceleration() { return acc; }

    int og_idx;  // kept for source compatibility; the engine's occupancy grid is a per-step counting sort, not a table of pointers

    void addConstraint(CONSTRAINT_TYPE type, const Vec &vec, double num);
    void clearConstraints(CONSTRAINT_TYPE type);
    void clearConstraints();
    void setDrag(double C);  // stored as int on the device, like the reference (object.h:213)
    void fix();
    void unfix();

private:
    Vec acc;
    Vec extern_force;
    bool valid;
    int ref_count;
    int32_t slot;           // index in the engine's arrays, -1 while not uploaded
    bool fixed;
    double drag_coefficient;
    std::vector<LocalConstraintRecord> local;   // in insertion order; the engine applies kinds in the reference's order

    Mass();                 // m = 1.0 (src/mass.cu:8-19); magnet fields zero-initialised (the reference leaves them indeterminate)
    void decrementRefCount();

    friend class Simulation;
    friend class Spring;
    friend class Container;
    friend class Lattice;
    friend class Beam;
    friend class Cube;
    friend class RobotLink;
};

}  // namespace titan

#endif
