// Titan/object.h -- global constraints and containers of the public API (drop-in for the reference's
// include/Titan/object.h:48-330; its CONSTRAINTS build, which is the one CMakeLists.txt produces).
//
// What is different underneath: constraints are plain host records here.  The reference keeps one
// thrust::device_vector per constraint kind PER MASS (object.h:181-201); this engine packs every mass's local
// constraints into one CSR table at upload time (tb_set_local_constraints, include/titan_b200.h).
// Device-side helper structs of the reference header (CudaBall, CudaContactPlane, ...) are not part of the user
// surface and do not exist here; graphics members are omitted (headless).
#ifndef TITAN_B200_OBJECT_H
#define TITAN_B200_OBJECT_H

#include <vector>

#include "vec.h"

namespace titan {

class Mass;
class Spring;
class Simulation;

class Constraint {
public:
    virtual ~Constraint() = default;
};

struct Ball : public Constraint {
    Ball(const Vec &center, double radius) : _radius(radius), _center(center) {}
    double _radius;
    Vec _center;
};

struct ContactPlane : public Constraint {
    // the normal is normalised at construction (reference object.h:101-109)
    ContactPlane(const Vec &normal, double offset)
        : _normal(normal / normal.norm()), _offset(offset), _FRICTION_K(0.0), _FRICTION_S(0.0) {}
    Vec _normal;
    double _offset;
    double _FRICTION_K;
    double _FRICTION_S;
};

enum CONSTRAINT_TYPE { CONSTRAINT_PLANE, CONTACT_PLANE, BALL, DIRECTION };

// One local constraint of a mass as the engine consumes it (mirrors tb_local in include/titan_b200.h):
// vec = normalised normal / tangent, or ball centre; num = friction / offset / radius.
struct LocalConstraintRecord {
    int kind;
    Vec vec;
    double num;
};

class Container {  // groups masses and springs; does not own them unless created by a Simulation builder
public:
    virtual ~Container() {}
    void translate(const Vec &displ);
    void rotate(const Vec &axis, double angle);  // about the centre of mass, radians
    void setMassValues(double m);                // ADDS m to every mass, like the reference (src/object.cu:148)
    void setSpringConstants(double k);
    void setRestLengths(double len);

    void addConstraint(CONSTRAINT_TYPE type, const Vec &v, double d);
    void clearConstraints();
    void fix();

    void add(Mass *m);
    void add(Spring *s);
    void add(Container *c);

    std::vector<Mass *> masses;
    std::vector<Spring *> springs;

protected:
    friend class Simulation;
    // Regular lattices tell the engine their brick decomposition (tile hint) at upload; 0 = not a lattice.
    int grid_nx = 0, grid_ny = 0, grid_nz = 0;
    void build_grid(const Vec &center, const Vec &dims, int nx, int ny, int nz, bool fix_first_plane);
};

class Cube : public Container {
public:
    ~Cube() {}
    Cube(const Vec &center, double side_length = 1.0);
    double _side_length;
    Vec _center;
};

class Lattice : public Container {
public:
    ~Lattice() {}
    Lattice(const Vec &center, const Vec &dims, int nx = 10, int ny = 10, int nz = 10);
    int nx, ny, nz;
    Vec _center, _dims;
};

class Beam : public Container {  // a lattice whose i == 0 face is fixed
public:
    ~Beam() {}
    Beam(const Vec &center, const Vec &dims, int nx = 10, int ny = 10, int nz = 10);
    int nx, ny, nz;
    Vec _center, _dims;
};

class RobotLink : public Container {  // two magnetic masses joined by an actuated spring
public:
    RobotLink(const Vec &pos1, const Vec &pos2, double mass, double max_exp_length, double min_exp_length,
              double expansion_rate, double k, double magnetic_force, double radius = 0.015);
    ~RobotLink() {}

    Mass *ml;
    Mass *mr;
    Spring *s;

    double exp_rate = 0.009;
    double max_length;
    double min_length;
    double k_link;
    double max_mag_force;

    bool expand();
    bool contract();
    bool setLength(double length);
    bool detach();
    bool attach();
    void setExpansionRate(double exp_rate);
    void setRobotMass(double mass);
    void setColor(Vec c);  // headless: no-op
    void setStiffness(double k);
};

}  // namespace titan

#endif
