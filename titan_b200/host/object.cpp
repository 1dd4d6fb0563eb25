// object.cpp -- host-side Mass / Spring / Container and the geometry builders of the titan:: façade.
//
// Bit-exact topology contract: mass order, coordinates, spring emission order and rest
// lengths equal the reference builders' (src/object.cu:182-199 Cube, :235-296 Lattice, :299-363 Beam,
// :368-384 RobotLink); tests/test_facade_host.py checks them against the golden vectors produced by the
// reference's own code.  The reference writes the lattice loop nest twice (Lattice and Beam); here both share
// one table-driven builder that also records the grid shape, so that start() can hand the engine a brick tile
// hint (tb_set_tile_hint).
#include <Titan/mass.h>
#include <Titan/object.h>
#include <Titan/spring.h>

#include <cassert>
#include <cmath>

namespace titan {

// ---------------------------------------------------------------- Mass ------------------------------------------
Mass::Mass()
    : m(1.0), T(0), rad(0.0), stiffness(0.0), max_mag_force(0.0), mag_scale_factor(0.0), og_idx(-1), valid(true),
      ref_count(0), slot(-1), fixed(false), drag_coefficient(0.0) {}

Mass::Mass(const Vec &position, double mass, bool fixed_, double radius, double mag_k, double maximum_magnet_force,
           double magnet_scale_factor)
    : m(mass), T(0), pos(position), rad(radius), stiffness(mag_k), max_mag_force(maximum_magnet_force),
      mag_scale_factor(magnet_scale_factor), og_idx(-1), valid(true), ref_count(0), slot(-1), fixed(fixed_),
      drag_coefficient(0.0) {}

// Normals and tangents are normalised when the constraint is created, as the reference's device-side
// constructors do (src/object.cu:61-66,110-115,128-133); a ball keeps its centre and radius.
void Mass::addConstraint(CONSTRAINT_TYPE type, const Vec &vec, double num) {
    LocalConstraintRecord r;
    r.kind = (int)type;
    r.num = num;
    if (type == BALL) {
        r.vec = vec;
    } else {
        assert(vec.norm() != 0.0);
        r.vec = vec / vec.norm();
    }
    local.push_back(r);
}

void Mass::clearConstraints(CONSTRAINT_TYPE type) {
    std::vector<LocalConstraintRecord> kept;
    for (const LocalConstraintRecord &r : local)
        if (r.kind != (int)type) kept.push_back(r);
    local.swap(kept);
}

void Mass::clearConstraints() { local.clear(); }
void Mass::fix() { fixed = true; }
void Mass::unfix() { fixed = false; }
void Mass::setDrag(double C) { drag_coefficient = C; }

void Mass::decrementRefCount() {
    if (--ref_count == 0) delete this;
}

// ---------------------------------------------------------------- Spring ----------------------------------------
void Spring::defaultLength() { _rest = (_left->pos - _right->pos).norm(); }

void Spring::setLeft(Mass *left) {
    if (_left) _left->decrementRefCount();
    _left = left;
    _left->ref_count++;
}

void Spring::setRight(Mass *right) {
    if (_right) _right->decrementRefCount();
    _right = right;
    _right->ref_count++;
}

// ---------------------------------------------------------------- Container -------------------------------------
void Container::translate(const Vec &displ) {
    for (Mass *q : masses) q->pos += displ;
}

// Rotation about the centre of mass.  The sequence of floating-point operations follows src/object.cu:207-233 so
// that a get / rotate / set cycle (test/physics/rotate_unittest.cpp) uploads the same doubles as the reference.
void Container::rotate(const Vec &axis, double angle) {
    Vec centroid(0, 0, 0);
    double total = 0;
    for (Mass *q : masses) {
        centroid += q->m * q->pos;
        total += q->m;
    }
    centroid = centroid / total;
    const Vec u = axis / axis.norm();
    for (Mass *q : masses) {
        const Vec rel = q->pos - centroid;
        const Vec radial = rel - dot(rel, u) * u;        // component perpendicular to the axis
        if (radial.norm() < 0.0001) continue;            // on the axis
        const Vec planar(-std::sin(angle) * radial.norm(), std::cos(angle) * radial.norm(), 0);
        q->pos = planar[0] * cross(u, radial / radial.norm()) + radial / radial.norm() * planar[1] + dot(rel, u) * u + centroid;
    }
}

void Container::setMassValues(double m) {
    for (Mass *q : masses) q->m += m;
}
void Container::setSpringConstants(double k) {
    for (Spring *s : springs) s->_k = k;
}
void Container::setRestLengths(double len) {
    for (Spring *s : springs) s->_rest = len;
}
void Container::addConstraint(CONSTRAINT_TYPE type, const Vec &v, double d) {
    for (Mass *q : masses) q->addConstraint(type, v, d);
}
void Container::clearConstraints() {
    for (Mass *q : masses) q->clearConstraints();
}
void Container::fix() {
    for (Mass *q : masses) q->fixed = true;
}
void Container::add(Mass *m) { masses.push_back(m); }
void Container::add(Spring *s) { springs.push_back(s); }
void Container::add(Container *c) {
    masses.insert(masses.end(), c->masses.begin(), c->masses.end());
    springs.insert(springs.end(), c->springs.begin(), c->springs.end());
}

// ---------------------------------------------------------------- builders --------------------------------------
Cube::Cube(const Vec &center, double side_length) : _side_length(side_length), _center(center) {
    for (int c = 0; c < 8; c++)
        masses.push_back(new Mass(side_length * (Vec(c & 1, (c >> 1) & 1, (c >> 2) & 1) - Vec(0.5, 0.5, 0.5)) + center));
    for (int a = 0; a < 8; a++)
        for (int b = a + 1; b < 8; b++) springs.push_back(new Spring(masses[a], masses[b]));
    for (Spring *s : springs) s->setRestLength((s->_right->pos - s->_left->pos).norm());
}

namespace {
// The 13 springs a lattice cell (i,j,k) emits, in the reference's order (src/object.cu:250-291): the seven
// forward neighbours in l -> m -> n order, the yz-, xz- and three body anti-diagonals, then the xy anti-diagonal.
// Each entry is (offset of the left mass, offset of the right mass); a spring exists when every +1 offset it
// uses stays inside the grid.
struct CellSpring { int a[3], b[3]; };
const CellSpring kCellSprings[13] = {
    {{0, 0, 0}, {0, 0, 1}}, {{0, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, 1, 1}}, {{0, 0, 0}, {1, 0, 0}},
    {{0, 0, 0}, {1, 0, 1}}, {{0, 0, 0}, {1, 1, 0}}, {{0, 0, 0}, {1, 1, 1}},
    {{0, 0, 1}, {0, 1, 0}}, {{0, 0, 1}, {1, 0, 0}}, {{0, 0, 1}, {1, 1, 0}}, {{1, 0, 1}, {0, 1, 0}},
    {{0, 1, 1}, {1, 0, 0}}, {{0, 1, 0}, {1, 0, 0}},
};
}  // namespace

void Container::build_grid(const Vec &center, const Vec &dims, int nx, int ny, int nz, bool fix_first_plane) {
    grid_nx = nx; grid_ny = ny; grid_nz = nz;
    masses.reserve((size_t)nx * ny * nz);
    for (int i = 0; i < nx; i++)
        for (int j = 0; j < ny; j++)
            for (int k = 0; k < nz; k++) {
                const Vec unit((nx > 1) ? (double)i / (nx - 1.0) - 0.5 : 0, (ny > 1) ? j / (ny - 1.0) - 0.5 : 0,
                               (nz > 1) ? k / (nz - 1.0) - 0.5 : 0);
                Mass *q = new Mass(unit * dims + center);
                if (fix_first_plane && i == 0) q->fixed = true;
                masses.push_back(q);
            }
    auto at = [&](int i, int j, int k) { return masses[(size_t)k + (size_t)j * nz + (size_t)i * ny * nz]; };
    for (int i = 0; i < nx; i++)
        for (int j = 0; j < ny; j++)
            for (int k = 0; k < nz; k++) {
                const bool room[3] = {i != nx - 1, j != ny - 1, k != nz - 1};
                for (const CellSpring &cs : kCellSprings) {
                    bool ok = true;
                    for (int c = 0; c < 3; c++) ok = ok && (room[c] || (cs.a[c] == 0 && cs.b[c] == 0));
                    if (!ok) continue;
                    springs.push_back(new Spring(at(i + cs.a[0], j + cs.a[1], k + cs.a[2]), at(i + cs.b[0], j + cs.b[1], k + cs.b[2])));
                }
            }
    for (Spring *s : springs) s->setRestLength((s->_right->pos - s->_left->pos).norm());
}

Lattice::Lattice(const Vec &center, const Vec &dims, int nx_, int ny_, int nz_)
    : nx(nx_), ny(ny_), nz(nz_), _center(center), _dims(dims) {
    build_grid(center, dims, nx_, ny_, nz_, false);
}

Beam::Beam(const Vec &center, const Vec &dims, int nx_, int ny_, int nz_)
    : nx(nx_), ny(ny_), nz(nz_), _center(center), _dims(dims) {
    build_grid(center, dims, nx_, ny_, nz_, true);
}

// ---------------------------------------------------------------- RobotLink -------------------------------------
RobotLink::RobotLink(const Vec &pos1, const Vec &pos2, double mass, double max_exp_length, double min_exp_length,
                     double expansion_rate, double k, double magnetic_force, double radius)
    : max_length(max_exp_length), min_length(min_exp_length), k_link(k), max_mag_force(magnetic_force) {
    ml = new Mass(pos1, mass, false, radius, 5000.0, max_mag_force, 1.0);   // shell stiffness 5000, src/object.cu:375-376
    mr = new Mass(pos2, mass, false, radius, 5000.0, max_mag_force, 1.0);
    s = new Spring(ml, mr, k_link, min_length, PASSIVE_SOFT, 0.0, max_length, min_length, expansion_rate);
    masses.push_back(ml);
    masses.push_back(mr);
    springs.push_back(s);
}

bool RobotLink::expand() {
    if (max_length <= s->_rest) { s->_type = PASSIVE_SOFT; return false; }
    s->_type = ACTUATED_EXPAND;
    attach();   // an expanding link is always magnetically attached
    return true;
}

bool RobotLink::contract() {
    if (min_length >= s->_rest) { s->_type = PASSIVE_SOFT; return false; }
    s->_type = ACTUATED_CONTRACT;
    return true;
}

bool RobotLink::setLength(double length) {
    const double slack = 0.01 * min_length;
    if (length - s->_rest > slack) { s->_type = ACTUATED_EXPAND; return true; }
    if (length - s->_rest < -slack) { s->_type = ACTUATED_CONTRACT; return true; }
    s->_type = PASSIVE_SOFT;
    return false;
}

bool RobotLink::detach() {
    if (contract()) return false;
    if (ml->isMagnetic()) ml->max_mag_force = 0.0;
    if (mr->isMagnetic()) mr->max_mag_force = 0.0;
    return true;
}

bool RobotLink::attach() {
    if (!ml->isMagnetic()) ml->max_mag_force = max_mag_force;
    if (!mr->isMagnetic()) mr->max_mag_force = max_mag_force;
    return false;
}

void RobotLink::setExpansionRate(double rate) { exp_rate = rate; s->_rate = rate; }
void RobotLink::setRobotMass(double mass) { ml->m = mass / 2; mr->m = mass / 2; }
void RobotLink::setColor(Vec) {}
void RobotLink::setStiffness(double k) { k_link = k; s->_k = k; }

}  // namespace titan
