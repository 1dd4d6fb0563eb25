/*
 * ref_host_shim.cu -- C entry points onto the REFERENCE's own host-callable code.  TEST INFRASTRUCTURE ONLY.
 *
 * Linked against the reference's vec.cu / mass.cu / spring.cu / object.cu (compiled from a headless-
 * patched scratch copy, see patch_ref.py) into oracle/_ref/libtitan_ref_host.so.  It lets the tests run
 * the reference's Vec math, every constraint applyForce (all __host__ __device__), the Lattice / Cube /
 * Beam builders and Container::rotate on the CPU and compare them with oracle/titan_oracle.c bit for bit.
 * Nothing here re-implements reference arithmetic: each function only marshals plain arrays in and out.
 */
#include <Titan/sim.h>

#include <unordered_map>

using namespace titan;

static Vec V(const double *p) { return Vec(p[0], p[1], p[2]); }
static void out3(const Vec &v, double *p) { p[0] = v[0]; p[1] = v[1]; p[2] = v[2]; }

extern "C" {

double ref_norm(const double *v) { return V(v).norm(); }
double ref_dot(const double *a, const double *b) { return dot(V(a), V(b)); }
void ref_cross(const double *a, const double *b, double *o) { out3(cross(V(a), V(b)), o); }

/* global contact plane: ContactPlane ctor -> CudaContactPlane(const ContactPlane&) -> applyForce */
void ref_contact_plane(const double *normal, double offset, double fk, double fs, const double *pos,
                       const double *vel, double *force, double *normal_out) {
    ContactPlane cp(V(normal), offset);
    cp._FRICTION_K = fk;
    cp._FRICTION_S = fs;
    CudaContactPlane p(cp);
    CUDA_MASS m;
    m.pos = V(pos); m.vel = V(vel); m.force = V(force);
    p.applyForce(&m);
    out3(m.force, force);
    if (normal_out) out3(p._normal, normal_out);
}

/* local contact plane as Mass::addConstraint(CONTACT_PLANE, v, num) builds it */
void ref_local_contact_plane(const double *normal, double offset, const double *pos, const double *vel,
                             double *force) {
    CudaContactPlane p(V(normal), offset);
    CUDA_MASS m;
    m.pos = V(pos); m.vel = V(vel); m.force = V(force);
    p.applyForce(&m);
    out3(m.force, force);
}

void ref_ball(const double *center, double radius, const double *pos, double *force) {
    CudaBall b(V(center), radius);
    CUDA_MASS m;
    m.pos = V(pos); m.force = V(force);
    b.applyForce(&m);
    out3(m.force, force);
}

void ref_constraint_plane(const double *normal, double friction, double *vel, double *force) {
    CudaConstraintPlane p(V(normal), friction);
    CUDA_MASS m;
    m.vel = V(vel); m.force = V(force);
    p.applyForce(&m);
    out3(m.vel, vel); out3(m.force, force);
}

void ref_direction(const double *tangent, double friction, double *vel, double *force) {
    CudaDirection d(V(tangent), friction);
    CUDA_MASS m;
    m.vel = V(vel); m.force = V(force);
    d.applyForce(&m);
    out3(m.vel, vel); out3(m.force, force);
}

static void dump_container(Container *c, double *pos, int *left, int *right, double *rest, double *k,
                           unsigned char *fixed) {
    std::unordered_map<Mass *, int> index;
    int n = 0;
    for (Mass *m : c->masses) {
        index[m] = n;
        out3(m->pos, pos + 3 * n);
        if (fixed) { CUDA_MASS cm(*m); fixed[n] = cm.constraints.fixed ? 1 : 0; }
        n++;
    }
    int e = 0;
    for (Spring *s : c->springs) {
        left[e] = index.at(s->_left);
        right[e] = index.at(s->_right);
        rest[e] = s->_rest;
        if (k) k[e] = s->_k;
        e++;
    }
}

static void free_container(Container *c) {
    for (Spring *s : c->springs) delete s;
    for (Mass *m : c->masses) delete m;
    delete c;
}

void ref_lattice_counts(int nx, int ny, int nz, long long *nm, long long *ns) {
    double z[3] = {0, 0, 0}, o[3] = {1, 1, 1};
    Lattice *l = new Lattice(V(z), V(o), nx, ny, nz);
    *nm = (long long)l->masses.size();
    *ns = (long long)l->springs.size();
    free_container(l);
}

void ref_lattice(const double *center, const double *dims, int nx, int ny, int nz, int beam, double *pos,
                 int *left, int *right, double *rest, double *k, unsigned char *fixed) {
    Container *c = beam ? (Container *)new Beam(V(center), V(dims), nx, ny, nz)
                        : (Container *)new Lattice(V(center), V(dims), nx, ny, nz);
    dump_container(c, pos, left, right, rest, k, fixed);
    free_container(c);
}

void ref_cube(const double *center, double side, double *pos, int *left, int *right, double *rest) {
    Cube *c = new Cube(V(center), side);
    dump_container(c, pos, left, right, rest, nullptr, nullptr);
    free_container(c);
}

void ref_rotate(double *pos, const double *mass, long long n, const double *axis, double angle) {
    Container c;
    for (long long i = 0; i < n; i++) c.masses.push_back(new Mass(V(pos + 3 * i), mass[i]));
    c.rotate(V(axis), angle);
    for (long long i = 0; i < n; i++) { out3(c.masses[i]->pos, pos + 3 * i); delete c.masses[i]; }
}

/* struct sizes the survey quotes; handy sanity check that the build flags match (CONSTRAINTS, no GRAPHICS) */
void ref_sizes(int *out) {
    out[0] = (int)sizeof(CUDA_MASS);
    out[1] = (int)sizeof(CUDA_SPRING);
    out[2] = (int)sizeof(Vec);
    out[3] = (int)sizeof(CudaContactPlane);
    out[4] = (int)sizeof(CudaBall);
}

} /* extern "C" */
