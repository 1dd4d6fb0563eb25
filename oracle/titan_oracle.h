/*
 * titan_oracle.h -- CPU oracle for the Titan mass-spring step.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of the reference's per-step algorithm.  It exists so that the
 * CUDA engine can be checked against something that follows the reference line by line; it is
 * never linked into, imported by or called from the product path (titan_b200/).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 *
 * Parity status: the reference ships NO golden vectors and has NO CPU implementation of the two
 * step kernels (they are __global__ only), so the step itself is pinned (a) piecewise against the
 * reference's own host-callable code compiled from /root/reference (oracle/_ref/libtitan_ref_host.so:
 * Vec math, every constraint applyForce, the Lattice/Cube/Beam builders, Container::rotate) and
 * (b) end-to-end against state dumps produced by the reference's own CUDA kernels run on a B200
 * (oracle/_ref/titan_ref_cuda, fixtures under tests/golden/).  See DESIGN.md "Oracle".
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference).
 * Compile with -ffp-contract=off: the oracle's arithmetic is unfused IEEE-754 double.
 */
#ifndef TITAN_ORACLE_H
#define TITAN_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* include/Titan/spring.h:17-18 */
enum {
    TO_PASSIVE_SOFT = 0,
    TO_PASSIVE_STIFF = 1,
    TO_ACTIVE_CONTRACT_THEN_EXPAND = 2,
    TO_ACTIVE_EXPAND_THEN_CONTRACT = 3,
    TO_ACTUATED_EXPAND = 4,
    TO_ACTUATED_CONTRACT = 5
};

/* include/Titan/object.h:225-227 (CONSTRAINT_TYPE) */
enum { TO_CONSTRAINT_PLANE = 0, TO_CONTACT_PLANE = 1, TO_BALL = 2, TO_DIRECTION = 3 };

/* CudaContactPlane, include/Titan/object.h:138-149.  normal is stored already normalised. */
typedef struct { double normal[3]; double offset; double friction_k; double friction_s; } to_plane;
/* CudaBall, include/Titan/object.h:90-99 */
typedef struct { double center[3]; double radius; } to_ball;
/* one local constraint record of a mass; `kind` is a CONSTRAINT_TYPE.  vec = normal / center /
 * tangent (already normalised where the reference normalises), num = friction / offset / radius
 * (src/mass.cu:104-122). */
typedef struct { int32_t kind; int32_t pad; double vec[3]; double num; } to_local;

/* All arrays are caller-owned.  xyz arrays are [n][3] row-major. */
typedef struct {
    int64_t nm, ns;
    /* masses (CUDA_MASS, include/Titan/mass.h:89-126) */
    double *pos, *vel, *acc, *force, *ext;   /* nm*3 */
    double *m, *T;                           /* nm */
    double *rad, *stiffness, *max_mag, *mag_scale; /* nm; may be NULL (= all zero) */
    uint8_t *fixed, *valid;                  /* nm */
    int32_t *drag;                           /* nm; CUDA_LOCAL_CONSTRAINTS::drag_coefficient is an int */
    /* springs (CUDA_SPRING, include/Titan/spring.h:77-97); left/right = mass index or -1 (nullptr) */
    int32_t *left, *right, *type;            /* ns */
    double *k, *rest, *damping, *omega, *l_max, *l_min, *rate; /* ns */
    /* global constraints, creation order */
    int32_t n_planes, n_balls;
    to_plane *planes;
    to_ball *balls;
    /* local constraints: CSR over masses; lc_ptr may be NULL (= none) */
    int64_t *lc_ptr;                         /* nm+1 */
    to_local *lc;
    double g[3];
    double dt;
    double T_sim;                            /* Simulation::T */
    int32_t magnets;                         /* run the occupancy-grid magnet pass (a7) */
    int32_t verlet;                          /* integrator: 0 = default Euler branch (src/sim.cu:1355-1362); 1 = the reference's
                                                -DVERLET build (:1350-1354), acc is then persistent state; 2 = its -DRK2 build
                                                (:1336-1349, loop :1778-1799 with the call sites repaired to pass the arguments
                                                the kernel declares -- as shipped they do not compile) */
    double *bk_pos, *bk_vel;                 /* nm*3 each, RK2 only: CUDA_MASS::__rk2_backup_pos/vel (include/Titan/mass.h:111-114) */
} to_scene;

/* ---- the step (src/sim.cu:1157-1200, 1285-1367, loop 1704-1714 + 1800-1811) ---- */
void to_spring_forces(to_scene *s, double t);              /* sequential, ascending spring index */
void to_spring_forces_order(to_scene *s, double t, const int64_t *order, int64_t n); /* explicit evaluation order */
void to_mass_update(to_scene *s);
void to_step(to_scene *s, int64_t nsteps);                 /* nsteps x {springs(T); masses; T += dt} */
int64_t to_run_until(to_scene *s, double breakpoint);      /* while !(bp <= T) step; returns #steps */
/* OpenMP transcription (scatter with `omp atomic`, like the reference's atomicAdd); timing baseline */
void to_step_omp(to_scene *s, int64_t nsteps);
int to_omp_threads(void);

/* ---- pieces exported for unit pinning against the reference's host-callable code ---- */
double to_norm(const double v[3]);                                            /* include/Titan/vec.h:149-151 */
double to_dot(const double a[3], const double b[3]);                          /* src/vec.cu:39-41 */
void to_cross(const double a[3], const double b[3], double out[3]);           /* src/vec.cu:43-45 */
void to_plane_init(to_plane *p, const double normal[3], double offset, double fk, double fs);
void to_contact_plane_apply(const to_plane *p, const double pos[3], const double vel[3], double force[3]);
void to_ball_apply(const to_ball *b, const double pos[3], double force[3]);
void to_constraint_plane_apply(const double n[3], double friction, double vel[3], double force[3]);
void to_direction_apply(const double t[3], double friction, double vel[3], double force[3]);

/* ---- builders (src/object.cu:182-199, 235-296, 299-363) ---- */
void to_lattice_counts(int nx, int ny, int nz, int64_t *nm, int64_t *ns);
/* fills pos[nm*3], left/right[ns], rest[ns]; fixed (may be NULL) gets the Beam rule when beam != 0 */
void to_lattice_build(const double center[3], const double dims[3], int nx, int ny, int nz, int beam,
                      double *pos, int32_t *left, int32_t *right, double *rest, uint8_t *fixed);
void to_cube_build(const double center[3], double side, double *pos /*8*3*/, int32_t *left /*28*/,
                   int32_t *right, double *rest);
/* Container::rotate / translate, src/object.cu:201-233 */
void to_rotate(double *pos, const double *m, int64_t n, const double axis[3], double angle);
void to_translate(double *pos, int64_t n, const double d[3]);

#ifdef __cplusplus
}
#endif
#endif
