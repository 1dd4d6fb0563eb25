/*
 * titan_oracle.c -- CPU oracle for the Titan mass-spring step.  TEST INFRASTRUCTURE ONLY.
 * See titan_oracle.h for the scope statement and parity status.  Never part of the product path.
 *
 * Build: gcc -O2 -std=c99 -ffp-contract=off -fopenmp -fPIC -shared titan_oracle.c -o libtitan_oracle.so -lm
 *
 * Arithmetic notes (they decide bit patterns):
 *  - Vec::norm() is sqrt(pow(x,2)+pow(y,2)+pow(z,2)) (include/Titan/vec.h:149-151).  pow(x,2) is x*x
 *    exactly both in glibc/gcc (folded) and in CUDA (__nv_powi), so it is written x*x here.
 *  - dot(a,b) is (a*b).sum() = (a0*b0 + a1*b1) + a2*b2 (src/vec.cu:39-41, vec.h:153-155).
 *  - scalar*Vec and Vec*scalar both evaluate component*scalar (vec.h:117-123); Vec/scalar is three
 *    true divisions (vec.h:133-135).
 */
#include "titan_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TO_NORMAL 20000.0 /* src/object.cu:29 */

/* ------------------------------------------------------------------ Vec ---------------------------------- */

double to_norm(const double v[3]) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

double to_dot(const double a[3], const double b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

void to_cross(const double a[3], const double b[3], double out[3]) {
    /* src/vec.cu:43-45 -- note the middle component's operand order */
    double x = a[1] * b[2] - a[2] * b[1];
    double y = b[0] * a[2] - a[0] * b[2];
    double z = a[0] * b[1] - a[1] * b[0];
    out[0] = x; out[1] = y; out[2] = z;
}

/* ------------------------------------------------------------------ constraints -------------------------- */

/* ContactPlane ctor, include/Titan/object.h:101-109: normal / normal.norm() */
void to_plane_init(to_plane *p, const double normal[3], double offset, double fk, double fs) {
    double n = to_norm(normal);
    for (int c = 0; c < 3; c++) p->normal[c] = normal[c] / n;
    p->offset = offset;
    p->friction_k = fk;
    p->friction_s = fs;
}

/* CudaContactPlane::applyForce, src/object.cu:76-109 */
void to_contact_plane_apply(const to_plane *p, const double pos[3], const double vel[3], double force[3]) {
    const double *n = p->normal;
    double disp = to_dot(pos, n) - p->offset;
    double fdn = to_dot(force, n);
    double f_normal[3] = { n[0] * fdn, n[1] * fdn, n[2] * fdn };

    if (disp < 0 && (p->friction_s > 0 || p->friction_k > 0)) {
        double vdn = to_dot(vel, n);
        double v_perp[3] = { vel[0] - n[0] * vdn, vel[1] - n[1] * vdn, vel[2] - n[2] * vdn };
        double v_norm = to_norm(v_perp);
        if (v_norm > 1e-16) {
            double friction_mag = p->friction_k * to_norm(f_normal);
            for (int c = 0; c < 3; c++) force[c] -= v_perp[c] * friction_mag / v_norm;
        } else {
            double f_perp[3] = { force[0] - f_normal[0], force[1] - f_normal[1], force[2] - f_normal[2] };
            if (p->friction_s * to_norm(f_normal) > to_norm(f_perp)) {
                for (int c = 0; c < 3; c++) force[c] -= f_perp[c];
            }
        }
    }
    /* (disp < 0) ? -disp * NORMAL * _normal : 0 * _normal   -- the 0*n term is added too (signed zeros) */
    double scale = (disp < 0) ? (-disp * TO_NORMAL) : 0.0;
    for (int c = 0; c < 3; c++) force[c] += n[c] * scale;
}

/* CudaBall::applyForce, src/object.cu:56-59 */
void to_ball_apply(const to_ball *b, const double pos[3], double force[3]) {
    double d[3] = { pos[0] - b->center[0], pos[1] - b->center[1], pos[2] - b->center[2] };
    double dist = to_norm(d);
    if (dist <= b->radius) {
        for (int c = 0; c < 3; c++) force[c] += d[c] * TO_NORMAL / dist;
    } else {
        for (int c = 0; c < 3; c++) force[c] += 0.0;
    }
}

/* CudaConstraintPlane::applyForce, src/object.cu:118-127 (mutates vel) */
void to_constraint_plane_apply(const double n[3], double friction, double vel[3], double force[3]) {
    double normal_force = to_dot(force, n);
    for (int c = 0; c < 3; c++) force[c] += (-n[c]) * normal_force;
    double v_norm = to_norm(vel);
    if (v_norm >= 1e-16) {
        double vdn = to_dot(vel, n);
        for (int c = 0; c < 3; c++) vel[c] += (-n[c]) * vdn;
        double coef = -friction * normal_force;
        for (int c = 0; c < 3; c++) force[c] += vel[c] * coef / v_norm;
    }
}

/* CudaDirection::applyForce, src/object.cu:136-144 (mutates vel) */
void to_direction_apply(const double t[3], double friction, double vel[3], double force[3]) {
    double fdt = to_dot(force, t);
    double nf[3] = { force[0] - t[0] * fdt, force[1] - t[1] * fdt, force[2] - t[2] * fdt };
    for (int c = 0; c < 3; c++) force[c] += -nf[c];
    if (to_norm(vel) >= 1e-16) {
        double vdt = to_dot(vel, t);
        for (int c = 0; c < 3; c++) vel[c] = t[c] * vdt;
        double coef = -to_norm(nf) * friction;
        for (int c = 0; c < 3; c++) force[c] += t[c] * coef;
    }
}

/* ------------------------------------------------------------------ spring force ------------------------- */

/* One spring of computeSpringForces, src/sim.cu:1161-1197.  Returns 0 when the spring is skipped
 * (null or invalid endpoint, :1163); otherwise writes the force that is ADDED to the right endpoint
 * (its negation goes to the left one) and performs the in-place _rest mutation of actuated springs. */
static int spring_force(to_scene *s, int64_t i, double t, double force[3]) {
    int32_t l = s->left[i], r = s->right[i];
    if (l < 0 || r < 0 || !s->valid[l] || !s->valid[r]) return 0;

    const double *pl = s->pos + 3 * (int64_t)l, *pr = s->pos + 3 * (int64_t)r;
    double temp[3] = { pr[0] - pl[0], pr[1] - pl[1], pr[2] - pl[2] };

    double scale = 1.0;
    int type = s->type[i];
    if (type == TO_ACTIVE_CONTRACT_THEN_EXPAND) {
        scale = (1 - 0.2 * sin(s->omega[i] * t));
    } else if (type == TO_ACTIVE_EXPAND_THEN_CONTRACT) {
        scale = (1 + 0.2 * sin(s->omega[i] * t));
    } else if (type == TO_ACTUATED_EXPAND) {
        if (s->rest[i] < s->l_max[i]) s->rest[i] += s->rate[i] * s->dt;
    } else if (type == TO_ACTUATED_CONTRACT) {
        if (s->rest[i] > s->l_min[i]) s->rest[i] -= s->rate[i] * s->dt;
    }

    double L = to_norm(temp);
    double dir[3] = { temp[0] / L, temp[1] / L, temp[2] / L };
    double hooke = s->k[i] * (s->rest[i] * scale - L);
    for (int c = 0; c < 3; c++) force[c] = dir[c] * hooke;

    const double *vl = s->vel + 3 * (int64_t)l, *vr = s->vel + 3 * (int64_t)r;
    double dv[3] = { vl[0] - vr[0], vl[1] - vr[1], vl[2] - vr[2] };
    double damp = to_dot(dv, dir) * s->damping[i];
    for (int c = 0; c < 3; c++) force[c] += dir[c] * damp;
    return 1;
}

static void spring_scatter(to_scene *s, int64_t i, const double force[3]) {
    int32_t l = s->left[i], r = s->right[i];
    if (!s->fixed[r]) {
        double *f = s->force + 3 * (int64_t)r;
        for (int c = 0; c < 3; c++) f[c] += force[c];
    }
    if (!s->fixed[l]) {
        double *f = s->force + 3 * (int64_t)l;
        for (int c = 0; c < 3; c++) f[c] += -force[c];
    }
}

void to_spring_forces(to_scene *s, double t) {
    double f[3];
    for (int64_t i = 0; i < s->ns; i++)
        if (spring_force(s, i, t, f)) spring_scatter(s, i, f);
}

void to_spring_forces_order(to_scene *s, double t, const int64_t *order, int64_t n) {
    double f[3];
    for (int64_t j = 0; j < n; j++)
        if (spring_force(s, order[j], t, f)) spring_scatter(s, order[j], f);
}

/* ------------------------------------------------------------------ magnet pass (a7) --------------------- */

#define OG_DIM 256      /* include/Titan/sim.h:179 */
#define OG_OFFSET 128   /* :180 */
#define OG_CAP 128      /* :181 */
#define OG_CELL 0.04    /* :182 */

/* get2DCellValue, src/sim.cu:824-834 */
static int og_cell(double raw) {
    int v = (int)floorf((float)(raw / OG_CELL)) + OG_OFFSET;
    if (v > OG_DIM - 1) v = OG_DIM - 1; else if (v < 0) v = 0;
    return v;
}

typedef struct { int32_t *counter; int32_t *cells; double *pos0; } og_t;

/* recomputeOG, src/sim.cu:893-906 / computeOgIdx :865-886.  Insertion order inside a cell is
 * scheduling-dependent in the reference (atomicAdd); the oracle inserts in ascending mass index. */
static void og_build(const to_scene *s, og_t *og) {
    memset(og->counter, 0, sizeof(int32_t) * OG_DIM * OG_DIM);
    for (int64_t i = 0; i < s->nm; i++) {
        if (!s->valid[i]) continue;
        int x = og_cell(s->pos[3 * i]), y = og_cell(s->pos[3 * i + 1]);
        int32_t *cnt = &og->counter[OG_DIM * x + y];
        if (*cnt < OG_CAP) { og->cells[(int64_t)(OG_DIM * x + y) * OG_CAP + *cnt] = (int32_t)i; (*cnt)++; }
    }
    memcpy(og->pos0, s->pos, sizeof(double) * 3 * s->nm);
}

/* computeExternalMagnetForcesOG + computeExternalMagnetForce, src/sim.cu:1223-1281; other masses are
 * read at their start-of-step positions (the reference reads them while they are being updated). */
static void magnet_forces(const to_scene *s, const og_t *og, int64_t i, double ext[3]) {
    const double *p1 = og->pos0 + 3 * i;
    double rad1 = s->rad ? s->rad[i] : 0.0, st1 = s->stiffness ? s->stiffness[i] : 0.0;
    double mm1 = s->max_mag ? s->max_mag[i] : 0.0;
    int cx = og_cell(p1[0]), cy = og_cell(p1[1]);
    for (int a = -1; a <= 1; a++) for (int b = -1; b <= 1; b++) {
        int x = cx + a, y = cy + b;
        if (x < 0 || y < 0 || !(x < OG_DIM) || !(y < OG_DIM)) continue;
        int32_t cnt = og->counter[OG_DIM * x + y];
        for (int32_t q = 0; q < cnt; q++) {
            int64_t j = og->cells[(int64_t)(OG_DIM * x + y) * OG_CAP + q];
            if (j == i) continue;
            const double *p2 = og->pos0 + 3 * j;
            double temp[3] = { p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2] };
            double tn = to_norm(temp);
            double inter = tn - (rad1 + (s->rad ? s->rad[j] : 0.0));
            double ef[3] = { 0, 0, 0 };
            if (tn < 0.14) {
                if (inter < 0.0) {
                    double c1 = fabs(inter) * st1;
                    for (int c = 0; c < 3; c++) ef[c] += (temp[c] / tn) * c1;
                }
                double tn2 = tn * tn;
                double c2 = (s->mag_scale ? s->mag_scale[j] : 0.0) * mm1 / (tn2 > 1e-12 ? tn2 : 1e-12);
                for (int c = 0; c < 3; c++) ef[c] -= (temp[c] / tn) * c2;
            }
            for (int c = 0; c < 3; c++) ext[c] += ef[c];
        }
    }
}

/* ------------------------------------------------------------------ mass update -------------------------- */

/* one mass of massForcesAndUpdate, src/sim.cu:1289-1366, default Euler branch */
/* rk2: 0 = not the RK2 build; 1 = massForcesAndUpdate<true> (half step, :1336-1343); 2 = <false> (full step from the
 * backup with the mid-point velocity and acceleration, :1344-1348) */
static void mass_update_one(to_scene *s, int64_t i, const og_t *og, int rk2) {
    if (s->fixed[i] == 1) return; /* :1294 -- nothing happens, not even the force reset or T */
    double *force = s->force + 3 * i, *pos = s->pos + 3 * i, *vel = s->vel + 3 * i;
    double *ext = s->ext + 3 * i, *acc = s->acc + 3 * i;
    double m = s->m[i];

    if (og) magnet_forces(s, og, i, ext);                         /* :1298 */
    for (int c = 0; c < 3; c++) force[c] += s->g[c] * m;          /* :1299 */
    for (int c = 0; c < 3; c++) force[c] += ext[c];               /* :1300 */

    for (int j = 0; j < s->n_planes; j++) to_contact_plane_apply(&s->planes[j], pos, vel, force); /* :1303 */
    for (int j = 0; j < s->n_balls; j++) to_ball_apply(&s->balls[j], pos, force);                 /* :1307 */

    if (s->lc_ptr) {                                              /* :1312-1326, kind by kind */
        int64_t b = s->lc_ptr[i], e = s->lc_ptr[i + 1];
        for (int64_t j = b; j < e; j++) if (s->lc[j].kind == TO_CONTACT_PLANE) {
            to_plane p; memcpy(p.normal, s->lc[j].vec, sizeof p.normal);
            p.offset = s->lc[j].num; p.friction_k = 0.0; p.friction_s = 0.0; /* src/object.cu:61-66 */
            to_contact_plane_apply(&p, pos, vel, force);
        }
        for (int64_t j = b; j < e; j++) if (s->lc[j].kind == TO_BALL) {
            to_ball bl; memcpy(bl.center, s->lc[j].vec, sizeof bl.center); bl.radius = s->lc[j].num;
            to_ball_apply(&bl, pos, force);
        }
        for (int64_t j = b; j < e; j++) if (s->lc[j].kind == TO_CONSTRAINT_PLANE)
            to_constraint_plane_apply(s->lc[j].vec, s->lc[j].num, vel, force);
        for (int64_t j = b; j < e; j++) if (s->lc[j].kind == TO_DIRECTION)
            to_direction_apply(s->lc[j].vec, s->lc[j].num, vel, force);
    }

    double vel_norm = to_norm(vel);                               /* :1329-1332, int drag coefficient */
    if (vel_norm != 0.0) {
        double coef = (double)(-s->drag[i]) * (vel_norm * vel_norm);
        for (int c = 0; c < 3; c++) force[c] += vel[c] * coef / vel_norm;
    }

    if (rk2 == 1) {                                               /* #ifdef RK2, step == true, :1337-1343 */
        double *bp = s->bk_pos + 3 * i, *bv = s->bk_vel + 3 * i;
        for (int c = 0; c < 3; c++) acc[c] = force[c] / m;
        for (int c = 0; c < 3; c++) { bv[c] = vel[c]; bp[c] = pos[c]; }
        for (int c = 0; c < 3; c++) pos[c] = pos[c] + (0.5 * vel[c]) * s->dt;
        for (int c = 0; c < 3; c++) vel[c] = vel[c] + (0.5 * acc[c]) * s->dt;
        s->T[i] += 0.5 * s->dt;
        for (int c = 0; c < 3; c++) { force[c] = 0.0; ext[c] = 0.0; }
        return;
    }
    if (rk2 == 2) {                                               /* step == false, :1344-1348 */
        const double *bp = s->bk_pos + 3 * i, *bv = s->bk_vel + 3 * i;
        for (int c = 0; c < 3; c++) acc[c] = force[c] / m;
        for (int c = 0; c < 3; c++) pos[c] = bp[c] + vel[c] * s->dt;
        for (int c = 0; c < 3; c++) vel[c] = bv[c] + acc[c] * s->dt;
        s->T[i] += 0.5 * s->dt;
        for (int c = 0; c < 3; c++) { force[c] = 0.0; ext[c] = 0.0; }
        return;
    }
    if (s->verlet == 1) {                                         /* #elif VERLET, :1350-1354 */
        double dt2 = s->dt * s->dt;                               /* pow(dt, 2) */
        for (int c = 0; c < 3; c++) vel[c] += (0.5 * (acc[c] + force[c] / m)) * s->dt;   /* :1351 */
        for (int c = 0; c < 3; c++) acc[c] = force[c] / m;                               /* :1352 */
        for (int c = 0; c < 3; c++) pos[c] += vel[c] * s->dt + (0.5 * acc[c]) * dt2;     /* :1353 */
    } else {
    for (int c = 0; c < 3; c++) acc[c] = force[c] / m;            /* :1356 */
    for (int c = 0; c < 3; c++) vel[c] = vel[c] + acc[c] * s->dt; /* :1357 */
    vel_norm = to_norm(vel);                                      /* :1359-1360 speed clamp */
    if (vel_norm > 1) for (int c = 0; c < 3; c++) vel[c] = vel[c] / vel_norm;
    for (int c = 0; c < 3; c++) pos[c] = pos[c] + vel[c] * s->dt; /* :1361 */
    }
    s->T[i] += s->dt;                                             /* :1362 */
    for (int c = 0; c < 3; c++) { force[c] = 0.0; ext[c] = 0.0; } /* :1364-1365 */
}

static og_t *og_alloc(const to_scene *s) {
    og_t *og = (og_t *)malloc(sizeof(og_t));
    og->counter = (int32_t *)malloc(sizeof(int32_t) * OG_DIM * OG_DIM);
    og->cells = (int32_t *)malloc(sizeof(int32_t) * (size_t)OG_DIM * OG_DIM * OG_CAP);
    og->pos0 = (double *)malloc(sizeof(double) * 3 * (size_t)s->nm);
    return og;
}
static void og_free(og_t *og) { if (og) { free(og->counter); free(og->cells); free(og->pos0); free(og); } }

void to_mass_update(to_scene *s) {
    og_t *og = NULL;
    if (s->magnets) { og = og_alloc(s); og_build(s, og); }
    for (int64_t i = 0; i < s->nm; i++) if (s->valid[i]) mass_update_one(s, i, og, 0);
    og_free(og);
}

/* Simulation::execute Euler branch, src/sim.cu:1800-1811: springs with t = T, masses, T += dt */
void to_step(to_scene *s, int64_t nsteps) {
    og_t *og = s->magnets ? og_alloc(s) : NULL;
    for (int64_t n = 0; n < nsteps; n++) {
        if (og) og_build(s, og);                                  /* recomputeOG, :1775 */
        if (s->verlet == 2) {                                     /* #ifdef RK2, :1778-1799 (call sites repaired) */
            to_spring_forces(s, s->T_sim);
            for (int64_t i = 0; i < s->nm; i++) if (s->valid[i]) mass_update_one(s, i, og, 1);
            s->T_sim += 0.5 * s->dt;
            to_spring_forces(s, s->T_sim);
            for (int64_t i = 0; i < s->nm; i++) if (s->valid[i]) mass_update_one(s, i, og, 2);
            s->T_sim += 0.5 * s->dt;
            continue;
        }
        to_spring_forces(s, s->T_sim);
        for (int64_t i = 0; i < s->nm; i++) if (s->valid[i]) mass_update_one(s, i, og, 0);
        s->T_sim += s->dt;
    }
    og_free(og);
}

/* src/sim.cu:1707: the loop body runs while NOT (breakpoint <= T) */
int64_t to_run_until(to_scene *s, double breakpoint) {
    int64_t n = 0;
    while (!(breakpoint <= s->T_sim)) { to_step(s, 1); n++; }
    return n;
}

int to_omp_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* OpenMP transcription used only as the timed CPU baseline: one iteration per spring with atomic
 * scatter (mirrors atomicVecAdd, src/vec.cu:33-37), then one iteration per mass. */
void to_step_omp(to_scene *s, int64_t nsteps) {
    og_t *og = s->magnets ? og_alloc(s) : NULL;
    for (int64_t n = 0; n < nsteps; n++) {
        double t = s->T_sim;
        if (og) og_build(s, og);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < s->ns; i++) {
            double f[3];
            if (!spring_force(s, i, t, f)) continue;
            int32_t l = s->left[i], r = s->right[i];
            if (!s->fixed[r]) {
                double *fr = s->force + 3 * (int64_t)r;
                for (int c = 0; c < 3; c++) {
#pragma omp atomic
                    fr[c] += f[c];
                }
            }
            if (!s->fixed[l]) {
                double *fl = s->force + 3 * (int64_t)l;
                for (int c = 0; c < 3; c++) {
#pragma omp atomic
                    fl[c] += -f[c];
                }
            }
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < s->nm; i++) if (s->valid[i]) mass_update_one(s, i, og, 0);
        s->T_sim += s->dt;
    }
    og_free(og);
}

/* ------------------------------------------------------------------ builders ----------------------------- */

void to_lattice_counts(int nx, int ny, int nz, int64_t *nm, int64_t *ns) {
    int64_t a = nx, b = ny, c = nz;
    *nm = a * b * c;
    /* axis springs + face diagonals (both) + body diagonals (all four), from the loops at src/object.cu:250-291 */
    *ns = (a - 1) * b * c + a * (b - 1) * c + a * b * (c - 1)
        + 2 * ((a - 1) * (b - 1) * c + (a - 1) * b * (c - 1) + a * (b - 1) * (c - 1))
        + 4 * (a - 1) * (b - 1) * (c - 1);
}

/* Lattice::Lattice src/object.cu:235-296 and Beam::Beam :299-363 (same topology; Beam fixes the i==0 face) */
void to_lattice_build(const double center[3], const double dims[3], int nx, int ny, int nz, int beam,
                      double *pos, int32_t *left, int32_t *right, double *rest, uint8_t *fixed) {
    int64_t q = 0;
    for (int i = 0; i < nx; i++) for (int j = 0; j < ny; j++) for (int k = 0; k < nz; k++) {
        /* :245 -- (double) i / (nx - 1.0) - 0.5, times dims, plus center */
        double u[3] = { (nx > 1) ? (double)i / (nx - 1.0) - 0.5 : 0,
                        (ny > 1) ? j / (ny - 1.0) - 0.5 : 0,
                        (nz > 1) ? k / (nz - 1.0) - 0.5 : 0 };
        for (int c = 0; c < 3; c++) pos[3 * q + c] = u[c] * dims[c] + center[c];
        if (fixed) fixed[q] = (beam && i == 0) ? 1 : 0;
        q++;
    }
#define IDX(i, j, k) ((int32_t)((k) + (j) * nz + (i) * ny * nz))
#define EMIT(a, b) do { left[e] = (a); right[e] = (b); e++; } while (0)
    int64_t e = 0;
    for (int i = 0; i < nx; i++) for (int j = 0; j < ny; j++) for (int k = 0; k < nz; k++) {
        for (int l = 0; l < ((i != nx - 1) ? 2 : 1); l++)
            for (int m = 0; m < ((j != ny - 1) ? 2 : 1); m++)
                for (int n = 0; n < ((k != nz - 1) ? 2 : 1); n++)
                    if (l != 0 || m != 0 || n != 0) EMIT(IDX(i, j, k), IDX(i + l, j + m, k + n));
        if (k != nz - 1) {
            if (j != ny - 1) EMIT(IDX(i, j, k + 1), IDX(i, j + 1, k));
            if (i != nx - 1) EMIT(IDX(i, j, k + 1), IDX(i + 1, j, k));
            if (j != ny - 1 && i != nx - 1) {
                EMIT(IDX(i, j, k + 1), IDX(i + 1, j + 1, k));
                EMIT(IDX(i + 1, j, k + 1), IDX(i, j + 1, k));
                EMIT(IDX(i, j + 1, k + 1), IDX(i + 1, j, k));
            }
        }
        if (j != ny - 1 && i != nx - 1) EMIT(IDX(i, j + 1, k), IDX(i + 1, j, k));
    }
#undef IDX
#undef EMIT
    for (int64_t s = 0; s < e; s++) { /* :293-295 rest = |right.pos - left.pos| */
        const double *a = pos + 3 * (int64_t)left[s], *b = pos + 3 * (int64_t)right[s];
        double d[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] };
        rest[s] = to_norm(d);
    }
}

/* Cube::Cube src/object.cu:182-199 */
void to_cube_build(const double center[3], double side, double *pos, int32_t *left, int32_t *right, double *rest) {
    for (int i = 0; i < 8; i++) {
        double u[3] = { (double)(i & 1) - 0.5, (double)((i >> 1) & 1) - 0.5, (double)((i >> 2) & 1) - 0.5 };
        for (int c = 0; c < 3; c++) pos[3 * i + c] = u[c] * side + center[c];
    }
    int e = 0;
    for (int i = 0; i < 8; i++) for (int j = i + 1; j < 8; j++) { left[e] = i; right[e] = j; e++; }
    for (int s = 0; s < e; s++) {
        const double *a = pos + 3 * left[s], *b = pos + 3 * right[s];
        double d[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] };
        rest[s] = to_norm(d);
    }
}

/* Container::translate src/object.cu:201-205 */
void to_translate(double *pos, int64_t n, const double d[3]) {
    for (int64_t i = 0; i < n; i++) for (int c = 0; c < 3; c++) pos[3 * i + c] += d[c];
}

/* Container::rotate src/object.cu:207-233 */
void to_rotate(double *pos, const double *m, int64_t n, const double axis[3], double angle) {
    double com[3] = { 0, 0, 0 }, total = 0;
    for (int64_t i = 0; i < n; i++) {
        for (int c = 0; c < 3; c++) com[c] += pos[3 * i + c] * m[i];
        total += m[i];
    }
    for (int c = 0; c < 3; c++) com[c] = com[c] / total;
    double an = to_norm(axis);
    double ax[3] = { axis[0] / an, axis[1] / an, axis[2] / an };
    for (int64_t i = 0; i < n; i++) {
        double *p = pos + 3 * i;
        double temp[3] = { p[0] - com[0], p[1] - com[1], p[2] - com[2] };
        double ta = to_dot(temp, ax);
        double y[3] = { temp[0] - ax[0] * ta, temp[1] - ax[1] * ta, temp[2] - ax[2] * ta };
        double yn = to_norm(y);
        if (yn < 0.0001) continue;
        double planar0 = -sin(angle) * yn, planar1 = cos(angle) * yn;
        double yh[3] = { y[0] / yn, y[1] / yn, y[2] / yn }, cr[3];
        to_cross(ax, yh, cr);
        /* planar[0]*cross + y/|y|*planar[1] + dot(temp,axis)*axis + com, summed left to right */
        for (int c = 0; c < 3; c++) p[c] = ((cr[c] * planar0 + yh[c] * planar1) + ax[c] * ta) + com[c];
    }
}
