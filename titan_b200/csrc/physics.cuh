// physics.cuh -- per-spring and per-mass arithmetic of the Titan step, shared by every step kernel.
//
// Two arithmetic policies, selected at engine creation (include/titan_b200.h TB_MATH_*):
//   Strict : every operation is an unfused IEEE-754 round-to-nearest intrinsic (__dmul_rn, __dadd_rn,
//            __ddiv_rn, __dsqrt_rn) in the reference's operation order, so results carry the same bit
//            patterns as the reference's expressions evaluated without FMA contraction.
//   Fast   : same formulas, FMA contraction allowed, one reciprocal square root per spring and one
//            reciprocal per mass instead of the reference's repeated true divisions.
// Reference lines: computeSpringForces src/sim.cu:1157-1200, massForcesAndUpdate src/sim.cu:1285-1367,
// constraint applyForce src/object.cu:56-144, Vec operators include/Titan/vec.h:109-155.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/titan_b200.h"

namespace tb {

#define TB_NORMAL 20000.0  // contact stiffness, src/object.cu:29

struct Strict {
    static constexpr bool strict = true;
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
};

struct Fast {
    static constexpr bool strict = false;
    static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
    static __device__ __forceinline__ double add(double a, double b) { return a + b; }
    static __device__ __forceinline__ double sub(double a, double b) { return a - b; }
    static __device__ __forceinline__ double div(double a, double b) { return a / b; }
    static __device__ __forceinline__ double sqrt(double a) { return ::sqrt(a); }
};

struct V3 { double x, y, z; };

// 1/x for the Fast policy: MUFU.RCP64H seed + two Newton steps (relative error ~1e-16 for normal x); no IEEE
// special-case path, a third of the dependent chain of a true division
__device__ __forceinline__ double rcp_fast(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}

// (a*b).sum(): (a0*b0 + a1*b1) + a2*b2, src/vec.cu:39-41
template <class P> __device__ __forceinline__ double dot3(const V3 &a, const V3 &b) {
    return P::add(P::add(P::mul(a.x, b.x), P::mul(a.y, b.y)), P::mul(a.z, b.z));
}
// sqrt(pow(x,2)+pow(y,2)+pow(z,2)), include/Titan/vec.h:149-151 (pow(x,2) == x*x exactly)
template <class P> __device__ __forceinline__ double norm3(const V3 &a) {
    return P::sqrt(P::add(P::add(P::mul(a.x, a.x), P::mul(a.y, a.y)), P::mul(a.z, a.z)));
}
template <class P> __device__ __forceinline__ V3 sub3(const V3 &a, const V3 &b) {
    return {P::sub(a.x, b.x), P::sub(a.y, b.y), P::sub(a.z, b.z)};
}
template <class P> __device__ __forceinline__ V3 scale3(const V3 &a, double s) {
    return {P::mul(a.x, s), P::mul(a.y, s), P::mul(a.z, s)};
}
template <class P> __device__ __forceinline__ V3 div3(const V3 &a, double s) {
    return {P::div(a.x, s), P::div(a.y, s), P::div(a.z, s)};
}
template <class P> __device__ __forceinline__ void acc3(V3 &a, const V3 &b) {
    a.x = P::add(a.x, b.x); a.y = P::add(a.y, b.y); a.z = P::add(a.z, b.z);
}
template <class P> __device__ __forceinline__ void dec3(V3 &a, const V3 &b) {
    a.x = P::sub(a.x, b.x); a.y = P::sub(a.y, b.y); a.z = P::sub(a.z, b.z);
}

// Spring parameters as one record; the actuation scale (1 -/+ 0.2 sin(omega t)) is resolved by the caller.
struct SpringArgs { double k, rest_scaled, damping; };

// Force added to the RIGHT endpoint of a spring (the left one receives its negation), src/sim.cu:1166-1185.
// pl/pr, vl/vr: positions and velocities of the left/right endpoints.
template <class P, bool DAMPING>
__device__ __forceinline__ V3 spring_force(const V3 &pl, const V3 &pr, const V3 &vl, const V3 &vr,
                                           const SpringArgs &s) {
    V3 temp = sub3<P>(pr, pl);
    if (P::strict) {
        double L = norm3<P>(temp);
        V3 dir = div3<P>(temp, L);
        double hooke = P::mul(s.k, P::sub(s.rest_scaled, L));
        V3 f = scale3<P>(dir, hooke);
        // the reference evaluates the damping term even when damping == 0 (signed zeros matter)
        double damp = P::mul(dot3<P>(sub3<P>(vl, vr), dir), s.damping);
        acc3<P>(f, scale3<P>(dir, damp));
        return f;
    } else {
        double L2 = temp.x * temp.x + temp.y * temp.y + temp.z * temp.z;
        double inv = rsqrt(L2);
        double L = L2 * inv;
        double coef = s.k * (s.rest_scaled - L);
        if (DAMPING) {
            V3 dv = {vl.x - vr.x, vl.y - vr.y, vl.z - vr.z};
            coef += (dv.x * temp.x + dv.y * temp.y + dv.z * temp.z) * inv * s.damping;
        }
        coef *= inv;
        return {temp.x * coef, temp.y * coef, temp.z * coef};
    }
}

// actuation scale, src/sim.cu:1168-1173; types 4/5 mutate _rest in a separate pre-pass (advance_rest).
template <class P> __device__ __forceinline__ double actuation_scale(int type, double omega, double t) {
    if (type == TB_ACTIVE_CONTRACT_THEN_EXPAND) return P::sub(1.0, P::mul(0.2, sin(P::mul(omega, t))));
    if (type == TB_ACTIVE_EXPAND_THEN_CONTRACT) return P::add(1.0, P::mul(0.2, sin(P::mul(omega, t))));
    return 1.0;
}

// CudaContactPlane::applyForce, src/object.cu:76-109
template <class P>
__device__ __forceinline__ void contact_plane(const tb_plane &p, const V3 &pos, const V3 &vel, V3 &force) {
    V3 n = {p.normal[0], p.normal[1], p.normal[2]};
    double disp = P::sub(dot3<P>(pos, n), p.offset);
    if (!P::strict && !(disp < 0)) return;  // outside: only +-0 would be added
    V3 f_normal = scale3<P>(n, dot3<P>(force, n));
    if (disp < 0 && (p.friction_s > 0 || p.friction_k > 0)) {
        V3 v_perp = sub3<P>(vel, scale3<P>(n, dot3<P>(vel, n)));
        double v_norm = norm3<P>(v_perp);
        if (v_norm > 1e-16) {
            double friction_mag = P::mul(p.friction_k, norm3<P>(f_normal));
            dec3<P>(force, div3<P>(scale3<P>(v_perp, friction_mag), v_norm));
        } else {
            V3 f_perp = sub3<P>(force, f_normal);
            if (P::mul(p.friction_s, norm3<P>(f_normal)) > norm3<P>(f_perp)) dec3<P>(force, f_perp);
        }
    }
    double s = (disp < 0) ? P::mul(-disp, TB_NORMAL) : 0.0;
    acc3<P>(force, scale3<P>(n, s));
}

// CudaBall::applyForce, src/object.cu:56-59
template <class P> __device__ __forceinline__ void ball(const tb_ball &b, const V3 &pos, V3 &force) {
    V3 d = sub3<P>(pos, V3{b.center[0], b.center[1], b.center[2]});
    double dist = norm3<P>(d);
    if (dist <= b.radius) acc3<P>(force, div3<P>(scale3<P>(d, TB_NORMAL), dist));
    else if (P::strict) acc3<P>(force, V3{0.0, 0.0, 0.0});
}

// CudaConstraintPlane::applyForce, src/object.cu:118-127
template <class P> __device__ __forceinline__ void constraint_plane(const tb_local &c, V3 &vel, V3 &force) {
    V3 n = {c.vec[0], c.vec[1], c.vec[2]}, nn = {-n.x, -n.y, -n.z};
    double normal_force = dot3<P>(force, n);
    acc3<P>(force, scale3<P>(nn, normal_force));
    double v_norm = norm3<P>(vel);
    if (v_norm >= 1e-16) {
        acc3<P>(vel, scale3<P>(nn, dot3<P>(vel, n)));
        acc3<P>(force, div3<P>(scale3<P>(vel, P::mul(-c.num, normal_force)), v_norm));
    }
}

// CudaDirection::applyForce, src/object.cu:136-144
template <class P> __device__ __forceinline__ void direction(const tb_local &c, V3 &vel, V3 &force) {
    V3 t = {c.vec[0], c.vec[1], c.vec[2]};
    V3 nf = sub3<P>(force, scale3<P>(t, dot3<P>(force, t)));
    acc3<P>(force, V3{-nf.x, -nf.y, -nf.z});
    if (norm3<P>(vel) >= 1e-16) {
        vel = scale3<P>(t, dot3<P>(vel, t));
        acc3<P>(force, scale3<P>(t, P::mul(-norm3<P>(nf), c.num)));
    }
}

// Global constraint set handed to the kernels by value: the first TB_INLINE_C of each kind live in the
// kernel parameter (constant) bank, any more are read from global memory.
#define TB_INLINE_C 4
struct Globals {
    double g[3];
    double dt;
    int32_t n_planes, n_balls;
    tb_plane planes[TB_INLINE_C];
    tb_ball balls[TB_INLINE_C];
    const tb_plane *more_planes;  // planes[TB_INLINE_C..n_planes)
    const tb_ball *more_balls;
};

struct MassExtras {                // rarely-used per-mass data; null pointers = feature absent
    const double *ext;             // [3][nm] one-shot external force, consumed by this step (or null)
    const int32_t *drag;           // [nm] (or null)
    const int32_t *lc_ptr;         // [nm+1] local-constraint CSR (or null)
    const tb_local *lc;
    int64_t nm;                    // plane stride of SoA arrays
    double *rk2_bvel = nullptr;    // [3][nm] RK2 build with local constraints: the velocity backup of stage 1 is taken AFTER the
                                   // constraints have projected the velocity (src/sim.cu:1312-1326 mutate mass.vel, :1338 copies it)
};

// Rarely-taken parts of the mass update live out of line: the step kernels' hot loop is instruction-cache bound
// (measured: +15 KB of inert code in the loop cost 9-18 % on B200), so only what every mass executes stays inline.
struct PVF { V3 pos, vel, force; };

// local constraints of one mass, src/sim.cu:1312-1326: contact planes, balls, constraint planes, directions --
// each kind in record order; always evaluated with the exact (Strict) forms
static __device__ __noinline__ PVF local_constraints(const MassExtras X, int64_t i, PVF s) {
    int b = X.lc_ptr[i], e = X.lc_ptr[i + 1];
#pragma unroll 1
    for (int j = b; j < e; j++) if (X.lc[j].kind == TB_CONTACT_PLANE) {
        tb_plane p = {{X.lc[j].vec[0], X.lc[j].vec[1], X.lc[j].vec[2]}, X.lc[j].num, 0.0, 0.0};
        contact_plane<Strict>(p, s.pos, s.vel, s.force);
    }
#pragma unroll 1
    for (int j = b; j < e; j++) if (X.lc[j].kind == TB_BALL) {
        tb_ball bl = {{X.lc[j].vec[0], X.lc[j].vec[1], X.lc[j].vec[2]}, X.lc[j].num};
        ball<Strict>(bl, s.pos, s.force);
    }
#pragma unroll 1
    for (int j = b; j < e; j++) if (X.lc[j].kind == TB_CONSTRAINT_PLANE) constraint_plane<Strict>(X.lc[j], s.vel, s.force);
#pragma unroll 1
    for (int j = b; j < e; j++) if (X.lc[j].kind == TB_DIRECTION) direction<Strict>(X.lc[j], s.vel, s.force);
    return s;
}

// drag, src/sim.cu:1329-1332 (C is an int on the device, include/Titan/object.h:213)
template <class P> static __device__ __noinline__ V3 drag_force(V3 vel, V3 force, int c) {
    double vn = norm3<P>(vel);
    if (vn != 0.0) {
        double coef = P::mul((double)(-c), P::mul(vn, vn));
        acc3<P>(force, div3<P>(scale3<P>(vel, coef), vn));
    }
    return force;
}

// Everything massForcesAndUpdate adds to the spring-force sum before it integrates (src/sim.cu:1299-1332): gravity,
// external force, global planes and balls, local constraints (may also change vel), drag.
template <class P>
__device__ __forceinline__ V3 mass_forces(const Globals &G, const MassExtras &X, int64_t i, bool has_local, double m,
                                          V3 &pos, V3 &vel, V3 force) {
    acc3<P>(force, scale3<P>(V3{G.g[0], G.g[1], G.g[2]}, m));                       // :1299
    if (X.ext) acc3<P>(force, V3{X.ext[i], X.ext[X.nm + i], X.ext[2 * X.nm + i]});   // :1300
    else if (P::strict) acc3<P>(force, V3{0.0, 0.0, 0.0});
    // global planes in creation order, then global balls (:1303-1309); one code instance each, looped
#pragma unroll 1
    for (int j = 0; j < G.n_planes; j++) contact_plane<P>(j < TB_INLINE_C ? G.planes[j] : G.more_planes[j - TB_INLINE_C], pos, vel, force);
#pragma unroll 1
    for (int j = 0; j < G.n_balls; j++) ball<P>(j < TB_INLINE_C ? G.balls[j] : G.more_balls[j - TB_INLINE_C], pos, force);

    if (has_local) {                                                                // :1312-1326
        PVF s = local_constraints(X, i, PVF{pos, vel, force});
        pos = s.pos; vel = s.vel; force = s.force;
    }

    if (P::strict || X.drag) force = drag_force<P>(vel, force, X.drag ? X.drag[i] : 0);   // :1329-1332
    return force;
}

// default integrator (src/sim.cu:1355-1362): acc = f/m; vel += acc dt; speed clamp; pos += vel dt
template <class P>
__device__ __forceinline__ void euler_tail(double dt, double m, const V3 &force, V3 &pos, V3 &vel, V3 &acc_out) {
    if (P::strict) {                                                                // :1356-1361
        acc_out = div3<P>(force, m);
        vel = V3{P::add(vel.x, P::mul(acc_out.x, dt)), P::add(vel.y, P::mul(acc_out.y, dt)), P::add(vel.z, P::mul(acc_out.z, dt))};
        double vn = norm3<P>(vel);
        if (vn > 1) vel = div3<P>(vel, vn);                                         // speed clamp :1359-1360
        pos = V3{P::add(pos.x, P::mul(vel.x, dt)), P::add(pos.y, P::mul(vel.y, dt)), P::add(pos.z, P::mul(vel.z, dt))};
    } else {
        double rm = rcp_fast(m);
        acc_out = V3{force.x * rm, force.y * rm, force.z * rm};
        vel = V3{vel.x + acc_out.x * dt, vel.y + acc_out.y * dt, vel.z + acc_out.z * dt};
        double v2 = vel.x * vel.x + vel.y * vel.y + vel.z * vel.z;
        if (v2 > 1.0) { double r = rsqrt(v2); vel = V3{vel.x * r, vel.y * r, vel.z * r}; }
        pos = V3{pos.x + vel.x * dt, pos.y + vel.y * dt, pos.z + vel.z * dt};
    }
}

// the reference's -DVERLET branch (src/sim.cu:1350-1354): vel += 0.5 * (acc + force / m) * dt;  acc = force / m;
// pos += vel * dt + 0.5 * acc * pow(dt, 2).  acc carries the previous step's acceleration in, the new one out.
template <class P>
__device__ __forceinline__ void verlet_tail(double dt, double m, const V3 &force, V3 &pos, V3 &vel, V3 &acc) {
    V3 a = P::strict ? div3<P>(force, m) : scale3<P>(force, rcp_fast(m));
    V3 half = scale3<P>(V3{P::add(acc.x, a.x), P::add(acc.y, a.y), P::add(acc.z, a.z)}, 0.5);
    acc3<P>(vel, scale3<P>(half, dt));
    acc = a;
    V3 step = scale3<P>(vel, dt);
    acc3<P>(step, scale3<P>(scale3<P>(a, 0.5), P::mul(dt, dt)));
    acc3<P>(pos, step);
}

// The reference's -DRK2 build (src/sim.cu:1336-1349 inside massForcesAndUpdate<step>, loop :1778-1799; as shipped its two
// call sites pass 6 arguments to the 12-argument kernel and do not compile -- "repaired" = the same call with the arguments
// the kernel declares).  Midpoint rule in two force evaluations, no speed clamp:
//   stage 1 (step == true):  acc = f/m; backup = (pos, vel); pos += 0.5 vel dt; vel += 0.5 acc dt; T += 0.5 dt
//   stage 2 (step == false): acc = f/m (forces at the midpoint state); pos = backup_pos + vel_mid dt; vel = backup_vel + acc dt
template <class P>
__device__ __forceinline__ void rk2_stage1_tail(double dt, double m, const V3 &force, V3 &pos, V3 &vel, V3 &acc) {
    acc = P::strict ? div3<P>(force, m) : scale3<P>(force, rcp_fast(m));
    acc3<P>(pos, scale3<P>(scale3<P>(vel, 0.5), dt));
    acc3<P>(vel, scale3<P>(scale3<P>(acc, 0.5), dt));
}
template <class P>
__device__ __forceinline__ void rk2_stage2_tail(double dt, double m, const V3 &force, const V3 &bpos, const V3 &bvel, V3 &pos, V3 &vel, V3 &acc) {
    acc = P::strict ? div3<P>(force, m) : scale3<P>(force, rcp_fast(m));
    pos = bpos; acc3<P>(pos, scale3<P>(vel, dt));
    vel = bvel; acc3<P>(vel, scale3<P>(acc, dt));
}

// Everything massForcesAndUpdate does after the spring forces are known (src/sim.cu:1299-1365).
// force: spring-force sum on entry.  Returns new pos/vel through the references; acc_out = force/m.
// VERLET: acc_out carries the mass's PREVIOUS acceleration on entry, and there is no speed clamp.
template <class P, bool VERLET = false>
__device__ __forceinline__ void mass_update(const Globals &G, const MassExtras &X, int64_t i, bool has_local,
                                            double m, V3 &pos, V3 &vel, V3 force, V3 &acc_out) {
    force = mass_forces<P>(G, X, i, has_local, m, pos, vel, force);
    if (VERLET) verlet_tail<P>(G.dt, m, force, pos, vel, acc_out);
    else euler_tail<P>(G.dt, m, force, pos, vel, acc_out);
}

}  // namespace tb
