// csr_step.cu -- generic fused step kernel: one thread per mass gathers the forces of its incident springs
// through the CSR incidence (deterministic: ascending spring index, right entry before left entry), then
// applies the whole mass update.  One launch == one time step; state ping-pongs between two buffers, which
// preserves the reference's Jacobi semantics (all spring forces are evaluated from the previous positions,
// kernel order src/sim.cu:1801-1809).  Replaces computeSpringForces' 6 fp64 atomics per spring
// (src/sim.cu:1187-1197, src/vec.cu:33-37) and massForcesAndUpdate (src/sim.cu:1285-1367).
//
// This kernel handles ANY topology and every feature; the tiled kernel (tiled.cu) is the fast path.
#include "engine.h"

namespace tb {

struct CsrArgs {
    int64_t nm;
    const double *pos_in, *vel_in;
    double *pos_out, *vel_out, *acc;
    const double *m;
    const uint8_t *flags;
    const int32_t *row_ptr, *adj, *left, *right, *type;
    const double *k, *rest, *damping, *omega;
    const double *t_in;
    double *t_out;
    int rk2_stage;           // the reference's RK2 build: 1 = half step, 2 = full step from the backup (the state before stage 1,
                             // which still sits in the output buffer: ping-pong)
};

template <class P, bool DAMPING, bool ACTIVE, bool WRITE_ACC, bool VERLET>
__global__ void __launch_bounds__(128) csr_step_kernel(CsrArgs A, Globals G, MassExtras X) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double t = *A.t_in;
    if (i == 0) *A.t_out = __dadd_rn(t, A.rk2_stage ? __dmul_rn(0.5, G.dt) : G.dt);      // T += dt, src/sim.cu:1811 (RK2: 0.5 dt per stage)
    if (i >= A.nm) return;
    const int64_t n = A.nm;
    uint8_t fl = A.flags[i];
    V3 pos = {A.pos_in[i], A.pos_in[n + i], A.pos_in[2 * n + i]};
    V3 vel = {A.vel_in[i], A.vel_in[n + i], A.vel_in[2 * n + i]};
    if (!(fl & F_VALID) || (fl & F_FIXED)) {          // fixed masses are frozen entirely, src/sim.cu:1292-1295
        A.pos_out[i] = pos.x; A.pos_out[n + i] = pos.y; A.pos_out[2 * n + i] = pos.z;
        A.vel_out[i] = vel.x; A.vel_out[n + i] = vel.y; A.vel_out[2 * n + i] = vel.z;
        return;
    }
    V3 force = {0.0, 0.0, 0.0};
    int e0 = A.row_ptr[i], e1 = A.row_ptr[i + 1];
    for (int e = e0; e < e1; e++) {
        int a = A.adj[e];
        int s = a >> 1;
        bool i_am_right = a & 1;
        int o = i_am_right ? A.left[s] : A.right[s];
        if (!(A.flags[o] & F_VALID)) continue;        // src/sim.cu:1163
        V3 po = {A.pos_in[o], A.pos_in[n + o], A.pos_in[2 * n + o]};
        V3 vo = {0.0, 0.0, 0.0};
        if (P::strict || DAMPING) vo = V3{A.vel_in[o], A.vel_in[n + o], A.vel_in[2 * n + o]};
        SpringArgs sp;
        sp.k = A.k[s];
        double rest = A.rest[s];
        sp.damping = (DAMPING && A.damping) ? A.damping[s] : 0.0;
        if (ACTIVE) rest = P::mul(rest, actuation_scale<P>(A.type[s], A.omega ? A.omega[s] : 0.0, t));
        else if (P::strict) rest = P::mul(rest, 1.0);
        sp.rest_scaled = rest;
        V3 f = i_am_right ? spring_force<P, DAMPING>(po, pos, vo, vel, sp) : spring_force<P, DAMPING>(pos, po, vel, vo, sp);
        if (i_am_right) acc3<P>(force, f);
        else acc3<P>(force, V3{-f.x, -f.y, -f.z});
    }
    V3 acc;
    if (VERLET) acc = V3{A.acc[i], A.acc[n + i], A.acc[2 * n + i]};   // the previous step's acceleration is state
    if (A.rk2_stage) {
        force = mass_forces<P>(G, X, i, fl & F_LOCAL, A.m[i], pos, vel, force);
        double *bvel = (fl & F_LOCAL) ? X.rk2_bvel : nullptr;      // backup taken after the constraints projected the velocity
        if (A.rk2_stage == 1) {
            if (bvel) { bvel[i] = vel.x; bvel[n + i] = vel.y; bvel[2 * n + i] = vel.z; }
            rk2_stage1_tail<P>(G.dt, A.m[i], force, pos, vel, acc);
        } else {
            const V3 bv = bvel ? V3{bvel[i], bvel[n + i], bvel[2 * n + i]} : V3{A.vel_out[i], A.vel_out[n + i], A.vel_out[2 * n + i]};
            rk2_stage2_tail<P>(G.dt, A.m[i], force, V3{A.pos_out[i], A.pos_out[n + i], A.pos_out[2 * n + i]}, bv, pos, vel, acc);
        }
    } else
    mass_update<P, VERLET>(G, X, i, fl & F_LOCAL, A.m[i], pos, vel, force, acc);
    A.pos_out[i] = pos.x; A.pos_out[n + i] = pos.y; A.pos_out[2 * n + i] = pos.z;
    A.vel_out[i] = vel.x; A.vel_out[n + i] = vel.y; A.vel_out[2 * n + i] = vel.z;
    if (WRITE_ACC || VERLET) { A.acc[i] = acc.x; A.acc[n + i] = acc.y; A.acc[2 * n + i] = acc.z; }
}

template <class P, bool DAMPING, bool ACTIVE>
static void launch_w(Engine *e, const CsrArgs &A, const MassExtras &X, bool write_acc) {
    unsigned grid = (unsigned)((e->nm + 127) / 128);
    if (e->flags & TB_INTEGRATOR_VERLET) csr_step_kernel<P, DAMPING, ACTIVE, true, true><<<grid, 128, 0, e->stream>>>(A, e->G, X);
    else if (write_acc) csr_step_kernel<P, DAMPING, ACTIVE, true, false><<<grid, 128, 0, e->stream>>>(A, e->G, X);
    else csr_step_kernel<P, DAMPING, ACTIVE, false, false><<<grid, 128, 0, e->stream>>>(A, e->G, X);
}
template <class P> static void launch_p(Engine *e, const CsrArgs &A, const MassExtras &X, bool write_acc) {
    bool d = e->any_damping, a = e->any_active;
    if (d && a) launch_w<P, true, true>(e, A, X, write_acc);
    else if (d) launch_w<P, true, false>(e, A, X, write_acc);
    else if (a) launch_w<P, false, true>(e, A, X, write_acc);
    else launch_w<P, false, false>(e, A, X, write_acc);
}

int launch_csr_step(Engine *e, bool use_ext, bool write_acc, int rk2_stage) {
    CsrArgs A;
    A.nm = e->nm;
    A.rk2_stage = rk2_stage;
    A.pos_in = e->pos[e->cur]; A.vel_in = e->vel[e->cur];
    A.pos_out = e->pos[e->cur ^ 1]; A.vel_out = e->vel[e->cur ^ 1];
    A.acc = e->acc; A.m = e->m; A.flags = e->mflags;
    A.row_ptr = e->row_ptr; A.adj = e->adj; A.left = e->left; A.right = e->right; A.type = e->type;
    A.k = e->k; A.rest = e->rest; A.damping = e->damping; A.omega = e->omega;
    A.t_in = e->t_dev + e->cur; A.t_out = e->t_dev + (e->cur ^ 1);
    MassExtras X;
    X.ext = e->any_magnet ? e->mag_force : ((use_ext && e->ext) ? e->ext : nullptr);   // magnet.cu folds the uploaded force in
    X.drag = e->drag;
    X.lc_ptr = e->any_local ? e->lc_ptr : nullptr;
    X.lc = e->lc;
    X.nm = e->nm;
    X.rk2_bvel = (rk2_stage && e->any_local) ? e->rk2_bvel : nullptr;
    if (e->any_active && !e->type) return fail(e, TB_ERR_INVALID, "active springs without type array");
    if (e->flags & TB_MATH_STRICT) launch_p<Strict>(e, A, X, write_acc);
    else launch_p<Fast>(e, A, X, write_acc);
    TB_CUDA(e, cudaGetLastError());
    return TB_OK;
}

}  // namespace tb
