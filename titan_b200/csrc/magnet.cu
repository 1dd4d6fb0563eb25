// magnet.cu -- the occupancy-grid pass and the short-range magnet / shell-contact forces of magnetic masses
// (RobotLink connectors): recomputeOG src/sim.cu:893-906, computeOgIdx :865-886, get2DCellValue :824-834,
// computeExternalMagnetForcesOG :1250-1281, computeExternalMagnetForce :1223-1241; grid constants
// include/Titan/sim.h:179-182.  SURVEY.md rows a7 / N2.
//
// The reference keeps a 256 x 256 x 128 table of CUDA_MASS pointers (64 MiB) that every mass appends itself to with
// an atomicAdd, so the order inside a cell -- and with it the fp64 summation order of the pair forces -- changes from
// run to run, and each mass reads its neighbours' positions while their own threads may already have moved them.
// Here the grid is a counting sort of the mass indices by cell, rebuilt every step from the start-of-step positions
// (the ping-pong buffer the step kernel reads):
//     og_cell_kernel   cell id per mass + population count per cell
//     og_scan_kernel   exclusive scan of the 65,536 populations (64 CTAs, each sums what lies in front of it)
//     og_fill_kernel   masses into their cell's segment (arbitrary order)
//     og_rank_kernel   ... brought into ascending mass index (rank inside the segment), positions copied alongside
//     magnet_force_kernel   per free mass (threads in sorted order): the 3 x 3 cells in the reference's (x offset, y offset) order, the
//                      first 128 masses of each cell in ascending index, one pair force at a time added to the
//                      mass's external force (the order the CPU oracle of the tests follows too)
// so results are reproducible and, under TB_MATH_STRICT, bit-identical to the oracle.  Cells hold 4 B per mass
// instead of 64 MiB; a cell with more than 128 masses keeps its 128 lowest indices (the reference drops whichever
// threads lose the race and prints from the device).
//
// The output is a full external-force array (uploaded one-shot force + pair forces) which the generic step kernel
// (csr_step.cu) or the tiled one consumes as `extern_force` (src/sim.cu:1298-1300).  In front of the tiled kernel the pass
// reads the tile-blocked state through the mass -> slot map (PosView); cells and pair order stay in API indices, so both
// kernels see the same forces bit for bit.
#include "engine.h"

namespace tb {

#define OG_DIM 256        // occupancy_grid_dim, include/Titan/sim.h:179
#define OG_OFFSET 128     // occupancy_grid_offset, :180
#define OG_CAP 128        // occupancy_grid_max_masses_per_cell, :181
#define OG_CELL_SIZE 0.04 // og_cell_size, :182
#define OG_CELLS (OG_DIM * OG_DIM)
#define MAG_CUTOFF 0.14   // src/sim.cu:1228

// get2DCellValue: int(floorf(raw / cell_size)) + offset, clamped; floorf takes the quotient rounded to float
__device__ __forceinline__ int og_coord(double raw) {
    int v = (int)floorf(__double2float_rn(__ddiv_rn(raw, OG_CELL_SIZE))) + OG_OFFSET;
    return v > OG_DIM - 1 ? OG_DIM - 1 : (v < 0 ? 0 : v);
}

// masses that are not valid take no part: their row of the output is just the uploaded force
__global__ void og_cell_kernel(int64_t nm, const PosView pos, const uint8_t *flags, int32_t *cell, int32_t *counter,
                               const double *ext_in, double *ext_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = -1;
    if (flags[i] & F_VALID) {
        c = OG_DIM * og_coord(pos.ld(i, 0)) + og_coord(pos.ld(i, 1));   // getOgCounterIdx, :835-837
        atomicAdd(&counter[c], 1);
    } else {
        for (int k = 0; k < 3; k++) ext_out[k * nm + i] = ext_in ? ext_in[k * nm + i] : 0.0;
    }
    cell[i] = c;
}

// exclusive scan of the cell populations: start[c] = sum of counter[0..c), start[OG_CELLS] = total; cursor = start.
// 64 CTAs of 1,024 cells each; a CTA adds up everything in front of its cells itself (coalesced, independent loads out
// of L2: at most 63 per thread) instead of waiting for its predecessors -- integers, so the grouping does not matter.
constexpr int OG_SCAN_T = 1024;
__global__ void __launch_bounds__(OG_SCAN_T) og_scan_kernel(const int32_t *counter, int32_t *start, int32_t *cursor) {
    __shared__ int32_t wsum[32], wpre[32];
    const int t = threadIdx.x, b = blockIdx.x, lane = t & 31, w = t >> 5;
    int prev = 0;
#pragma unroll 8
    for (int q = 0; q < b; q++) prev += counter[q * OG_SCAN_T + t];
    const int own = counter[b * OG_SCAN_T + t];
    int incl = own;
    for (int off = 1; off < 32; off <<= 1) { int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
    for (int off = 16; off; off >>= 1) prev += __shfl_xor_sync(0xffffffffu, prev, off);
    if (lane == 31) wsum[w] = incl;
    if (lane == 0) wpre[w] = prev;
    __syncthreads();
    if (w == 0) {
        const int v = wsum[lane];
        int iv = v, pv = wpre[lane];
        for (int off = 1; off < 32; off <<= 1) { int u = __shfl_up_sync(0xffffffffu, iv, off); if (lane >= off) iv += u; }
        for (int off = 16; off; off >>= 1) pv += __shfl_xor_sync(0xffffffffu, pv, off);
        __syncwarp();
        wsum[lane] = iv - v;                  // cells of this CTA in front of warp `lane`
        if (lane == 0) wpre[0] = pv;          // cells in front of this CTA
    }
    __syncthreads();
    const int run = wpre[0] + wsum[w] + incl - own;
    start[b * OG_SCAN_T + t] = run; cursor[b * OG_SCAN_T + t] = run;
    if (b == gridDim.x - 1 && t == OG_SCAN_T - 1) start[OG_CELLS] = run + own;
}

__global__ void og_fill_kernel(int64_t nm, const int32_t *cell, int32_t *cursor, int32_t *items) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = cell[i];
    if (c >= 0) items[atomicAdd(&cursor[c], 1)] = (int32_t)i;
}

// ... and the positions with them (spos, [3][nm] by place in the sorted list): the pair loop then reads addresses it
// knows in advance instead of chasing sorted[q] -> slot -> position
__global__ void og_rank_kernel(int64_t nm, const PosView pos, const int32_t *cell, const int32_t *start, const int32_t *items,
                               int32_t *sorted, double *spos, float4 *sposf) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = cell[i];
    if (c < 0) return;
    int b = start[c], e = start[c + 1], rank = 0;
    for (int q = b; q < e; q++) rank += items[q] < (int32_t)i;
    sorted[b + rank] = (int32_t)i;
    double p[3];
    for (int k = 0; k < 3; k++) { p[k] = pos.ld(i, k); spos[k * nm + b + rank] = p[k]; }
    // single-precision copy + the mass index in one 16-byte word: what the neighbour scan of the pair kernel reads
    sposf[b + rank] = make_float4(__double2float_rn(p[0]), __double2float_rn(p[1]), __double2float_rn(p[2]), __int_as_float((int)i));
}

struct MagnetArgs {
    int64_t nm;
    PosView pos;                // start-of-step positions
    const uint8_t *flags;
    const double *rad, *stiffness, *max_mag, *mag_scale;
    const int32_t *start, *sorted;
    const double *spos;         // positions by place in `sorted`, [3][nm]
    const float4 *sposf;        // (float x, y, z, mass index) by place in `sorted`
    const double *ext_in;       // uploaded one-shot external force or null (= 0)
    double *ext_out;            // [3][nm]
};

#define MAG_D2_GATE 0.0197   // > 0.14^2 with room for the rounding of the sum: farther pairs need no square root
constexpr int MAG_BATCH = 8;
#define MAG_D2_GATE_F 0.0225f // 0.15^2: the single-precision pre-test of the neighbour scan

// computeExternalMagnetForce, src/sim.cu:1223-1241, for one neighbour inside the gate: shell contact + inverse-square
// magnet.  No component of the result is ever -0 (it is 0 + a - b with a, b possibly signed zeros).
template <class P>
__device__ __forceinline__ V3 pair_force(const MagnetArgs &A, const V3 &temp, double d2, double rad1, double st1, double mm1, int j) {
    V3 ef = {0.0, 0.0, 0.0};
    const double tn = P::sqrt(d2);                                                    // norm3: d2 is its radicand
    if (tn < MAG_CUTOFF) {
        const double inter = P::sub(tn, P::add(rad1, A.rad[j]));
        // Fast: reciprocals instead of the four IEEE divisions (whose zero numerators -- a neighbour straight above --
        // take the division's slow path)
        V3 dir = P::strict ? div3<P>(temp, tn) : scale3<P>(temp, rcp_fast(tn));
        if (inter < 0.0) acc3<P>(ef, scale3<P>(dir, P::mul(fabs(inter), st1)));       // shells intersect, :1230-1232
        double tn2 = P::mul(tn, tn);
        tn2 = tn2 > 1e-12 ? tn2 : 1e-12;
        double c2 = P::strict ? P::div(P::mul(A.mag_scale[j], mm1), tn2) : P::mul(P::mul(A.mag_scale[j], mm1), rcp_fast(tn2));
        dec3<P>(ef, scale3<P>(dir, c2));                                              // :1235-1237
    }
    return ef;
}

// One thread per place in the sorted list: the threads of a warp sit in the same cell (or a few cells) and need the same
// nine lists.  Nearly every neighbour of a 2-D cell list is beyond the 0.14 m cut-off (a whole z column shares a cell)
// and adds exactly zero, but in a warp SOME lane is near the current neighbour at almost every step, so a
// straightforward loop runs the expensive branch (square root, four divisions) every time at 1/32 efficiency.  Hence
// two phases per batch of neighbours: a cheap test (single-precision squared distance against a gate above the
// cut-off, on 16-byte records) marks each lane's candidates, then the warp evaluates candidates together, one round
// per candidate.  Order of a mass's sum: unchanged (its candidates in list order); the skipped zero terms only matter
// for a sum that is still -0, which the reference's first `+ 0` would turn into +0: `seen`.
constexpr int MAG_LOCKSTEP_MIN = 4;   // lanes of a warp in the same cell from which the warp walks that cell's lists together

template <class P> __global__ void __launch_bounds__(128) magnet_force_kernel(MagnetArgs A) {
    __shared__ float4 stage[4][32];                               // one batch of neighbour records per warp
    constexpr unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t place = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = place < A.nm && place < A.start[OG_CELLS];   // idle lanes stay for the warp-wide steps
    const int64_t n = A.nm;
    int i = -1;
    V3 ext = {0.0, 0.0, 0.0}, p1 = {0.0, 0.0, 0.0};
    double rad1 = 0.0, st1 = 0.0, mm1 = 0.0;
    bool more = false;
    if (active) {
        i = A.sorted[place];
        if (A.ext_in) ext = V3{A.ext_in[i], A.ext_in[n + i], A.ext_in[2 * n + i]};
        p1 = V3{A.spos[place], A.spos[n + place], A.spos[2 * n + place]};
        rad1 = A.rad[i]; st1 = A.stiffness[i]; mm1 = A.max_mag[i];
        more = !(A.flags[i] & F_FIXED);                           // fixed masses return before the pass, :1292-1298
    }
    const int cx = og_coord(p1.x), cy = og_coord(p1.y);
    // the scan's test runs in single precision on the float copies: gate 0.15 m, i.e. 1 cm of slack for their rounding,
    // enough while this mass is within 10 km of the origin; farther out (or not finite) every neighbour is a candidate
    const float x1f = __double2float_rn(p1.x), y1f = __double2float_rn(p1.y), z1f = __double2float_rn(p1.z);
    const bool wide = !(fabsf(x1f) + fabsf(y1f) + fabsf(z1f) < 1e4f);
    bool seen = false;
    auto evaluate = [&](int qc, int j) {                          // the exact test and the pair force, in fp64
        const V3 temp = sub3<P>(p1, V3{A.spos[qc], A.spos[n + qc], A.spos[2 * n + qc]});
        const double d2 = P::add(P::add(P::mul(temp.x, temp.x), P::mul(temp.y, temp.y)), P::mul(temp.z, temp.z));
        if (d2 < MAG_D2_GATE) acc3<P>(ext, pair_force<P>(A, temp, d2, rad1, st1, mm1, j));       // :1240
    };

    // ---- lanes that share a cell (a z column of a lattice: dozens) walk its nine lists together: a batch of 32 records
    // is loaded once by the warp, every lane tests all of them (a bit per candidate), then the candidates are evaluated
    // in list order, one warp-wide round per bit
    const int key = more ? OG_DIM * cx + cy : -1;
    unsigned todo = __ballot_sync(FULL, key >= 0);
    while (todo) {
        const int lk = __shfl_sync(FULL, key, __ffs(todo) - 1);
        const bool mine = key == lk;
        const unsigned group = __ballot_sync(FULL, mine);
        todo &= ~group;
        if (__popc(group) < MAG_LOCKSTEP_MIN) continue;           // few lanes: each scans for itself below
        const int gx = lk / OG_DIM, gy = lk % OG_DIM;
#pragma unroll 1
        for (int ci = 0; ci < 9; ci++) {                          // 3 * (a + 1) + (b + 1): the reference's loop order, :1256-1258
            const int x = gx + ci / 3 - 1, y = gy + ci % 3 - 1;
            if (x < 0 || y < 0 || !(x < OG_DIM) || !(y < OG_DIM)) continue;           // :1262
            const int c = OG_DIM * x + y;
            const int q0 = A.start[c];
            int q1 = A.start[c + 1];
            if (q1 - q0 > OG_CAP) q1 = q0 + OG_CAP;
#pragma unroll 1
            for (int base = q0; base < q1; base += 32) {
                const int cnt = q1 - base < 32 ? q1 - base : 32;
                __syncwarp();
                if (lane < cnt) stage[w][lane] = A.sposf[base + lane];
                __syncwarp();
                unsigned m = 0;
                if (mine) {
#pragma unroll 4
                    for (int k = 0; k < cnt; k++) {
                        const float4 f = stage[w][k];
                        const float dx = x1f - f.x, dy = y1f - f.y, dz = z1f - f.z;
                        const bool other = __float_as_int(f.w) != i;                  // :1273
                        seen |= other;
                        const bool near = fmaf(dz, dz, fmaf(dy, dy, dx * dx)) < MAG_D2_GATE_F;
                        m |= (unsigned)(other & (wide | near)) << k;                  // no branches: one predicate per record
                    }
                }
                while (__any_sync(FULL, m != 0)) {
                    if (m) {
                        const int k = __ffs(m) - 1;
                        m &= m - 1;
                        evaluate(base + k, __float_as_int(stage[w][k].w));
                    }
                }
            }
        }
        if (mine) more = false;
    }

    // ---- the rest (sparse scenes: a lane's cell is its own): every lane scans its own lists
    int ci = -1, q = 0, q1 = 0;
    while (__any_sync(FULL, more)) {
        int qc = 0, j = 0;
        bool found = false;
        while (more) {                                            // cheap scan to this lane's next candidate
            if (q >= q1) {
                if (++ci >= 9) { more = false; break; }
                const int x = cx + ci / 3 - 1, y = cy + ci % 3 - 1;
                if (x < 0 || y < 0 || !(x < OG_DIM) || !(y < OG_DIM)) continue;       // :1262
                const int c = OG_DIM * x + y;
                q = A.start[c]; q1 = A.start[c + 1];
                if (q1 - q > OG_CAP) q1 = q + OG_CAP;
                continue;
            }
            // MAG_BATCH records (one 128-byte line) per trip: the list is read once, straight out of L2, and a trip that
            // waits for one record at a time pays that latency for every 32-byte sector
            float4 f[MAG_BATCH];
#pragma unroll
            for (int u = 0; u < MAG_BATCH; u++) f[u] = A.sposf[q + u];   // past q1: another cell's records or the array's padding, masked below
            int hit = -1;
#pragma unroll
            for (int u = MAG_BATCH - 1; u >= 0; u--) {            // the FIRST candidate of the batch wins
                const int jj = __float_as_int(f[u].w);
                const float dx = x1f - f[u].x, dy = y1f - f[u].y, dz = z1f - f[u].z;
                const bool other = q + u < q1 && jj != i;                             // :1273
                seen |= other;
                const bool cand = other & (wide | (fmaf(dz, dz, fmaf(dy, dy, dx * dx)) < MAG_D2_GATE_F));
                hit = cand ? u : hit; j = cand ? jj : j;
            }
            if (hit >= 0) {
                // records of the batch in front of the hit are done; `seen` may have looked past it, which is harmless:
                // the candidate itself is a neighbour, so a neighbour has been seen either way
                found = true; qc = q + hit; q = qc + 1;
                break;
            }
            q += MAG_BATCH;
        }
        __syncwarp();                                             // everyone has its candidate (or is through): evaluate together
        if (found) evaluate(qc, j);
    }
    if (!active) return;
    if (seen) ext = V3{P::add(ext.x, 0.0), P::add(ext.y, 0.0), P::add(ext.z, 0.0)};
    A.ext_out[i] = ext.x; A.ext_out[n + i] = ext.y; A.ext_out[2 * n + i] = ext.z;
}

int magnet_alloc(Engine *e) {
    if (e->mag_rad) return TB_OK;
    size_t nm = (size_t)e->nm;
    double **scal[4] = {&e->mag_rad, &e->mag_stiff, &e->mag_max, &e->mag_scale};
    for (auto p : scal) {
        TB_TRY(dev_alloc(e, p, nm));
        TB_CUDA(e, cudaMemsetAsync(*p, 0, nm * 8, e->stream));
    }
    TB_TRY(dev_alloc(e, &e->mag_force, nm * 3));
    TB_TRY(dev_alloc(e, &e->og_pos, nm * 3));
    TB_TRY(dev_alloc(e, &e->og_posf, nm + MAG_BATCH));          // the pair kernel reads whole batches
    TB_CUDA(e, cudaMemsetAsync(e->og_posf, 0, (nm + MAG_BATCH) * sizeof(float4), e->stream));
    TB_TRY(dev_alloc(e, &e->og_cell, nm));
    TB_TRY(dev_alloc(e, &e->og_items, nm));
    TB_TRY(dev_alloc(e, &e->og_sorted, nm));
    TB_TRY(dev_alloc(e, &e->og_counter, (size_t)3 * OG_CELLS + 1));   // counter | start[+1] | cursor
    return TB_OK;
}

void magnet_free(Engine *e) {
    auto F = [&](auto *&p, size_t bytes) { dev_free(e, (void *)p, bytes); p = nullptr; };
    size_t nm = (size_t)e->nm;
    F(e->mag_rad, nm * 8); F(e->mag_stiff, nm * 8); F(e->mag_max, nm * 8); F(e->mag_scale, nm * 8);
    F(e->mag_force, nm * 24); F(e->og_pos, nm * 24); F(e->og_posf, (nm + MAG_BATCH) * 16);
    F(e->og_cell, nm * 4); F(e->og_items, nm * 4); F(e->og_sorted, nm * 4);
    F(e->og_counter, ((size_t)3 * OG_CELLS + 1) * 4);
}

// one occupancy-grid rebuild + magnet pass over the current state (capturable); 5 launches + 1 memset
int launch_magnet_pass(Engine *e, bool use_ext, bool tiled) {
    if (!e->mag_rad) return fail(e, TB_ERR_INVALID, "magnet pass without magnet fields");
    const int64_t nm = e->nm;
    const unsigned grid = (unsigned)((nm + 255) / 256);
    int32_t *counter = e->og_counter, *start = counter + OG_CELLS, *cursor = start + OG_CELLS + 1;
    PosView pos;
    pos.base = e->pos[e->cur]; pos.nm = nm;
    if (tiled) tiled_pos_view(e, &pos);
    TB_CUDA(e, cudaMemsetAsync(counter, 0, (size_t)OG_CELLS * 4, e->stream));
    og_cell_kernel<<<grid, 256, 0, e->stream>>>(nm, pos, e->mflags, e->og_cell, counter, (use_ext && e->ext) ? e->ext : nullptr, e->mag_force);
    og_scan_kernel<<<OG_CELLS / OG_SCAN_T, OG_SCAN_T, 0, e->stream>>>(counter, start, cursor);
    og_fill_kernel<<<grid, 256, 0, e->stream>>>(nm, e->og_cell, cursor, e->og_items);
    og_rank_kernel<<<grid, 256, 0, e->stream>>>(nm, pos, e->og_cell, start, e->og_items, e->og_sorted, e->og_pos, e->og_posf);
    MagnetArgs A;
    A.nm = nm; A.pos = pos; A.flags = e->mflags;
    A.rad = e->mag_rad; A.stiffness = e->mag_stiff; A.max_mag = e->mag_max; A.mag_scale = e->mag_scale;
    A.start = start; A.sorted = e->og_sorted; A.spos = e->og_pos; A.sposf = e->og_posf;
    A.ext_in = (use_ext && e->ext) ? e->ext : nullptr;
    A.ext_out = e->mag_force;
    const unsigned g2 = (unsigned)((nm + 127) / 128);
    if (e->flags & TB_MATH_STRICT) magnet_force_kernel<Strict><<<g2, 128, 0, e->stream>>>(A);
    else magnet_force_kernel<Fast><<<g2, 128, 0, e->stream>>>(A);
    TB_CUDA(e, cudaGetLastError());
    e->launches += 5;
    return TB_OK;
}

}  // namespace tb
