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
//     og_scan_kernel   exclusive scan of the 65,536 populations (one CTA)
//     og_fill_kernel   masses into their cell's segment (arbitrary order)
//     og_rank_kernel   ... brought into ascending mass index (rank inside the segment)
//     magnet_force_kernel   per free mass: the 3 x 3 cells in the reference's (x offset, y offset) order, the
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

__global__ void og_cell_kernel(int64_t nm, const PosView pos, const uint8_t *flags, int32_t *cell, int32_t *counter) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = -1;
    if (flags[i] & F_VALID) {
        c = OG_DIM * og_coord(pos.ld(i, 0)) + og_coord(pos.ld(i, 1));   // getOgCounterIdx, :835-837
        atomicAdd(&counter[c], 1);
    }
    cell[i] = c;
}

// one CTA of 1024 threads, 64 cells each: start[c] = sum of counter[0..c), start[OG_CELLS] = total; cursor = start
__global__ void __launch_bounds__(1024) og_scan_kernel(const int32_t *counter, int32_t *start, int32_t *cursor) {
    __shared__ int32_t part[1024];
    const int per = OG_CELLS / 1024;
    int t = threadIdx.x, base = t * per, sum = 0;
    for (int q = 0; q < per; q++) sum += counter[base + q];
    part[t] = sum;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        int v = (t >= off) ? part[t - off] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int run = part[t] - sum;
    for (int q = 0; q < per; q++) {
        start[base + q] = run; cursor[base + q] = run;
        run += counter[base + q];
    }
    if (t == 1023) start[OG_CELLS] = run;
}

__global__ void og_fill_kernel(int64_t nm, const int32_t *cell, int32_t *cursor, int32_t *items) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = cell[i];
    if (c >= 0) items[atomicAdd(&cursor[c], 1)] = (int32_t)i;
}

__global__ void og_rank_kernel(int64_t nm, const int32_t *cell, const int32_t *start, const int32_t *items, int32_t *sorted) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    int c = cell[i];
    if (c < 0) return;
    int b = start[c], e = start[c + 1], rank = 0;
    for (int q = b; q < e; q++) rank += items[q] < (int32_t)i;
    sorted[b + rank] = (int32_t)i;
}

struct MagnetArgs {
    int64_t nm;
    PosView pos;                // start-of-step positions
    const uint8_t *flags;
    const double *rad, *stiffness, *max_mag, *mag_scale;
    const int32_t *start, *sorted;
    const double *ext_in;       // uploaded one-shot external force or null (= 0)
    double *ext_out;            // [3][nm]
};

template <class P> __global__ void __launch_bounds__(128) magnet_force_kernel(MagnetArgs A) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.nm) return;
    const int64_t n = A.nm;
    V3 ext = {0.0, 0.0, 0.0};
    if (A.ext_in) ext = V3{A.ext_in[i], A.ext_in[n + i], A.ext_in[2 * n + i]};
    uint8_t fl = A.flags[i];
    if ((fl & F_VALID) && !(fl & F_FIXED)) {                      // fixed masses return before the pass, :1292-1298
        V3 p1 = {A.pos.ld(i, 0), A.pos.ld(i, 1), A.pos.ld(i, 2)};
        double rad1 = A.rad[i], st1 = A.stiffness[i], mm1 = A.max_mag[i];
        int cx = og_coord(p1.x), cy = og_coord(p1.y);
#pragma unroll 1
        for (int a = -1; a <= 1; a++) {
#pragma unroll 1
            for (int b = -1; b <= 1; b++) {
                int x = cx + a, y = cy + b;
                if (x < 0 || y < 0 || !(x < OG_DIM) || !(y < OG_DIM)) continue;   // :1262
                int c = OG_DIM * x + y;
                int q0 = A.start[c], q1 = A.start[c + 1];
                if (q1 - q0 > OG_CAP) q1 = q0 + OG_CAP;
#pragma unroll 1
                for (int q = q0; q < q1; q++) {
                    int j = A.sorted[q];
                    if (j == (int)i) continue;                                     // :1273
                    V3 temp = sub3<P>(p1, V3{A.pos.ld(j, 0), A.pos.ld(j, 1), A.pos.ld(j, 2)});
                    double tn = norm3<P>(temp);
                    double inter = P::sub(tn, P::add(rad1, A.rad[j]));
                    V3 ef = {0.0, 0.0, 0.0};
                    if (tn < MAG_CUTOFF) {
                        V3 dir = div3<P>(temp, tn);
                        if (inter < 0.0) acc3<P>(ef, scale3<P>(dir, P::mul(fabs(inter), st1)));   // shells intersect, :1230-1232
                        double tn2 = P::mul(tn, tn);
                        double c2 = P::div(P::mul(A.mag_scale[j], mm1), tn2 > 1e-12 ? tn2 : 1e-12);
                        dec3<P>(ef, scale3<P>(dir, c2));                                          // :1235-1237
                    }
                    acc3<P>(ext, ef);                                                             // :1240
                }
            }
        }
    }
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
    F(e->mag_force, nm * 24);
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
    og_cell_kernel<<<grid, 256, 0, e->stream>>>(nm, pos, e->mflags, e->og_cell, counter);
    og_scan_kernel<<<1, 1024, 0, e->stream>>>(counter, start, cursor);
    og_fill_kernel<<<grid, 256, 0, e->stream>>>(nm, e->og_cell, cursor, e->og_items);
    og_rank_kernel<<<grid, 256, 0, e->stream>>>(nm, e->og_cell, start, e->og_items, e->og_sorted);
    MagnetArgs A;
    A.nm = nm; A.pos = pos; A.flags = e->mflags;
    A.rad = e->mag_rad; A.stiffness = e->mag_stiff; A.max_mag = e->mag_max; A.mag_scale = e->mag_scale;
    A.start = start; A.sorted = e->og_sorted;
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
