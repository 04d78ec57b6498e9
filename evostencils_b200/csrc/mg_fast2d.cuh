// mg_fast2d.cuh -- specialised 2D kernels for scalar real 5-point star stencils (2D Poisson).
//
// Design (B200, HBM-bound; 24 B/DoF for a sweep, 18 B/DoF for the inter-grid stages):
//   * every thread owns two adjacent columns (one aligned 16-byte double2 per row and array);
//     a CTA of NT threads covers a strip of 2*NT columns and marches down `rows` rows keeping the
//     live rows of u in registers -> each u value is read from memory once per CTA (plus the halo
//     rows of the strip segment), f once, the result written once;
//   * left/right neighbours come from the adjacent lane by warp shuffle; only the two edge lanes of
//     a warp issue extra 8-byte loads (L1/L2 hits: the neighbouring warp streams those lines);
//   * loads of the next row are issued one iteration ahead (software prefetch);
//   * strips cover the columns [0, n-1) = 2^L exactly, so there is no partially filled CTA; the
//     last Dirichlet column n-1 is handled by the thread that owns column n-3;
//   * operand order is the canonical stencil order S, W, C, E, N with separate multiply and add
//     (the translation unit is compiled with --fmad=false), so results are bit-identical to the
//     generic kernels and to the oracle.
// Applicable when (n-1) is a multiple of 2*NT (levels >= 8 for NT = 128); smaller levels are
// latency-bound and use the generic kernels.
#pragma once
#include "mg_kernels_generic.cuh"

namespace mg {

struct Star2D {
    double cs, cw, cc;     // south/north, west/east, centre coefficients
};

__device__ __forceinline__ double2 ld2(const double* p) { return *reinterpret_cast<const double2*>(p); }
__device__ __forceinline__ void st2(double* p, double2 v) { *reinterpret_cast<double2*>(p) = v; }
// "up-front" loads of the tile kernels: volatile asm keeps them ahead of the first store in program
// order (the compiler otherwise sinks them next to their use to save registers, which serialises
// the memory latency of a band); read-only data path.
#ifndef MG_PLAIN_LOADS
__device__ __forceinline__ double2 ldf2(const double* p) {
    double2 v;
    asm volatile("ld.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ double ldf1(const double* p) {
    double v;
    asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
#else
__device__ __forceinline__ double2 ldf2(const double* p) { return ld2(p); }
__device__ __forceinline__ double ldf1(const double* p) { return *p; }
#endif

// L2 prefetch of one contiguous row segment (bytes % 16 == 0, 16-byte aligned): issued by the CTA
// that runs `pf` bands ahead of the one that will load the segment, so that band finds its rows in
// L2 instead of paying the DRAM latency (each line is still fetched from DRAM once).
__device__ __forceinline__ void l2_prefetch_rows(const double* p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// (A u) at one node, canonical order S, W, C, E, N
__device__ __forceinline__ double star_at(const Star2D& s, double S, double W, double Cc, double E, double N) {
    double a = s.cs * S;
    a = a + s.cw * W;
    a = a + s.cc * Cc;
    a = a + s.cw * E;
    a = a + s.cs * N;
    return a;
}

// left neighbour of column x0 and right neighbour of column x0+1 in `row` (uc = own two values)
__device__ __forceinline__ void row_neighbours(const double* __restrict__ row, int x0, double2 uc, int lane,
                                               double& left, double& right) {
    left = __shfl_up_sync(0xffffffffu, uc.y, 1);
    right = __shfl_down_sync(0xffffffffu, uc.x, 1);
    if (lane == 0) left = x0 > 0 ? row[x0 - 1] : 0.0;
    if (lane == 31) right = row[x0 + 2];
}

enum { STAR_JACOBI = 0, STAR_RESIDUAL = 1, STAR_NORM = 2 };

__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

// MODE 0: out = u + w (f - A u) on interior nodes, out = u on the boundary   (Jacobi sweep)
// MODE 1: out = f - A u on interior nodes                                      (residual)
// MODE 2: partial[cta] = sum (f - A u)^2                                        (residual norm)
//
// One CTA = strip of 2*NT columns x band of R rows (R small, compile time).  Every load of the
// band (R+2 rows of u, R rows of f, the edge lanes' halo column) is issued before the first use,
// so a thread has 2R+2 16-byte loads in flight and the band costs one memory latency; the halo
// rows shared with the neighbouring bands are served by L2 (measured DRAM traffic = algorithmic).
template <int MODE, int NT, int R>
__global__ void __launch_bounds__(NT) k_star2d(Geom g, Star2D s, const double* __restrict__ u,
                                                const double* __restrict__ f, double* __restrict__ out, double w,
                                                double* __restrict__ partial) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // 0 <= x0 <= n-3
    const int n = g.n;
    const long long pitch = g.pitch;
    const int y0 = 1 + blockIdx.y * R;                         // rows y0 .. y0+R-1 (those <= n-2)
    const bool in0 = x0 >= 1;                                  // column x0+1 is always interior
    const bool last = x0 == n - 3;                             // also owns the Dirichlet column n-1
    const bool zin = u == nullptr;                             // u == 0 everywhere (pending zero fill): no loads of u
    const bool edge = ((lane == 0 && x0 > 0) || lane == 31) && !zin;
    const int xe = lane == 0 ? x0 - 1 : x0 + 2;
    double2 uu[R + 2], ff[R];
    double ue[R];
#pragma unroll
    for (int i = 0; i < R + 2; ++i) {
        uu[i] = make_double2(0.0, 0.0);
        if (!zin) uu[i] = ldf2(u + (long long)imin(y0 - 1 + i, n) * pitch + x0);   // row n: guard
    }
#pragma unroll
    for (int i = 0; i < R; ++i) ff[i] = ldf2(f + (long long)imin(y0 + i, n) * pitch + x0);
#pragma unroll
    for (int i = 0; i < R; ++i) {
        ue[i] = 0.0;
        if (edge) ue[i] = ldf1(u + (long long)imin(y0 + i, n) * pitch + xe);
    }
    __syncthreads();      // scheduling fence: keeps every load above ahead of the first use / store
    if (MODE == STAR_JACOBI && blockIdx.y == 0) {              // Dirichlet row 0
        st2(out + x0, uu[0]);
        if (last) out[n - 1] = zin ? 0.0 : u[n - 1];
    }
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const int y = y0 + i;
        const double2 um = uu[i], uc = uu[i + 1], un = uu[i + 2];
        double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
        double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
        if (lane == 0) left = ue[i];
        if (lane == 31) right = ue[i];
        if (y <= n - 2) {
            const long long base = (long long)y * pitch + x0;
            const double r0 = ff[i].x - star_at(s, um.x, left, uc.x, uc.y, un.x);
            const double r1 = ff[i].y - star_at(s, um.y, uc.x, uc.y, right, un.y);
            if (MODE == STAR_JACOBI) {
                double2 o;
                o.x = in0 ? uc.x + w * r0 : uc.x;
                o.y = uc.y + w * r1;
                st2(out + base, o);
                if (last) out[base + 2] = right;                   // Dirichlet column n-1
            } else if (MODE == STAR_RESIDUAL) {
                if (in0) {
                    double2 o;
                    o.x = r0; o.y = r1;
                    st2(out + base, o);
                } else {
                    out[base + 1] = r1;
                }
            } else {
                if (in0) acc = acc + r0 * r0;
                acc = acc + r1 * r1;
            }
        }
    }
    if (MODE == STAR_JACOBI && y0 + R >= n - 1) {              // this band reaches the Dirichlet row n-1
#pragma unroll
        for (int i = 1; i < R + 2; ++i)
            if (y0 - 1 + i == n - 1) st2(out + (long long)(n - 1) * pitch + x0, uu[i]);
        if (last) out[(long long)(n - 1) * pitch + n - 1] = zin ? 0.0 : u[(long long)(n - 1) * pitch + n - 1];
    }
    if (MODE == STAR_NORM) {
        __shared__ double s32[32];
        double v = warp_sum<double>(acc);
        if (lane == 0) s32[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = s32[0];
            for (int i = 1; i < NT / 32; ++i) t = t + s32[i];
            partial[blockIdx.y * gridDim.x + blockIdx.x] = t;
        }
    }
}

// ---- fused two-colour Gauss-Seidel sweep, out of place ------------------------------------------
// R_j = row j after the red half-sweep (red nodes (x+j) even updated from the old black
// neighbours); the black half-sweep of row y reads R_{y-1}, R_y, R_{y+1}.  One pass over memory.
// A red node only reads black nodes, which the red half-sweep leaves unchanged, so the red rows
// are formed in place in the registers that held the old rows.  The red values a warp needs just
// outside its 64 columns are recomputed by its edge lanes from a halo column pair they load with
// the rest of the band (rows y0-2 .. y0+R+1 of u, y0-1 .. y0+R of f: all loads up front).
__device__ __forceinline__ double red_value(const Star2D& s, double w, double S, double W, double Cc, double E,
                                            double N, double fv) {
    return Cc + w * (fv - star_at(s, S, W, Cc, E, N));
}

#ifndef MG_RBGS_MIN_CTAS
#define MG_RBGS_MIN_CTAS 1
#endif
// ZIN: u == 0 everywhere (pending zero fill, `u` is not dereferenced); a template parameter because the
// run-time flag cost 8 registers and with them the fifth resident CTA of the R = 4 kernel.
template <int NT, int R, bool ZIN>
__global__ void __launch_bounds__(NT, R <= 4 ? MG_RBGS_MIN_CTAS : 1) k_rbgs2d(Geom g, Star2D s, const double* __restrict__ u,
                                                const double* __restrict__ f, double* __restrict__ out, double w,
                                                int pf) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // even column; x0+1 odd
    const int n = g.n;
    const long long pitch = g.pitch;
    const int y0 = 1 + blockIdx.y * R;
    constexpr bool zin = ZIN;
    if (pf > 0 && threadIdx.x < 2 * R && !(zin && threadIdx.x < R)) {      // own rows of the band `pf` bands ahead
        const int row = y0 + pf * R + (threadIdx.x % R);
        if (row <= n - 1)
            l2_prefetch_rows((threadIdx.x < R ? u : f) + (long long)row * pitch + blockIdx.x * NT * 2, NT * 16);
    }
    const bool in0 = x0 >= 1;
    const bool last = x0 == n - 3;
    const bool edge = (lane == 0 && x0 > 0) || lane == 31;
    const bool edge_u = edge && !zin;
    const int xe = lane == 0 ? x0 - 2 : x0 + 2;                // halo column pair (xe, xe+1)
    double2 o[R + 4];          // row y0-2+k
    double2 fr[R + 2];         // row y0-1+k
    double2 eh[R + 2];         // halo pair of row y0-1+k (edge lanes)
    double ef[R];              // f at the halo red node of black row y0+k
#pragma unroll
    for (int k = 0; k < R + 4; ++k) {
        o[k] = make_double2(0.0, 0.0);
        if (!zin) o[k] = ldf2(u + (long long)imin(y0 - 2 + k, n) * pitch + x0);      // row -1 / n: guard rows
    }
#pragma unroll
    for (int k = 0; k < R + 2; ++k) fr[k] = ldf2(f + (long long)imin(y0 - 1 + k, n) * pitch + x0);
#pragma unroll
    for (int k = 0; k < R + 2; ++k) {
        eh[k] = make_double2(0.0, 0.0);
        if (edge_u) eh[k] = ldf2(u + (long long)imin(y0 - 1 + k, n) * pitch + xe);
    }
#pragma unroll
    for (int k = 0; k < R; ++k) {
        ef[k] = 0.0;
        if (edge) ef[k] = ldf1(f + (long long)imin(y0 + k, n) * pitch + (lane == 0 ? x0 - 1 : x0 + 2));
    }
    __syncthreads();      // scheduling fence (see k_star2d)
    // red half-sweep of rows y0-1 .. y0+R (k = 1 .. R+2), in place
#pragma unroll
    for (int k = 1; k < R + 3; ++k) {
        const int j = y0 - 2 + k;
        double left = __shfl_up_sync(0xffffffffu, o[k].y, 1);
        double right = __shfl_down_sync(0xffffffffu, o[k].x, 1);
        if (lane == 0) left = eh[k - 1].y;                     // old (x0-1, j)
        if (lane == 31) right = eh[k - 1].x;                   // old (x0+2, j)
        if (j >= 1 && j <= n - 2) {
            if ((j & 1) == 0) {                                 // column x0 (even) is red on even rows
                if (in0) o[k].x = red_value(s, w, o[k - 1].x, left, o[k].x, o[k].y, o[k + 1].x, fr[k - 1].x);
            } else {                                            // column x0+1 (odd) is red on odd rows
                o[k].y = red_value(s, w, o[k - 1].y, o[k].x, o[k].y, right, o[k + 1].y, fr[k - 1].y);
            }
        }
    }
    if (blockIdx.y == 0) {                                      // Dirichlet row 0 (k = 1)
        st2(out + x0, o[1]);
        if (last) out[n - 1] = zin ? 0.0 : u[n - 1];
    }
    // black half-sweep of rows y0 .. y0+R-1 (k = 2 .. R+1)
#pragma unroll
    for (int i = 0; i < R; ++i) {
        const int k = i + 2;
        const int y = y0 + i;
        double left = __shfl_up_sync(0xffffffffu, o[k].y, 1);
        double right = __shfl_down_sync(0xffffffffu, o[k].x, 1);
        if (y <= n - 2) {
            if (y & 1) {               // black node: column x0 on odd rows -> needs the red value at x0-1
                if (lane == 0)
                    left = x0 > 0 ? red_value(s, w, eh[i].y, eh[i + 1].x, eh[i + 1].y, o[k].x, eh[i + 2].y, ef[i]) : 0.0;
            } else {                   // black node: column x0+1 on even rows -> needs the red value at x0+2
                if (lane == 31)
                    right = x0 + 2 > n - 2 ? eh[i + 1].x
                                           : red_value(s, w, eh[i].x, o[k].y, eh[i + 1].x, eh[i + 1].y, eh[i + 2].x, ef[i]);
            }
            double2 v = o[k];
            if (y & 1) {
                if (in0) v.x = o[k].x + w * (fr[i + 1].x - star_at(s, o[k - 1].x, left, o[k].x, o[k].y, o[k + 1].x));
            } else {
                v.y = o[k].y + w * (fr[i + 1].y - star_at(s, o[k - 1].y, o[k].x, o[k].y, right, o[k + 1].y));
            }
            const long long base = (long long)y * pitch + x0;
            st2(out + base, v);
            if (last) out[base + 2] = eh[i + 1].x;                // Dirichlet column n-1 (old value)
        }
    }
    if (y0 + R >= n - 1) {                                      // this band reaches the Dirichlet row n-1
#pragma unroll
        for (int k = 2; k < R + 3; ++k)
            if (y0 - 2 + k == n - 1) st2(out + (long long)(n - 1) * pitch + x0, o[k]);
        if (last) out[(long long)(n - 1) * pitch + n - 1] = zin ? 0.0 : u[(long long)(n - 1) * pitch + n - 1];
    }
}

// ---- block (tile) Jacobi sweep, out of place -------------------------------------------------------
// `solve locally ... with jacobi` over non-overlapping px x py tiles anchored at node 0
// (stencils/multiple.py:204-217, exastencils.py:769-822): u_s += sum_t W[s][t] r_t over the interior
// nodes t of the tile of s, r = f - A u_old, W = omega * inv(L_tile) per boundary class of the tile.
// One CTA = strip of cw columns x band of rb rows, both whole numbers of tiles (cw % px == 0,
// rb % py == 0, anchored at 0), so every tile lives in one CTA: node-parallel and coalesced
// (thread = 2 adjacent columns), residuals exchanged through shared memory.  All loads up front.
// Arithmetic identical to k_smooth_tile (ascending slot order, first term initialises the sum).
struct Tile2D {
    int px, py, m, cw, rb, full_index, pf;
};

// PX, PY are template parameters: slot positions, the row's position in its tile and the W row of
// a complete tile are compile-time, only tiles cut by the Dirichlet boundary take the masked path
// (with run-time tile shapes the index arithmetic made the sweep instruction-bound: 0.52 of the copy
// peak at 4096^2).
template <int NT, int PX, int PY>
__global__ void __launch_bounds__(NT) k_tile2d(Geom g, Star2D s, Tile2D t, const SmootherTable<double>* __restrict__ st,
                                                const double* __restrict__ wtab, const double* __restrict__ u,
                                                const double* __restrict__ f, double* __restrict__ out) {
    pdl_sync();
    constexpr int M = PX * PY;
    constexpr int RB = PY == 3 ? 3 : 4;                        // rows per band: a whole number of tiles
    constexpr int CW = PX == 3 ? (2 * NT / 6) * 6 : 2 * NT;    // columns per strip: an even multiple of PX
    __shared__ __align__(16) double r_s[RB][2 * NT];
    __shared__ double w_s[M * M];
    const int lane = threadIdx.x & 31;
    const int n = g.n;
    const long long pitch = g.pitch;
    const int c0 = 2 * threadIdx.x;                            // column inside the strip
    const int x0 = blockIdx.x * CW + c0;
    const int xl = imin(x0, n - 1);                            // load column (even, in bounds)
    const int yb = blockIdx.y * RB;                            // first row of the band (row 0 included)
    const bool active = c0 < CW && x0 <= n - 2;
    const int xe = lane == 0 ? imax(xl - 1, 0) : imin(xl + 2, n - 1);
    if (threadIdx.x < M * M) w_s[threadIdx.x] = wtab[(long long)t.full_index * M * M + threadIdx.x];
    if (t.pf > 0 && threadIdx.x >= 32 && threadIdx.x < 32 + 2 * RB) {      // L2 prefetch, see k_rbgs2d
        const int k = threadIdx.x - 32;
        const int row = yb + t.pf * RB + (k % RB);
        const long long col = (long long)blockIdx.x * CW;
        if (row <= n - 1 && col + 2 * NT <= pitch) l2_prefetch_rows((k < RB ? u : f) + (long long)row * pitch + col, NT * 16);
    }
    double2 uu[RB + 2], ff[RB];
    double ue[RB];
#pragma unroll
    for (int i = 0; i < RB + 2; ++i) uu[i] = ldf2(u + (long long)imin(yb - 1 + i, n) * pitch + xl);   // row -1 / n: guard rows
#pragma unroll
    for (int i = 0; i < RB; ++i) ff[i] = ldf2(f + (long long)imin(yb + i, n) * pitch + xl);
#pragma unroll
    for (int i = 0; i < RB; ++i) {
        ue[i] = 0.0;
        if (lane == 0 || lane == 31) ue[i] = ldf1(u + (long long)imin(yb + i, n) * pitch + xe);
    }
    __syncthreads();      // scheduling fence (see k_star2d); also publishes w_s
    const bool in0 = x0 >= 1 && x0 <= n - 2, in1 = x0 + 1 <= n - 2;
#pragma unroll
    for (int i = 0; i < RB; ++i) {
        const int y = yb + i;
        const double2 um = uu[i], uc = uu[i + 1], un = uu[i + 2];
        double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
        double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
        if (lane == 0) left = ue[i];
        if (lane == 31) right = ue[i];
        const bool rowin = y >= 1 && y <= n - 2;
        double r0 = 0.0, r1 = 0.0;
        if (rowin && in0) r0 = ff[i].x - star_at(s, um.x, left, uc.x, uc.y, un.x);
        if (rowin && in1) r1 = ff[i].y - star_at(s, um.y, uc.x, uc.y, right, un.y);
        *reinterpret_cast<double2*>(&r_s[i][c0]) = make_double2(r0, r1);      // one 16-byte store: conflict-free
    }
    __syncthreads();
    if (!active) return;
    constexpr int FULL = (1 << M) - 1;
    const int sx0 = c0 % PX, sx1 = (c0 + 1) % PX;              // strips start at a multiple of PX
#pragma unroll
    for (int i = 0; i < RB; ++i) {
        const int y = yb + i;
        if (y > n - 1) continue;
        const int sy = i % PY;                                  // bands start at a multiple of PY; i is unrolled
        double2 o = uu[i + 1];
        if (y >= 1 && y <= n - 2) {
            const int by = y - sy;
            const bool rows_full = by >= 1 && by + PY - 1 <= n - 2;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int x = x0 + c;
                if (x < 1 || x > n - 2) continue;
                const int sx = c ? sx1 : sx0;
                const int bx = x - sx;                          // tile origin (global)
                const int slot = sx * PY + sy;
                const double* rbase = &r_s[i - sy][c0 + c - sx];
                double delta = 0.0;
                if (rows_full && bx >= 1 && bx + PX - 1 <= n - 2) {
                    const double* wrow = w_s + slot * M;
#pragma unroll
                    for (int q = 0; q < M; ++q) {               // slot q at (q / PY, q % PY), x outer
                        const double term = wrow[q] * rbase[(q % PY) * 2 * NT + q / PY];
                        delta = q == 0 ? term : delta + term;
                    }
                } else {
                    int mask = 0;
#pragma unroll
                    for (int q = 0; q < M; ++q) {
                        const int qx = bx + q / PY, qy = by + q % PY;
                        if (qx >= 1 && qx <= n - 2 && qy >= 1 && qy <= n - 2) mask |= 1 << q;
                    }
                    const double* wrow = mask == FULL ? w_s + slot * M
                                                      : wtab + ((long long)st->mask_index[mask] * M + slot) * M;
                    bool first = true;
#pragma unroll
                    for (int q = 0; q < M; ++q) {
                        if (!((mask >> q) & 1)) continue;
                        const double term = wrow[q] * rbase[(q % PY) * 2 * NT + q / PY];
                        delta = first ? term : delta + term;
                        first = false;
                    }
                }
                if (c) o.y = o.y + delta; else o.x = o.x + delta;
            }
        }
        double* dst = out + (long long)y * pitch + x0;
        st2(dst, o);
        if (x0 == n - 3) dst[2] = u[(long long)y * pitch + n - 1];  // Dirichlet column n-1
    }
}

// ---- fused residual + full-weighting restriction ----------------------------------------------------
// Thread owns coarse columns X0, X0+1 (one double2 store per coarse row) = fine columns 2X0 .. 2X0+3;
// the fine residual at column 2X0-1 comes from the left lane (its column 2X0'+3), lane 0 recomputes it.
// The CTA marches over fine rows; every fine residual row is computed once and feeds the coarse rows
// above and below.  Accumulation order per coarse node: fine rows 2Y-1, 2Y, 2Y+1, columns -1, 0, +1
// (memory order), term = (w_x * w_y) * r  -- identical to the generic kernel.
template <int NT>
__global__ void __launch_bounds__(NT) k_resrestrict2d(Geom gc, Geom gf, Star2D s, const double* __restrict__ u,
                                                       const double* __restrict__ f, double* __restrict__ coarse,
                                                       int rows) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int X0 = (blockIdx.x * NT + threadIdx.x) * 2;        // coarse columns X0, X0+1 ; X0 <= nc-3
    const int x0 = 2 * X0;                                      // fine columns x0 .. x0+3
    const int nc = gc.n, nf = gf.n;
    const long long pf = gf.pitch, pc = gc.pitch;
    const int Y_begin = 1 + blockIdx.y * rows;
    int Y_end = Y_begin + rows;
    if (Y_end > nc - 1) Y_end = nc - 1;
    const bool in0 = X0 >= 1;                                   // coarse column X0 interior; X0+1 always is

    // residuals of fine row y at columns x0-1 .. x0+3 (r[0] = column x0-1)
    struct Row { double2 a, b; };
    auto load_row = [&](int y) -> Row {
        Row r;
        const double* q = u + (long long)y * pf + x0;
        r.a = ld2(q); r.b = ld2(q + 2);
        return r;
    };
    auto residual_row = [&](int y, const Row& m, const Row& c, const Row& n_, double r[5]) {
        const double* frow = f + (long long)y * pf + x0;
        double2 fa = ld2(frow), fb = ld2(frow + 2);
        const double* urow = u + (long long)y * pf;
        // own columns x0 .. x0+3
        double wl = __shfl_up_sync(0xffffffffu, c.b.y, 1);      // u at x0-1
        double er = __shfl_down_sync(0xffffffffu, c.a.x, 1);    // u at x0+4
        if (lane == 0) wl = x0 > 0 ? urow[x0 - 1] : 0.0;
        if (lane == 31) er = urow[x0 + 4];
        r[1] = fa.x - star_at(s, m.a.x, wl, c.a.x, c.a.y, n_.a.x);
        r[2] = fa.y - star_at(s, m.a.y, c.a.x, c.a.y, c.b.x, n_.a.y);
        r[3] = fb.x - star_at(s, m.b.x, c.a.y, c.b.x, c.b.y, n_.b.x);
        r[4] = fb.y - star_at(s, m.b.y, c.b.x, c.b.y, er, n_.b.y);
        double rl = __shfl_up_sync(0xffffffffu, r[4], 1);       // residual at x0-1 from the left lane
        if (lane == 0) {
            rl = 0.0;
            if (x0 > 0) {
                const double* c1 = urow + x0 - 1;
                rl = f[(long long)y * pf + x0 - 1] - star_at(s, c1[-pf], c1[-1], c1[0], c.a.x, c1[pf]);
            }
        }
        r[0] = rl;
    };

    int y = 2 * Y_begin - 1;                                    // first fine row needed
    Row um = load_row(y - 1), uc = load_row(y), un = load_row(y + 1);
    double rm[5];
    residual_row(y, um, uc, un, rm);                            // fine row 2Y-1
    for (int Y = Y_begin; Y < Y_end; ++Y) {
        // fine rows 2Y and 2Y+1
        Row u2 = load_row(2 * Y + 1), u3 = load_row(2 * Y + 2);
        double rc[5], rn[5];
        residual_row(2 * Y, uc, un, u2, rc);
        residual_row(2 * Y + 1, un, u2, u3, rn);
        double2 o;
        {
            double a = 0.0625 * rm[0];
            a = a + 0.125 * rm[1];
            a = a + 0.0625 * rm[2];
            a = a + 0.125 * rc[0];
            a = a + 0.25 * rc[1];
            a = a + 0.125 * rc[2];
            a = a + 0.0625 * rn[0];
            a = a + 0.125 * rn[1];
            a = a + 0.0625 * rn[2];
            o.x = a;
            double b = 0.0625 * rm[2];
            b = b + 0.125 * rm[3];
            b = b + 0.0625 * rm[4];
            b = b + 0.125 * rc[2];
            b = b + 0.25 * rc[3];
            b = b + 0.125 * rc[4];
            b = b + 0.0625 * rn[2];
            b = b + 0.125 * rn[3];
            b = b + 0.0625 * rn[4];
            o.y = b;
        }
        double* dst = coarse + (long long)Y * pc + X0;
        if (in0) st2(dst, o); else dst[1] = o.y;
#pragma unroll
        for (int k = 0; k < 5; ++k) rm[k] = rn[k];
        uc = u2; un = u3;
    }
    (void)nf;
}

// ---- fused prolongation + correction: u += omega * (P e) --------------------------------------------
// Thread owns fine columns x0 .. x0+3 (x0 = 4 t) <- coarse columns X0 = x0/2, X0+1 and, for the odd fine
// column x0+3, X0+2 (from the right lane; lane 31 loads it).  The CTA marches over coarse rows: coarse
// row Y and Y+1 give the fine rows 2Y (even) and 2Y+1 (odd).  In place (each fine node is touched by
// exactly one thread).
template <int NT>
__global__ void __launch_bounds__(NT) k_prolong2d(Geom gf, Geom gc, const double* __restrict__ e,
                                                   double* __restrict__ u, double omega, int rows) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 4;        // fine columns x0..x0+3 ; x0 <= nf-5
    const int X0 = x0 >> 1;
    const int nf = gf.n, nc = gc.n;
    const long long pf = gf.pitch, pc = gc.pitch;
    const int Y_begin = blockIdx.y * rows;                      // coarse rows [Y_begin, Y_end)
    int Y_end = Y_begin + rows;
    if (Y_end > nc - 1) Y_end = nc - 1;                         // coarse row nc-1 only serves fine row nf-2
    auto load_c = [&](int Y, double& c0, double& c1, double& c2) {
        const double* row = e + (long long)Y * pc;
        double2 v = ld2(row + X0);
        c0 = v.x; c1 = v.y;
        c2 = __shfl_down_sync(0xffffffffu, v.x, 1);
        if (lane == 31) c2 = row[X0 + 2];
    };
    auto update = [&](int y, double p0, double p1, double p2, double p3) {
        double* q = u + (long long)y * pf + x0;
        double2 a = ld2(q), b = ld2(q + 2);
        if (x0 >= 1) a.x = a.x + omega * p0;                    // column 0 is Dirichlet
        a.y = a.y + omega * p1;
        b.x = b.x + omega * p2;
        b.y = b.y + omega * p3;
        st2(q, a); st2(q + 2, b);
    };
    double a0, a1, a2;
    load_c(Y_begin, a0, a1, a2);
    for (int Y = Y_begin; Y < Y_end; ++Y) {
        double b0, b1, b2;
        load_c(Y + 1, b0, b1, b2);
        if (Y >= 1)                                             // even fine row 2Y (row 0 is Dirichlet)
            update(2 * Y, a0, 0.5 * (a0 + a1), a1, 0.5 * (a1 + a2));
        // odd fine row 2Y+1 <= nf-2
        update(2 * Y + 1, 0.5 * (a0 + b0), 0.25 * (((a0 + a1) + b0) + b1), 0.5 * (a1 + b1),
               0.25 * (((a1 + a2) + b1) + b2));
        a0 = b0; a1 = b1; a2 = b2;
    }
    (void)nf;
}

// ---- tile versions of the two inter-grid stages (all loads of a band issued up front) ----------------
//
// Fused residual + full weighting.  A thread owns the fine columns x0 = 2X, x0+1 of coarse column X
// and a band of RC coarse rows: it forms the fine residuals at (x0, x0+1) of the 2RC+1 fine rows
// the band needs and takes the residual at x0-1 from the lane to its left.  Warps overlap by one
// lane (lane 0 of every warp only donates, a warp stores 31 coarse columns), so that no lane has
// to recompute a neighbour's residual from extra halo loads; the overlap costs 1/32 more L2 reads.
// Accumulation order per coarse node: fine rows 2Y-1, 2Y, 2Y+1, columns -1, 0, +1 (memory order),
// term = (w_x * w_y) * r -- identical to the generic kernel.
template <int NT, int RC>
__global__ void __launch_bounds__(NT) k_resrestrict2d_t(Geom gc, Geom gf, Star2D s, const double* __restrict__ u,
                                                         const double* __restrict__ f, double* __restrict__ coarse,
                                                         int ahead) {
    pdl_sync();
    constexpr int NR = 2 * RC + 1, NU = NR + 2;
    if (ahead > 0 && threadIdx.x < 4 * RC) {                   // the 2RC fine rows 2Y, 2Y+1 of the band `ahead` bands on
        const int row = 2 * (1 + (blockIdx.y + ahead) * RC) + (threadIdx.x % (2 * RC));
        const long long col = (long long)blockIdx.x * (NT / 32) * 62;      // first fine column of the CTA (16-byte aligned)
        if (row <= gf.n - 1 && col + NT * 2 <= gf.pitch)
            l2_prefetch_rows((threadIdx.x < 2 * RC ? u : f) + (long long)row * gf.pitch + col, NT * 16);
    }
    const int lane = threadIdx.x & 31;
    const int nc = gc.n, nf = gf.n;
    const long long pf = gf.pitch, pc = gc.pitch;
    const int X = (blockIdx.x * (NT / 32) + (threadIdx.x >> 5)) * 31 + lane;    // coarse column
    const int x0 = 2 * imin(X, nc - 1);                                          // <= nf-1 (x0+1 < pitch)
    const bool store = lane >= 1 && X <= nc - 2;
    const int Y0 = 1 + blockIdx.y * RC;                                          // coarse rows Y0 .. Y0+RC-1
    const int yr0 = 2 * Y0 - 1;                                                  // first fine residual row
    double2 uu[NU], ff[NR];
    double er[NR];
#pragma unroll
    for (int i = 0; i < NU; ++i) uu[i] = ldf2(u + (long long)imin(yr0 - 1 + i, nf) * pf + x0);    // row nf: guard
#pragma unroll
    for (int j = 0; j < NR; ++j) ff[j] = ldf2(f + (long long)imin(yr0 + j, nf) * pf + x0);
#pragma unroll
    for (int j = 0; j < NR; ++j) {
        er[j] = 0.0;
        if (lane == 31) er[j] = ldf1(u + (long long)imin(yr0 + j, nf) * pf + x0 + 2);
    }
    __syncthreads();      // scheduling fence (see k_star2d)
    double rl[NR], r0[NR], r1[NR];
#pragma unroll
    for (int j = 0; j < NR; ++j) {
        const double2 um = uu[j], uc = uu[j + 1], un = uu[j + 2];
        const double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
        double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
        if (lane == 31) right = er[j];
        r0[j] = ff[j].x - star_at(s, um.x, left, uc.x, uc.y, un.x);
        r1[j] = ff[j].y - star_at(s, um.y, uc.x, uc.y, right, un.y);
        rl[j] = __shfl_up_sync(0xffffffffu, r1[j], 1);
    }
#pragma unroll
    for (int i = 0; i < RC; ++i) {
        const int Y = Y0 + i;
        if (store && Y <= nc - 2) {
            const int m = 2 * i, c = 2 * i + 1, n_ = 2 * i + 2;
            double a = 0.0625 * rl[m];
            a = a + 0.125 * r0[m];
            a = a + 0.0625 * r1[m];
            a = a + 0.125 * rl[c];
            a = a + 0.25 * r0[c];
            a = a + 0.125 * r1[c];
            a = a + 0.0625 * rl[n_];
            a = a + 0.125 * r0[n_];
            a = a + 0.0625 * r1[n_];
            coarse[(long long)Y * pc + X] = a;
        }
    }
}

// Fused prolongation + correction, u += omega * (P e), in place.  A thread owns the fine columns
// x0 = 2X, x0+1 and the fine rows 2Y, 2Y+1 of RC coarse rows Y; e(., X+1) comes from the lane to
// the right (lane 31 loads it).
template <int NT, int RC>
__global__ void __launch_bounds__(NT) k_prolong2d_t(Geom gf, Geom gc, const double* __restrict__ e,
                                                     double* __restrict__ u, double omega) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;        // <= nf-3
    const int X = x0 >> 1;
    const int nf = gf.n, nc = gc.n;
    const long long pf = gf.pitch, pc = gc.pitch;
    const int Y0 = blockIdx.y * RC;                             // coarse rows Y0 .. Y0+RC-1 (<= nc-2)
    double c0[RC + 1], c1[RC + 1];
    double2 uu[2 * RC];
#pragma unroll
    for (int i = 0; i <= RC; ++i) c0[i] = ldf1(e + (long long)imin(Y0 + i, nc - 1) * pc + X);
#pragma unroll
    for (int i = 0; i <= RC; ++i) {
        c1[i] = 0.0;
        if (lane == 31) c1[i] = ldf1(e + (long long)imin(Y0 + i, nc - 1) * pc + X + 1);
    }
#pragma unroll
    for (int k = 0; k < 2 * RC; ++k) uu[k] = ldf2(u + (long long)imin(2 * Y0 + k, nf) * pf + x0);
    __syncthreads();      // scheduling fence
#pragma unroll
    for (int i = 0; i <= RC; ++i) {
        const double r = __shfl_down_sync(0xffffffffu, c0[i], 1);
        if (lane != 31) c1[i] = r;
    }
    const bool in0 = x0 >= 1;                                   // column 0 is Dirichlet
#pragma unroll
    for (int i = 0; i < RC; ++i) {
        const int Y = Y0 + i;
        if (Y <= nc - 2) {
            const double a0 = c0[i], a1 = c1[i], b0 = c0[i + 1], b1 = c1[i + 1];
            if (Y >= 1) {                                       // even fine row 2Y (row 0 is Dirichlet)
                double2 v = uu[2 * i];
                if (in0) v.x = v.x + omega * a0;
                v.y = v.y + omega * (0.5 * (a0 + a1));
                st2(u + (long long)(2 * Y) * pf + x0, v);
            }
            double2 v = uu[2 * i + 1];                          // odd fine row 2Y+1 <= nf-2
            if (in0) v.x = v.x + omega * (0.5 * (a0 + b0));
            v.y = v.y + omega * (0.25 * (((a0 + a1) + b0) + b1));
            st2(u + (long long)(2 * Y + 1) * pf + x0, v);
        }
    }
}

}  // namespace mg
