// mg_restrict3d.cuh -- fused residual + full-weighting restriction for the 3D 7-point star (3D Poisson).
//
// coarse(X,Y,Z) = sum over (oz, oy, ox) in memory order of w_oz w_oy w_ox r(2X+ox, 2Y+oy, 2Z+oz),
// r = f - A u on the fine grid; the fine residual is never written to global memory (17 B per fine
// DoF instead of 24 + 9).  Same structure as k_rbgs3d (mg_rbgs3d.cuh):
//   * a CTA owns 30 coarse columns x RYC coarse rows and marches over the fine planes of a chunk of
//     coarse planes; plane tiles of u (2 RYC + 3 rows x 64 columns) and f (2 RYC + 1 rows) are staged by
//     tensor-map TMA into shared-memory rings, one box per plane, completion on the plane's mbarrier;
//   * warp k owns fine row yb-1+k and carries its column pairs of the planes z-1, z, z+1 in registers;
//     per fine plane it forms the residual of its row (both columns of every lane's pair) and puts it
//     into a double-buffered residual plane in shared memory;
//   * the warps of the even fine rows (= coarse rows) add the 9 terms of the plane to the running sum of
//     their coarse node -- sequentially in (oz, oy, ox) order, i.e. bit-identical to the generic
//     kernel and the oracle: an odd fine plane finishes the coarse plane below (oz = +1) and starts
//     the one above (oz = -1), an even plane adds the centre terms; left neighbours come by shuffle.
#pragma once
#include "mg_rbgs3d.cuh"

namespace mg {

constexpr int RS3_TXC = 30;     // coarse columns per tile (fine columns xb-1 .. xb+59, staged xb-2 .. xb+61)

template <int RYC, int NS, int NF>
constexpr size_t resrestrict3d_smem_bytes() {
    return (size_t)(NS * (2 * RYC + 3) + (NF + 2) * (2 * RYC + 1)) * RB3_W * sizeof(double) + 16 * sizeof(unsigned long long) + 128;
}

// nine terms of one fine plane, rows y-1, y, y+1 x columns x-1, x, x+1, in memory order
__device__ __forceinline__ double rs3_plane_terms(double acc, bool first, double wz, double lm, const double2 ym, double l0,
                                                  const double2 y0, double lp, const double2 yp) {
    const double w_c = wz * 0.0625, w_e = wz * 0.125, w_m = wz * 0.25;      // wz * (w_ox * w_oy): exact powers of two
    double t = w_c * lm;
    acc = first ? t : acc + t;
    acc = acc + w_e * ym.x;
    acc = acc + w_c * ym.y;
    acc = acc + w_e * l0;
    acc = acc + w_m * y0.x;
    acc = acc + w_e * y0.y;
    acc = acc + w_c * lp;
    acc = acc + w_e * yp.x;
    acc = acc + w_c * yp.y;
    return acc;
}

// grid = (ceil((nc-2) / 30), ceil((nc-2) / RYC), ceil(planes / zchunk)), block = 32 * (2 RYC + 1).
// z-slabs: the tensor maps describe the fine slab array including its ghost planes, their plane 0 being the
// global fine plane `zoff_f`; `coarse` points at the global coarse plane `zoff_c` of the coarse slab; the
// coarse planes Zlo .. Zhi (global, the owned ones) are written.  Whole grids: offsets 0, Zlo = 1, Zhi = nc-2.
template <int RYC, int NS, int NF>
__global__ void __launch_bounds__(32 * (2 * RYC + 1)) k_resrestrict3d(Geom gc, Geom gf, Star3D st,
                                                                     const __grid_constant__ CUtensorMap tm_u,
                                                                     const __grid_constant__ CUtensorMap tm_f,
                                                                     double* __restrict__ coarse, int zchunk,
                                                                     int zoff_f, int zoff_c, int Zlo, int Zhi) {
    constexpr int RF = 2 * RYC, UR = RF + 3, FR = RF + 1, NW = RF + 1;
    constexpr uint32_t U_BYTES = UR * RB3_W * 8, F_BYTES = FR * RB3_W * 8;
    constexpr int U_SLOT = UR * RB3_W, F_SLOT = FR * RB3_W;
    extern __shared__ __align__(128) unsigned char rs3_smem[];
    double* su = reinterpret_cast<double*>(rs3_smem + ((128u - (smem_u32(rs3_smem) & 127u)) & 127u));   // [NS][UR][64]
    double* sf = su + NS * U_SLOT;                                          // [NF][FR][64]
    double* sr = sf + NF * F_SLOT;                                          // [2][FR][64] residual planes
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sr + 2 * F_SLOT);     // [NS] u, [NF] f
    pdl_sync();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nc = gc.n;
    const int X0 = 1 + blockIdx.x * RS3_TXC, Y0 = 1 + blockIdx.y * RYC, Zb = Zlo + blockIdx.z * zchunk;
    const int Ze = imin(Zb + zchunk, Zhi + 1);                              // coarse planes Zb .. Ze-1 (global indices)
    const int xb = 2 * X0, yb = 2 * Y0;
    const int z0 = 2 * Zb - 1, z1 = 2 * Ze - 1;                             // fine planes of the march
    const int p0 = z0 - 1;                                                  // first staged plane of u (f: z0)
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    auto stage_u = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * slot;
            mbar_arrive_expect_tx(bar, U_BYTES);
            tma_load_3d(su_a + (uint32_t)(slot * U_SLOT * 8), &tm_u, xb - 2, yb - 2, p - zoff_f, bar);
        }
    };
    auto stage_f = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * (NS + slot);
            mbar_arrive_expect_tx(bar, F_BYTES);
            tma_load_3d(sf_a + (uint32_t)(slot * F_SLOT * 8), &tm_f, xb - 2, yb - 1, p - zoff_f, bar);
        }
    };

    if (threadIdx.x == 0) {
        for (int i = 0; i < NS + NF; ++i) mbar_init(bar_a + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == NW - 1) {
        for (int i = 0; i < NS; ++i)
            if (p0 + i <= z1 + 1) stage_u(p0 + i, i);
        for (int i = 0; i < NF; ++i)
            if (z0 + i <= z1) stage_f(z0 + i, i);
    }
    mbar_wait(bar_a, 0);            // planes p0 and p0+1 of u
    mbar_wait(bar_a + 8, 0);

    const int o_row = (warp + 1) * RB3_W + 2 * lane;                         // the warp's row inside a slot of u
    const int o_f = warp * RB3_W + 2 * lane;                                 // ... of f and of the residual planes
    const bool coarse_warp = (warp & 1) != 0;                                // fine row yb-1+warp is even
    const int Y = Y0 + (warp >> 1), X = X0 + lane - 1;
    const bool store_ok = coarse_warp && lane >= 1 && lane <= RS3_TXC && X <= nc - 2 && Y <= nc - 2;
    double* cptr = coarse + (long long)Y * gc.pitch + X;

    double2 rzm = *reinterpret_cast<const double2*>(su + o_row);            // plane z0-1 (slot 0)
    double2 rc = *reinterpret_cast<const double2*>(su + U_SLOT + o_row);    // plane z0 (slot 1)
    double acc = 0.0;

    // ring state of step z = z0: slots of planes z, z+1 of u and of plane z of f
    int e_c = U_SLOT, e_p = 2 * U_SLOT, e_f = 0;
    int u_slot = 0, f_slot = 0;                                             // slot numbers of plane z-1 of u / plane z of f
    uint32_t b_u = bar_a + 16, b_f = bar_a + 8 * NS;
    uint32_t par_u = 0, par_f = 0;
    // one fine plane z of parity ODD (compile time: z0 is odd, the planes alternate); the carried rows are
    // explicit operands so that the loop below can be unrolled by two without moving them:
    // (zm, c) = the row in planes z-1, z -> nxt = plane z+1
    auto step = [&](auto tag, const int z, const double2& zm, const double2& c, double2& nxt) {
        constexpr bool ODD = decltype(tag)::value != 0;
        mbar_wait(b_u, par_u);
        mbar_wait(b_f, par_f);
        // ---- residual of the warp's row in plane z ----
        nxt = *reinterpret_cast<const double2*>(su + e_p + o_row);
        const double2 s2 = *reinterpret_cast<const double2*>(su + e_c + o_row - RB3_W);
        const double2 n2 = *reinterpret_cast<const double2*>(su + e_c + o_row + RB3_W);
        const double2 f2 = *reinterpret_cast<const double2*>(sf + e_f + o_f);
        const double left = __shfl_up_sync(0xffffffffu, c.y, 1), right = __shfl_down_sync(0xffffffffu, c.x, 1);
        double2 r2;
        r2.x = f2.x - star3_at(st, zm.x, s2.x, left, c.x, c.y, n2.x, nxt.x);
        r2.y = f2.y - star3_at(st, zm.y, s2.y, c.x, c.y, right, n2.y, nxt.y);
        double* const rb = sr + (ODD ? F_SLOT : 0);
        *reinterpret_cast<double2*>(rb + o_f) = r2;
        __syncthreads();
        if (warp == NW - 1) {
            // plane z-1 of u and plane z of f are dead now (the row values live on in registers): refill the slots
            fence_proxy_async();
            if (z - 1 + NS <= z1 + 1) stage_u(z - 1 + NS, u_slot);
            if (z + NF <= z1) stage_f(z + NF, f_slot);
        }
        // ---- the nine terms of plane z for the coarse nodes of this row ----
        if (coarse_warp) {
            const double2 ym = *reinterpret_cast<const double2*>(rb + o_f - RB3_W);
            const double2 yp = *reinterpret_cast<const double2*>(rb + o_f + RB3_W);
            const double lm = __shfl_up_sync(0xffffffffu, ym.y, 1), l0 = __shfl_up_sync(0xffffffffu, r2.y, 1),
                         lp = __shfl_up_sync(0xffffffffu, yp.y, 1);
            if (ODD) {
                const int Zdone = (z - 1) >> 1;                              // oz = +1 finishes the coarse plane below
                if (Zdone >= Zb) {
                    const double v = rs3_plane_terms(acc, false, 0.25, lm, ym, l0, r2, lp, yp);
                    if (store_ok) cptr[(long long)(Zdone - zoff_c) * gc.plane] = v;
                }
                acc = rs3_plane_terms(0.0, true, 0.25, lm, ym, l0, r2, lp, yp);      // oz = -1 starts the one above
            } else {
                acc = rs3_plane_terms(acc, false, 0.5, lm, ym, l0, r2, lp, yp);
            }
        }
        e_c = e_p;
        e_p += U_SLOT; b_u += 8;
        if (e_p == NS * U_SLOT) { e_p = 0; b_u = bar_a; par_u ^= 1; }
        e_f += F_SLOT; b_f += 8;
        if (e_f == NF * F_SLOT) { e_f = 0; b_f = bar_a + 8 * NS; par_f ^= 1; }
        if (++u_slot == NS) u_slot = 0;
        if (++f_slot == NF) f_slot = 0;
    };
    const std::integral_constant<int, 0> even;
    const std::integral_constant<int, 1> odd;
    double2 n1, n2;
    step(odd, z0, rzm, rc, n1);                                             // z0 and z1 are odd
    for (int z = z0 + 1; z < z1; z += 2) {
        step(even, z, rc, n1, n2);
        step(odd, z + 1, n1, n2, rzm);
        rc = n2; n1 = rzm;
    }
}

// ---- the same stage with two fine rows per warp (round 2, after k_rbgs3d2) ---------------------------------------
// Warp w owns the fine rows a = yb-1+2w (odd) and b = a+1 = 2 (Y0+w) (even = the coarse row Y0+w); the last warp
// owns the tile's last odd row alone.  Inside the pair the y neighbours of the residual come from registers; of the
// three residual rows a coarse node needs, two (a, b) are the warp's own and only the third (the next warp's row a)
// goes through shared memory: only odd rows are stored to the residual planes.  Per row pair 7 + 1 shared-memory
// accesses instead of 10 + 2, the waits / ring bookkeeping / barrier once.  Arithmetic and order as k_resrestrict3d.
// grid as k_resrestrict3d, block = 32 * (RYC + 1).
template <int RYC, int NS, int NF>
__global__ void __launch_bounds__(32 * (RYC + 1)) k_resrestrict3d2(Geom gc, Geom gf, Star3D st,
                                                                   const __grid_constant__ CUtensorMap tm_u,
                                                                   const __grid_constant__ CUtensorMap tm_f,
                                                                   double* __restrict__ coarse, int zchunk,
                                                                   int zoff_f, int zoff_c, int Zlo, int Zhi) {
    constexpr int RF = 2 * RYC, UR = RF + 3, FR = RF + 1, NW = RYC + 1;
    constexpr uint32_t U_BYTES = UR * RB3_W * 8, F_BYTES = FR * RB3_W * 8;
    constexpr int U_SLOT = UR * RB3_W, F_SLOT = FR * RB3_W;
    extern __shared__ __align__(128) unsigned char rs3_smem[];
    double* su = reinterpret_cast<double*>(rs3_smem + ((128u - (smem_u32(rs3_smem) & 127u)) & 127u));   // [NS][UR][64]
    double* sf = su + NS * U_SLOT;                                          // [NF][FR][64]
    double* sr = sf + NF * F_SLOT;                                          // [2][FR][64] residual planes (odd rows only)
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sr + 2 * F_SLOT);     // [NS] u, [NF] f
    pdl_sync();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nc = gc.n;
    const int X0 = 1 + blockIdx.x * RS3_TXC, Y0 = 1 + blockIdx.y * RYC, Zb = Zlo + blockIdx.z * zchunk;
    const int Ze = imin(Zb + zchunk, Zhi + 1);
    const int xb = 2 * X0, yb = 2 * Y0;
    const int z0 = 2 * Zb - 1, z1 = 2 * Ze - 1;
    const int p0 = z0 - 1;
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    auto stage_u = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * slot;
            mbar_arrive_expect_tx(bar, U_BYTES);
            tma_load_3d(su_a + (uint32_t)(slot * U_SLOT * 8), &tm_u, xb - 2, yb - 2, p - zoff_f, bar);
        }
    };
    auto stage_f = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * (NS + slot);
            mbar_arrive_expect_tx(bar, F_BYTES);
            tma_load_3d(sf_a + (uint32_t)(slot * F_SLOT * 8), &tm_f, xb - 2, yb - 1, p - zoff_f, bar);
        }
    };

    if (threadIdx.x == 0) {
        for (int i = 0; i < NS + NF; ++i) mbar_init(bar_a + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == NW - 1) {
        for (int i = 0; i < NS; ++i)
            if (p0 + i <= z1 + 1) stage_u(p0 + i, i);
        for (int i = 0; i < NF; ++i)
            if (z0 + i <= z1) stage_f(z0 + i, i);
    }
    mbar_wait(bar_a, 0);            // planes p0 and p0+1 of u
    mbar_wait(bar_a + 8, 0);

    const bool has_b = warp < RYC;                                           // the last warp owns a single (odd) row
    const int o_a = (2 * warp + 1) * RB3_W + 2 * lane;                       // row a inside a slot of u
    const int o_b = has_b ? o_a + RB3_W : o_a;                               // row b (the last warp re-reads row a: unused)
    const int of_a = 2 * warp * RB3_W + 2 * lane;                            // ... of f and of the residual planes
    const int of_b = has_b ? of_a + RB3_W : of_a;
    const int Y = Y0 + warp, X = X0 + lane - 1;
    const bool store_ok = has_b && lane >= 1 && lane <= RS3_TXC && X <= nc - 2 && Y <= nc - 2;
    double* cptr = coarse + (long long)Y * gc.pitch + X;

    double2 azm = *reinterpret_cast<const double2*>(su + o_a), bzm = *reinterpret_cast<const double2*>(su + o_b);   // plane z0-1
    double2 ac = *reinterpret_cast<const double2*>(su + U_SLOT + o_a), bc = *reinterpret_cast<const double2*>(su + U_SLOT + o_b);
    double acc = 0.0;

    int e_c = U_SLOT, e_p = 2 * U_SLOT, e_f = 0;
    int u_slot = 0, f_slot = 0;
    uint32_t b_u = bar_a + 16, b_f = bar_a + 8 * NS;
    uint32_t par_u = 0, par_f = 0;
    // one fine plane z of parity ODD; (zm, c) = a row in planes z-1, z -> nxt = plane z+1
    auto step = [&](auto tag, const int z, const double2& zma, const double2& ca, double2& nxa, const double2& zmb,
                    const double2& cb, double2& nxb) {
        constexpr bool ODD = decltype(tag)::value != 0;
        mbar_wait(b_u, par_u);
        mbar_wait(b_f, par_f);
        // ---- residuals of the rows a and b in plane z ----
        nxa = *reinterpret_cast<const double2*>(su + e_p + o_a);
        nxb = *reinterpret_cast<const double2*>(su + e_p + o_b);
        const double2 sa = *reinterpret_cast<const double2*>(su + e_c + o_a - RB3_W);       // row a-1: another warp's
        const double2 nb = *reinterpret_cast<const double2*>(su + e_c + o_b + RB3_W);       // row b+1 (last warp: row a+1)
        const double2 fa = *reinterpret_cast<const double2*>(sf + e_f + of_a);
        const double2 fb = *reinterpret_cast<const double2*>(sf + e_f + of_b);
        const double2 na = has_b ? cb : nb;                                                 // row a+1
        const double la = __shfl_up_sync(0xffffffffu, ca.y, 1), ra_ = __shfl_down_sync(0xffffffffu, ca.x, 1);
        double2 ra;
        ra.x = fa.x - star3_at(st, zma.x, sa.x, la, ca.x, ca.y, na.x, nxa.x);
        ra.y = fa.y - star3_at(st, zma.y, sa.y, ca.x, ca.y, ra_, na.y, nxa.y);
        const double lb = __shfl_up_sync(0xffffffffu, cb.y, 1), rb_ = __shfl_down_sync(0xffffffffu, cb.x, 1);
        double2 rb;
        rb.x = fb.x - star3_at(st, zmb.x, ca.x, lb, cb.x, cb.y, nb.x, nxb.x);
        rb.y = fb.y - star3_at(st, zmb.y, ca.y, cb.x, cb.y, rb_, nb.y, nxb.y);
        double* const rp = sr + (ODD ? F_SLOT : 0);
        *reinterpret_cast<double2*>(rp + of_a) = ra;                                        // odd rows: read by the warp below
        __syncthreads();
        if (warp == NW - 1) {
            fence_proxy_async();
            if (z - 1 + NS <= z1 + 1) stage_u(z - 1 + NS, u_slot);
            if (z + NF <= z1) stage_f(z + NF, f_slot);
        }
        // ---- the nine terms of plane z for the coarse nodes of row b: rows a, b (own), b+1 (the next warp's row a) ----
        if (has_b) {
            const double2 yp = *reinterpret_cast<const double2*>(rp + of_b + RB3_W);
            const double lm = __shfl_up_sync(0xffffffffu, ra.y, 1), l0 = __shfl_up_sync(0xffffffffu, rb.y, 1),
                         lp = __shfl_up_sync(0xffffffffu, yp.y, 1);
            if (ODD) {
                const int Zdone = (z - 1) >> 1;
                if (Zdone >= Zb) {
                    const double v = rs3_plane_terms(acc, false, 0.25, lm, ra, l0, rb, lp, yp);
                    if (store_ok) cptr[(long long)(Zdone - zoff_c) * gc.plane] = v;
                }
                acc = rs3_plane_terms(0.0, true, 0.25, lm, ra, l0, rb, lp, yp);
            } else {
                acc = rs3_plane_terms(acc, false, 0.5, lm, ra, l0, rb, lp, yp);
            }
        }
        e_c = e_p;
        e_p += U_SLOT; b_u += 8;
        if (e_p == NS * U_SLOT) { e_p = 0; b_u = bar_a; par_u ^= 1; }
        e_f += F_SLOT; b_f += 8;
        if (e_f == NF * F_SLOT) { e_f = 0; b_f = bar_a + 8 * NS; par_f ^= 1; }
        if (++u_slot == NS) u_slot = 0;
        if (++f_slot == NF) f_slot = 0;
    };
    const std::integral_constant<int, 0> even;
    const std::integral_constant<int, 1> odd;
    double2 an1, an2, bn1, bn2;
    step(odd, z0, azm, ac, an1, bzm, bc, bn1);                               // z0 and z1 are odd
    for (int z = z0 + 1; z < z1; z += 2) {
        step(even, z, ac, an1, an2, bc, bn1, bn2);
        step(odd, z + 1, an1, an2, azm, bn1, bn2, bzm);
        ac = an2; an1 = azm; bc = bn2; bn1 = bzm;
    }
}

}  // namespace mg
