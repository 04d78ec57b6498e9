// mg_rbgs3d2.cuh -- k_rbgs3d with TWO tile rows per warp (round 2).
//
// ncu of k_rbgs3d<8,7,6> at 513^3 (profiles/r2/ncu_3d_513.md): 166 warp instructions and 28 shared-memory
// wavefronts per row and plane, L1/shared pipe 84 % busy, 42 % of its wavefronts bank conflicts (the 8-byte
// accesses to one colour of a row have stride 16 B) -- the kernel is bound by the shared-memory pipe and by
// instruction issue, not by HBM (67 % of peak).  Here a thread owns the column pair of two adjacent rows
// (a, b = a + 1):
//   * the y neighbours inside the row pair come from registers (the node above a red node of row a is a black
//     node of row b, which the red half-sweep leaves alone; the node above a black node of row a is the red
//     node of row b this thread has just formed), so a row pair reads 2 + 2 neighbour values from shared
//     memory instead of 4 + 4: 20 wavefronts per row and plane instead of 28;
//   * the barrier waits, the ring bookkeeping, the __syncthreads and the predicates are paid once per row pair.
// Arithmetic, operand order and therefore results are those of k_rbgs3d (and of two colour passes): the
// parity tests compare both with the oracle.  Staging, rings, slab arguments: see mg_rbgs3d.cuh.
#pragma once
#include "mg_rbgs3d.cuh"

namespace mg {

template <int C>
__device__ __forceinline__ void rb3_set(double2& v, double x) {
    if (C) v.y = x; else v.x = x;
}

// registers of one row of the pair along the march: original planes (zm, c, nxt), red planes (bz, b, pr), f (fq, fp);
// two sets of the "next" values because the plane loop is unrolled by two
struct RB3Row {
    double2 zm, c, n1, n2, bz, b, p1, p2, fq, f1, f2;
};

// grid = (ceil(n / 60), ceil(n / RY), ceil(planes / zchunk)), block = 16 * (RY + 2): warp k owns the tile rows
// yb-1+2k and yb+2k
template <int RY, int RB3_NS, int RB3_NF>
__global__ void __launch_bounds__(16 * (RY + 2)) k_rbgs3d2(Geom g, Star3D st, const __grid_constant__ CUtensorMap tm_u,
                                                           const __grid_constant__ CUtensorMap tm_f,
                                                           double* __restrict__ out, double w, int zchunk, int pf,
                                                           int zoff, int zlo, int zhi) {
    constexpr int UR = RY + 4, FR = RY + 2, NW = (RY + 2) / 2;
    static_assert(RY % 2 == 0, "tile rows start on even y");
    constexpr uint32_t U_BYTES = UR * RB3_W * 8, F_BYTES = FR * RB3_W * 8;
    constexpr int U_SLOT = UR * RB3_W, F_SLOT = FR * RB3_W;
    extern __shared__ __align__(128) unsigned char rb3_smem[];
    double* su = reinterpret_cast<double*>(rb3_smem + ((128u - (smem_u32(rb3_smem) & 127u)) & 127u));   // [NS][UR][64]
    double* sf = su + RB3_NS * U_SLOT;                                      // [NF][FR][64]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sf + RB3_NF * F_SLOT);   // [NS] u, [NF] f
    pdl_sync();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = g.n;
    const long long pitch = g.pitch, plane = g.plane;
    const int xb = blockIdx.x * RB3_TW, yb = blockIdx.y * RY, zb = zlo + blockIdx.z * zchunk;
    const int ze = imin(zb + zchunk, zhi + 1);
    const int p0 = zb - 2, fp0 = zb - 1;
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    auto stage_u = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * slot;
            mbar_arrive_expect_tx(bar, U_BYTES);
            tma_load_3d(su_a + (uint32_t)(slot * U_SLOT * 8), &tm_u, xb - 2, yb - 2, p - zoff, bar);
        }
    };
    auto stage_f = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * (RB3_NS + slot);
            mbar_arrive_expect_tx(bar, F_BYTES);
            tma_load_3d(sf_a + (uint32_t)(slot * F_SLOT * 8), &tm_f, xb - 2, yb - 1, p - zoff, bar);
        }
    };
    auto back = [](int slot, int k, int ring) { const int v = slot - k; return v < 0 ? v + ring : v; };

    if (threadIdx.x == 0) {
        for (int i = 0; i < RB3_NS + RB3_NF; ++i) mbar_init(bar_a + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == NW - 1) {
        for (int i = 0; i < RB3_NS; ++i)
            if (p0 + i <= ze + 1) stage_u(p0 + i, i);
        for (int i = 0; i < RB3_NF; ++i)
            if (fp0 + i <= ze) stage_f(fp0 + i, i);
    }
    mbar_wait(bar_a, 0);            // planes p0 and p0+1 of u
    mbar_wait(bar_a + 8, 0);

    // ---- per-thread constants of the march ----
    const int x0 = xb - 2 + 2 * lane;                                       // first column of the lane's pair
    const int ya = yb - 1 + 2 * warp;                                        // rows a = ya, b = ya + 1
    const int o_a = (2 * warp + 1) * RB3_W + 2 * lane, o_b = o_a + RB3_W;    // the pairs inside a slot of u
    const int of_a = 2 * warp * RB3_W + 2 * lane, of_b = of_a + RB3_W;       // ... of f (the f tile starts one row later)
    // per row, as in k_rbgs3d: bit c: the red update of component c is kept; bit 2+c: component c is an interior
    // node in x and y (black update); bit 4: the row is stored as whole pairs, bit 5: only its first column
    unsigned oka = 0, okb = 0;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int y = ya + r;
        unsigned ok = 0;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int x = x0 + c;
            if (y >= 1 && y <= n - 2 && x >= imax(1, xb - 1) && x <= imin(n - 2, xb + RB3_TW)) ok |= 1u << c;
            if (y >= 1 && y <= n - 2 && x >= 1 && x <= n - 2) ok |= 4u << c;
        }
        const bool out_row = y >= yb && y <= yb + RY - 1;                    // rows yb-1 and yb+RY are halo rows
        if (out_row && lane >= 1 && lane <= RB3_TW / 2 && y <= n - 1) {
            if (x0 + 1 <= n - 1) ok |= 16u;
            else if (x0 <= n - 1) ok |= 32u;
        }
        if (r == 0) oka = ok; else okb = ok;
    }
    const int par_row = ya & 1;                                              // + s: component of row a's red node in plane s
    double* optr = out + (long long)(zb - 2 - zoff) * plane + (long long)ya * pitch + x0;   // row a, plane s-1 at s = zb-1

    RB3Row A, B;
    A.zm = *reinterpret_cast<const double2*>(su + o_a);                     // plane zb-2 (slot 0)
    B.zm = *reinterpret_cast<const double2*>(su + o_b);
    A.c = *reinterpret_cast<const double2*>(su + U_SLOT + o_a);             // plane zb-1 (slot 1)
    B.c = *reinterpret_cast<const double2*>(su + U_SLOT + o_b);
    A.bz = A.b = A.fq = B.bz = B.b = B.fq = make_double2(0.0, 0.0);

    int e_m2 = 0, e_m = 0, e_s = U_SLOT, e_p = 2 * U_SLOT, e_f1 = 0, e_f = 0;
    uint32_t b_u = bar_a + 16, b_f = bar_a + 8 * RB3_NS;
    uint32_t par_u = 0, par_f = 0;
    // one step: C = component of the red node of row a in plane s (row b: 1 - C) = of the black nodes in plane s-1.
    // (zm, c) original planes s-1, s -> nxt = plane s+1; (bz, b) red planes s-2, s-1 -> pr = red plane s; fq = f(s-1) -> fp = f(s)
    auto step = [&](auto tag, const int s, const double2& zma, const double2& ca, double2& nxa, const double2& bza,
                    const double2& ba, double2& pra, const double2& fqa, double2& fpa, const double2& zmb,
                    const double2& cb, double2& nxb, const double2& bzb, const double2& bb, double2& prb,
                    const double2& fqb, double2& fpb) {
        constexpr int C = decltype(tag)::value, D = 1 - C;
        mbar_wait(b_u, par_u);
        mbar_wait(b_f, par_f);
        double* const us = su + e_s;
        const double* const um = su + e_m;
        nxa = *reinterpret_cast<const double2*>(su + e_p + o_a);
        nxb = *reinterpret_cast<const double2*>(su + e_p + o_b);
        fpa = *reinterpret_cast<const double2*>(sf + e_f + of_a);
        fpb = *reinterpret_cast<const double2*>(sf + e_f + of_b);
        const int q = s - 1;
        pra = ca;
        prb = cb;
        // ---- red(s): row a component C, row b component D ----
        if (s >= 1 && s <= n - 2) {
            {
                const double wv = C ? ca.x : __shfl_up_sync(0xffffffffu, ca.y, 1);
                const double ev = C ? __shfl_down_sync(0xffffffffu, ca.x, 1) : ca.y;
                const double sv = us[o_a + C - RB3_W], nv = rb3_get<C>(cb);
                const double cen = rb3_get<C>(ca);
                const double r = rb3_get<C>(fpa) - star3_at(st, rb3_get<C>(zma), sv, wv, cen, ev, nv, rb3_get<C>(nxa));
                const double v = cen + w * r;
                if (oka & (1u << C)) {
                    us[o_a + C] = v;
                    rb3_set<C>(pra, v);
                }
            }
            {
                const double wv = D ? cb.x : __shfl_up_sync(0xffffffffu, cb.y, 1);
                const double ev = D ? __shfl_down_sync(0xffffffffu, cb.x, 1) : cb.y;
                const double sv = rb3_get<D>(ca), nv = us[o_b + D + RB3_W];
                const double cen = rb3_get<D>(cb);
                const double r = rb3_get<D>(fpb) - star3_at(st, rb3_get<D>(zmb), sv, wv, cen, ev, nv, rb3_get<D>(nxb));
                const double v = cen + w * r;
                if (okb & (1u << D)) {
                    us[o_b + D] = v;
                    rb3_set<D>(prb, v);
                }
            }
        }
        __syncthreads();
        if (warp == NW - 1) {
            fence_proxy_async();
            if (s - 3 >= p0 && s - 3 + RB3_NS <= ze + 1) stage_u(s - 3 + RB3_NS, back(e_m2 / U_SLOT, 1, RB3_NS));
            if (s - 2 >= fp0 && s - 2 + RB3_NF <= ze) stage_f(s - 2 + RB3_NF, back(e_f1 / F_SLOT, 1, RB3_NF));
            if (pf > 0 && lane == 0 && s + pf <= imin(ze + 1, n - 1)) {
                tma_prefetch_3d(&tm_u, xb - 2, yb - 2, s + pf - zoff);
                if (s + pf <= ze) tma_prefetch_3d(&tm_f, xb - 2, yb - 1, s + pf - zoff);
            }
        }
        // ---- black(s-1): row a component C, row b component D; the finished rows of plane s-1 are stored ----
        if (q >= zb && q < ze) {
            const bool inner = q >= 1 && q <= n - 2;
            {
                const double wv = C ? ba.x : __shfl_up_sync(0xffffffffu, ba.y, 1);
                const double ev = C ? __shfl_down_sync(0xffffffffu, ba.x, 1) : ba.y;
                const double sv = um[o_a + C - RB3_W], nv = rb3_get<C>(bb);
                const double cen = rb3_get<C>(ba);
                const double r = rb3_get<C>(fqa) - star3_at(st, rb3_get<C>(bza), sv, wv, cen, ev, nv, rb3_get<C>(pra));
                const double v = cen + w * r;
                double2 o2 = ba;
                if (inner && (oka & (4u << C))) rb3_set<C>(o2, v);
                if (oka & 16u) st2(optr, o2);
                else if (oka & 32u) *optr = o2.x;
            }
            {
                const double wv = D ? bb.x : __shfl_up_sync(0xffffffffu, bb.y, 1);
                const double ev = D ? __shfl_down_sync(0xffffffffu, bb.x, 1) : bb.y;
                const double sv = rb3_get<D>(ba), nv = um[o_b + D + RB3_W];
                const double cen = rb3_get<D>(bb);
                const double r = rb3_get<D>(fqb) - star3_at(st, rb3_get<D>(bzb), sv, wv, cen, ev, nv, rb3_get<D>(prb));
                const double v = cen + w * r;
                double2 o2 = bb;
                if (inner && (okb & (4u << D))) rb3_set<D>(o2, v);
                if (okb & 16u) st2(optr + pitch, o2);
                else if (okb & 32u) optr[pitch] = o2.x;
            }
        }
        optr += plane;
        e_m2 = e_m; e_m = e_s; e_s = e_p;
        e_p += U_SLOT; b_u += 8;
        if (e_p == RB3_NS * U_SLOT) { e_p = 0; b_u = bar_a; par_u ^= 1; }
        e_f1 = e_f;
        e_f += F_SLOT; b_f += 8;
        if (e_f == RB3_NF * F_SLOT) { e_f = 0; b_f = bar_a + 8 * RB3_NS; par_f ^= 1; }
    };
    const std::integral_constant<int, 0> even;
    const std::integral_constant<int, 1> odd;
    int s = zb - 1;
    if ((par_row + s) & 1) {                                                 // peel: the loop starts on component 0
        step(odd, s, A.zm, A.c, A.n1, A.bz, A.b, A.p1, A.fq, A.f1, B.zm, B.c, B.n1, B.bz, B.b, B.p1, B.fq, B.f1);
        A.zm = A.c; A.c = A.n1; A.bz = A.b; A.b = A.p1; A.fq = A.f1;
        B.zm = B.c; B.c = B.n1; B.bz = B.b; B.b = B.p1; B.fq = B.f1;
        ++s;
    }
    for (; s + 1 <= ze; s += 2) {
        step(even, s, A.zm, A.c, A.n1, A.bz, A.b, A.p1, A.fq, A.f1, B.zm, B.c, B.n1, B.bz, B.b, B.p1, B.fq, B.f1);
        step(odd, s + 1, A.c, A.n1, A.n2, A.b, A.p1, A.p2, A.f1, A.f2, B.c, B.n1, B.n2, B.b, B.p1, B.p2, B.f1, B.f2);
        A.zm = A.n1; A.c = A.n2; A.bz = A.p1; A.b = A.p2; A.fq = A.f2;
        B.zm = B.n1; B.c = B.n2; B.bz = B.p1; B.b = B.p2; B.fq = B.f2;
    }
    if (s <= ze) step(even, s, A.zm, A.c, A.n1, A.bz, A.b, A.p1, A.fq, A.f1, B.zm, B.c, B.n1, B.bz, B.b, B.p1, B.fq, B.f1);
}

}  // namespace mg
