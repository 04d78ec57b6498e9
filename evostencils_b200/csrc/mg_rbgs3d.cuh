// mg_rbgs3d.cuh -- one-pass red-black Gauss-Seidel sweep for the 3D 7-point star (3D Poisson).
//
// Two out-of-place colour passes move 2 x 24 B per DoF.  This kernel moves 24: it forms the red
// half-sweep in shared memory and applies the black half-sweep on top of it before anything is
// written, like k_rbgs2d does in registers for 2D.
//
//   * A CTA owns a tile of TW = 60 columns x RY rows and marches over a chunk of z planes.  Per plane
//     it keeps RY+4 rows x 64 columns of u (halo of 2: the black update of a node needs the red
//     values at distance 1, which need the original values at distance 2) and RY+2 rows of f.
//   * Planes are staged by TMA through a tensor map of the whole array (host: cuTensorMapEncodeTiled,
//     mg_kernels_fast.cuh): one `cp.async.bulk.tensor.3d` box per plane tile of u and one of f, issued by
//     one lane, completing on the ring slot's mbarrier; rows, columns and planes outside the grid arrive
//     as zeros.  Rings of NS planes of u and NF of f; a slot is re-armed right after the barrier that
//     retires its plane, so about 3 planes are in flight while the warps compute.  No thread ever
//     waits on a global load.  (One 512-byte `cp.async.bulk` per row -- 22 copies per plane, serialised
//     lane by lane through the uniform datapath -- cost 1.03 ms per sweep at 513^3 against 0.66.)
//   * Step s of the march: red(s) updates the red nodes of plane s in place in shared memory (red
//     nodes read only black nodes, which the red half-sweep leaves alone, so planes s-1, s, s+1 may be
//     in any red state); after one __syncthreads, black(s-1) reads the red nodes of planes s-2, s-1, s
//     around every black node of plane s-1 and writes the finished plane s-1 (both colours, boundary
//     nodes copied) to `out`.  Black results never go back to shared memory: nothing reads them.
//   * One warp per tile row, in both half-sweeps: lane j holds the aligned column pair (xb-2+2j,
//     xb-1+2j) of its row in the planes around s in registers (see k_rbgs3d); x neighbours come from the
//     adjacent lane by shuffle, only the y neighbours (rows of other warps) from shared memory.  The
//     component of the pair that is red in plane s (= black in plane s-1) alternates from plane to
//     plane: the plane loop is unrolled by two with the component as a template argument.
//   * Red values in the halo (rows yb-1 and yb+RY, columns xb-1 and xb+60, planes zb-1 and ze) are
//     recomputed by this CTA from the same inputs as their owner uses: bit-identical to two
//     sequential colour passes (operand order of star3_at, separate multiply and add).
#pragma once
#include <cuda.h>

#include <type_traits>

#include "mg_fast3d.cuh"

namespace mg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// one box of a 3D tensor map (x fastest) -> shared memory; out-of-range coordinates are filled with zeros
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, int x, int y, int z, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
        "l"(tm), "r"(x), "r"(y), "r"(z), "r"(bar)
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap* tm, int x, int y, int z) {
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(tm), "r"(x), "r"(y), "r"(z)
                 : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}
__device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity) {
    for (uint32_t tries = 0; !mbar_try_wait(bar, parity); ++tries)
        if (tries == (1u << 20)) __trap();          // a transfer that never completes: fail, do not hang
}
// The first try_wait (which itself suspends the thread for a hardware-defined time) almost always succeeds: the
// planes are staged three steps ahead.  The bounded retry loop is kept out of line -- inlined, its counter, YIELD and
// reconvergence barrier cost nine instructions per wait, two waits per plane and warp.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (!mbar_try_wait(bar, parity)) mbar_wait_slow(bar, parity);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int RB3_TW = 60;     // output columns per tile
constexpr int RB3_W = 64;      // staged columns per row (xb-2 .. xb+61)

template <int RY, int RB3_NS, int RB3_NF>      // RB3_NS / RB3_NF: ring sizes of u / f planes
constexpr size_t rbgs3d_smem_bytes() {
    return (size_t)(RB3_NS * (RY + 4) + RB3_NF * (RY + 2)) * RB3_W * sizeof(double) + 16 * sizeof(unsigned long long) + 128;
}

template <int C>
__device__ __forceinline__ double rb3_get(const double2 v) { return C ? v.y : v.x; }

// out = RB-GS sweep of u (red = (x+y+z) even first, then black), whole grid (g.nz == g.n, g.z0 == 0).
// grid = (ceil(n / 60), ceil(n / RY), ceil(n / zchunk)), block = 32 * (RY + 2): warp k owns tile row k+1
// (y = yb-1+k) in both half-sweeps and carries its column pairs along z in registers:
//   rzm, rc, rn   the row in the original planes s-1, s, s+1 (rn is read one step ahead: it is the z+
//                 neighbour now and the centre of the next step; rzm is the z- neighbour),
//   bzm, bc, pr   the row after the red half-sweep in planes s-2, s-1, s (pr is formed by this thread in
//                 red(s), so the black half-sweep never re-reads its own row),
//   fp, fq        the right-hand side of the row in planes s and s-1.
// Per step a thread reads two 16-byte pairs (rn, fp) and four 8-byte values (the y neighbours of the red
// and of the black node, which other warps own) from shared memory.
// z-slabs (domain decomposition): the tensor maps and `out` describe the slab array including its ghost planes,
// plane 0 of them being the global plane `zoff`; the sweep writes the global planes zlo .. zhi (the owned ones,
// plus the global Dirichlet plane on the first / last slab) and reads two planes beyond them -- the slab
// carries two ghost planes on either side for this kernel.  Whole grids: zoff = 0, zlo = 0, zhi = n-1.
template <int RY, int RB3_NS, int RB3_NF>
__global__ void __launch_bounds__(32 * (RY + 2)) k_rbgs3d(Geom g, Star3D st, const __grid_constant__ CUtensorMap tm_u,
                                                          const __grid_constant__ CUtensorMap tm_f,
                                                          double* __restrict__ out, double w, int zchunk, int pf,
                                                          int zoff, int zlo, int zhi) {
    constexpr int UR = RY + 4, FR = RY + 2, NW = RY + 2;
    static_assert(RY % 2 == 0, "tile rows start on even y");
    constexpr uint32_t U_BYTES = UR * RB3_W * 8, F_BYTES = FR * RB3_W * 8;
    constexpr int U_SLOT = UR * RB3_W, F_SLOT = FR * RB3_W;                  // doubles per ring slot
    extern __shared__ __align__(128) unsigned char rb3_smem[];
    // TMA destinations must be 128-byte aligned
    double* su = reinterpret_cast<double*>(rb3_smem + ((128u - (smem_u32(rb3_smem) & 127u)) & 127u));   // [NS][UR][64]
    double* sf = su + RB3_NS * U_SLOT;                                      // [NF][FR][64]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sf + RB3_NF * F_SLOT);   // [NS] u, [NF] f
    pdl_sync();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = g.n;
    const long long pitch = g.pitch, plane = g.plane;
    const int xb = blockIdx.x * RB3_TW, yb = blockIdx.y * RY, zb = zlo + blockIdx.z * zchunk;
    const int ze = imin(zb + zchunk, zhi + 1);                              // output planes zb .. ze-1 (global indices)
    const int p0 = zb - 2, fp0 = zb - 1;                                    // first staged plane of u / f
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    // stage plane p of u (rows yb-2 .. yb+RY+1, columns xb-2 .. xb+61) into ring slot `slot` / of f (rows
    // yb-1 .. yb+RY): one TMA box each; rows, columns and planes outside the grid arrive as zeros and are
    // only read for nodes that are not updated
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
    const int y = yb - 1 + warp;                                             // the warp's row
    const int o_row = (warp + 1) * RB3_W + 2 * lane;                         // its pair inside a slot of u
    const int o_f = warp * RB3_W + 2 * lane;                                 // ... of f (the f tile starts one row later)
    // bit c: the red update of component c is kept (interior node inside the red region); bit 2+c:
    // component c is an interior node in x and y (black update); bit 4: the row is stored as whole pairs,
    // bit 5: only its first column (x0 == n-1)
    unsigned ok = 0;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int x = x0 + c;
        if (y >= 1 && y <= n - 2 && x >= imax(1, xb - 1) && x <= imin(n - 2, xb + RB3_TW)) ok |= 1u << c;
        if (y >= 1 && y <= n - 2 && x >= 1 && x <= n - 2) ok |= 4u << c;
    }
    if (warp >= 1 && warp <= RY && lane >= 1 && lane <= RB3_TW / 2 && y <= n - 1) {
        if (x0 + 1 <= n - 1) ok |= 16u;
        else if (x0 <= n - 1) ok |= 32u;
    }
    const bool black_warp = warp >= 1 && warp <= RY;
    const int par_row = y & 1;                                               // + s: component of the red node (plane s)
                                                                             //      = of the black node (plane s-1)
    double* optr = out + (long long)(zb - 2 - zoff) * plane + (long long)y * pitch + x0;   // plane s-1 at s = zb-1

    double2 rzm = *reinterpret_cast<const double2*>(su + o_row);            // plane zb-2 (slot 0)
    double2 rc = *reinterpret_cast<const double2*>(su + U_SLOT + o_row);    // plane zb-1 (slot 1)
    double2 bzm = make_double2(0.0, 0.0), bc = bzm, fq = bzm;

    // ring state of step s = zb-1: element offsets of the slots of planes s-2, s-1, s, s+1 of u and of
    // planes s-1, s of f; barrier address and phase parity of plane s+1 of u and of plane s of f
    int e_m2 = 0, e_m = 0, e_s = U_SLOT, e_p = 2 * U_SLOT, e_f1 = 0, e_f = 0;
    uint32_t b_u = bar_a + 16, b_f = bar_a + 8 * RB3_NS;
    uint32_t par_u = 0, par_f = 0;
    // one step of the march with the red / black node in component C of the pair; the carried rows are
    // explicit operands so that the loop below can be unrolled by two without moving them:
    // (zm, c) original planes s-1, s -> nxt = plane s+1; (bz, b) red planes s-2, s-1 -> pr = red plane s;
    // fq = f(s-1) -> fp = f(s)
    auto step = [&](auto tag, const int s, const double2& zm, const double2& c, double2& nxt, const double2& bz,
                    const double2& b, double2& pr, const double2& fq, double2& fp) {
        constexpr int C = decltype(tag)::value;
        mbar_wait(b_u, par_u);
        mbar_wait(b_f, par_f);
        double* const us = su + e_s;
        const double* const um = su + e_m;
        nxt = *reinterpret_cast<const double2*>(su + e_p + o_row);
        fp = *reinterpret_cast<const double2*>(sf + e_f + o_f);
        const int q = s - 1;
        pr = c;
        // ---- red(s) ----
        if (s >= 1 && s <= n - 2) {
            const double wv = C ? c.x : __shfl_up_sync(0xffffffffu, c.y, 1);
            const double ev = C ? __shfl_down_sync(0xffffffffu, c.x, 1) : c.y;
            const double sv = us[o_row + C - RB3_W], nv = us[o_row + C + RB3_W];
            const double cen = rb3_get<C>(c);
            const double r = rb3_get<C>(fp) - star3_at(st, rb3_get<C>(zm), sv, wv, cen, ev, nv, rb3_get<C>(nxt));
            const double v = cen + w * r;
            if (ok & (1u << C)) {
                us[o_row + C] = v;
                if (C) pr.y = v; else pr.x = v;
            }
        }
        __syncthreads();
        if (warp == NW - 1) {
            // planes <= s-3 of u and <= s-2 of f are dead now: refill their slots (the proxy fence orders
            // the red stores other warps made into them, visible here through the barrier, before the
            // asynchronous writes)
            fence_proxy_async();
            if (s - 3 >= p0 && s - 3 + RB3_NS <= ze + 1) stage_u(s - 3 + RB3_NS, back(e_m2 / U_SLOT, 1, RB3_NS));
            if (s - 2 >= fp0 && s - 2 + RB3_NF <= ze) stage_f(s - 2 + RB3_NF, back(e_f1 / F_SLOT, 1, RB3_NF));
            if (pf > 0 && lane == 0 && s + pf <= imin(ze + 1, n - 1)) {     // the boxes of plane s+pf -> L2
                tma_prefetch_3d(&tm_u, xb - 2, yb - 2, s + pf - zoff);
                if (s + pf <= ze) tma_prefetch_3d(&tm_f, xb - 2, yb - 1, s + pf - zoff);
            }
        }
        // ---- black(s-1), then the finished row of plane s-1 is stored ----
        if (q >= zb && q < ze && black_warp) {
            const double wv = C ? b.x : __shfl_up_sync(0xffffffffu, b.y, 1);
            const double ev = C ? __shfl_down_sync(0xffffffffu, b.x, 1) : b.y;
            const double sv = um[o_row + C - RB3_W], nv = um[o_row + C + RB3_W];
            const double cen = rb3_get<C>(b);
            const double r = rb3_get<C>(fq) - star3_at(st, rb3_get<C>(bz), sv, wv, cen, ev, nv, rb3_get<C>(pr));
            const double v = cen + w * r;
            double2 o2 = b;
            if (q >= 1 && q <= n - 2 && (ok & (4u << C))) {
                if (C) o2.y = v; else o2.x = v;
            }
            if (ok & 16u) st2(optr, o2);
            else if (ok & 32u) *optr = o2.x;
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
    double2 n1, n2, p1, p2, f1, f2;
    int s = zb - 1;
    if ((par_row + s) & 1) {                                                 // peel: the loop starts on component 0
        step(odd, s, rzm, rc, n1, bzm, bc, p1, fq, f1);
        rzm = rc; rc = n1; bzm = bc; bc = p1; fq = f1;
        ++s;
    }
    for (; s + 1 <= ze; s += 2) {
        step(even, s, rzm, rc, n1, bzm, bc, p1, fq, f1);
        step(odd, s + 1, rc, n1, n2, bc, p1, p2, f1, f2);
        rzm = n1; rc = n2; bzm = p1; bc = p2; fq = f2;
    }
    if (s <= ze) step(even, s, rzm, rc, n1, bzm, bc, p1, fq, f1);
}

}  // namespace mg
