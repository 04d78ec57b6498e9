// mg_rbgs3d.cuh -- one-pass red-black Gauss-Seidel sweep for the 3D 7-point star (3D Poisson).
//
// Two out-of-place colour passes move 2 x 24 B per DoF.  This kernel moves 24: it forms the red
// half-sweep in shared memory and applies the black half-sweep on top of it before anything is
// written, like k_rbgs2d does in registers for 2D.
//
//   * A CTA owns a tile of TW = 60 columns x RY rows and marches over a chunk of z planes.  Per plane
//     it keeps RY+4 rows x 64 columns of u (halo of 2: the black update of a node needs the red
//     values at distance 1, which need the original values at distance 2) and RY+2 rows of f.
//   * Planes are staged by TMA: every row of a plane tile is one `cp.async.bulk` (512 B, global ->
//     shared) that completes on the plane's mbarrier; NS = 6 planes of u and NF = 5 of f form rings,
//     a slot is refilled as soon as its plane is dead, so 2-3 planes are always in flight while
//     the warps compute.  No thread ever waits on a global load.
//   * Step s of the march: red(s) updates the red nodes of plane s in place in shared memory (red
//     nodes read only black nodes, which the red half-sweep leaves alone, so planes s-1, s, s+1 may be
//     in any red state); after one __syncthreads, black(s-1) reads the red nodes of planes s-2, s-1, s
//     around every black node of plane s-1 and writes the finished plane s-1 (both colours, boundary
//     nodes copied) to `out`.  Black results never go back to shared memory: nothing reads them.
//   * One warp per tile row: lane j holds the aligned column pair (xb-2+2j, xb-1+2j); x neighbours come
//     from the adjacent lane by shuffle, y and z neighbours from shared memory.
//   * Red values in the halo (rows yb-1 and yb+RY, columns xb-1 and xb+60, planes zb-1 and ze) are
//     recomputed by this CTA from the same inputs as their owner uses: bit-identical to two
//     sequential colour passes (operand order of star3_at, separate multiply and add).
#pragma once
#include <cuda.h>

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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr int RB3_TW = 60;     // output columns per tile
constexpr int RB3_W = 64;      // staged columns per row (xb-2 .. xb+61): 512 B, one bulk copy

template <int RY, int RB3_NS, int RB3_NF>      // RB3_NS / RB3_NF: ring sizes of u / f planes
constexpr size_t rbgs3d_smem_bytes() {
    return (size_t)(RB3_NS * (RY + 4) + RB3_NF * (RY + 2)) * RB3_W * sizeof(double) + 16 * sizeof(unsigned long long) + 128;
}

// u + w (f - A u) at component C of a lane's column pair c2; o = element offset of the pair in a ring
// slot (um / uq / up: the staged planes below, at and above the node; fq: the right-hand side, whose tile
// starts one row later).  The parity of a row is warp-uniform, so the two components are two code paths:
// one shuffle each, no selects.
template <int C>
__device__ __forceinline__ double rb3_relax(const Star3D& st, double w, const double* um, const double* uq, const double* up,
                                            const double* fq, int o, const double2 c2) {
    double cen, wv, ev;
    if (C == 0) {
        cen = c2.x; ev = c2.y;
        wv = __shfl_up_sync(0xffffffffu, c2.y, 1);
    } else {
        cen = c2.y; wv = c2.x;
        ev = __shfl_down_sync(0xffffffffu, c2.x, 1);
    }
    const double sv = uq[o + C - RB3_W], nv = uq[o + C + RB3_W];
    const double zm = um[o + C], zp = up[o + C];
    const double fv = fq[o + C - RB3_W];
    const double r = fv - star3_at(st, zm, sv, wv, cen, ev, nv, zp);
    return cen + w * r;
}

// out = RB-GS sweep of u (red = (x+y+z) even first, then black), whole grid (g.nz == g.n, g.z0 == 0).
// grid = (ceil(n / 60), ceil(n / RY), ceil(n / zchunk)), block = 32 * (RY + 2) / RW (RW tile rows per warp).
template <int RY, int RB3_NS, int RB3_NF, int RW>
__global__ void __launch_bounds__(32 * (RY + 2) / RW) k_rbgs3d(Geom g, Star3D st, const __grid_constant__ CUtensorMap tm_u,
                                                               const __grid_constant__ CUtensorMap tm_f,
                                                               double* __restrict__ out, double w, int zchunk) {
    constexpr int UR = RY + 4, FR = RY + 2, NW = (RY + 2) / RW;
    static_assert((RY + 2) % RW == 0 && RY % RW == 0 && RY % 2 == 0, "rows per warp must divide the red and the black rows");
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
    const int xb = blockIdx.x * RB3_TW, yb = blockIdx.y * RY, zb = blockIdx.z * zchunk;
    const int ze = imin(zb + zchunk, n);                                    // output planes zb .. ze-1
    const int p0 = zb - 2, fp0 = zb - 1;                                    // first staged plane of u / f
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    // stage plane p of u (rows yb-2 .. yb+RY+1, columns xb-2 .. xb+61) into ring slot `slot` / of f (rows
    // yb-1 .. yb+RY): one TMA box each; rows, columns and planes outside the grid arrive as zeros and are
    // only read for nodes that are not updated
    auto stage_u = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * slot;
            mbar_arrive_expect_tx(bar, U_BYTES);
            tma_load_3d(su_a + (uint32_t)(slot * U_SLOT * 8), &tm_u, xb - 2, yb - 2, p, bar);
        }
    };
    auto stage_f = [&](int p, int slot) {
        if (lane == 0) {
            const uint32_t bar = bar_a + 8 * (RB3_NS + slot);
            mbar_arrive_expect_tx(bar, F_BYTES);
            tma_load_3d(sf_a + (uint32_t)(slot * F_SLOT * 8), &tm_f, xb - 2, yb - 1, p, bar);
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
    const int o_red = (warp * RW + 1) * RB3_W + 2 * lane;                    // pair of the first red row in a slot
    const int o_blk = o_red + RB3_W;                                         // ... of the first black row
    // bit 2i+c: the red update of component c in red row i is kept (interior node inside the red region);
    // bit 8+2i+c: component c of black row i is an interior node (in x and y)
    unsigned ok = 0;
#pragma unroll
    for (int i = 0; i < RW; ++i)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int x = x0 + c, yr = yb - 1 + warp * RW + i, yk = yr + 1;
            if (yr >= 1 && yr <= n - 2 && x >= imax(1, xb - 1) && x <= imin(n - 2, xb + RB3_TW)) ok |= 1u << (2 * i + c);
            if (yk >= 1 && yk <= n - 2 && x >= 1 && x <= n - 2) ok |= 1u << (8 + 2 * i + c);
        }
    // store of black row i: bit 16+i a whole pair, bit 20+i only its first column (x0 == n-1)
    const int y_blk = yb + warp * RW;
#pragma unroll
    for (int i = 0; i < RW; ++i)
        if (lane >= 1 && lane <= RB3_TW / 2 && y_blk + i <= n - 1) {
            if (x0 + 1 <= n - 1) ok |= 1u << (16 + i);
            else if (x0 <= n - 1) ok |= 1u << (20 + i);
        }
    const bool black_warp = warp * RW < RY;
    const int par_red = (yb - 1 + warp * RW) & 1;                            // + s: component of the red node in red row 0
    double* optr = out + (long long)(zb - 2) * plane + (long long)y_blk * pitch + x0;   // plane s-1 at s = zb-1

    // ring state of step s = zb-1: element offsets of the slots of planes s-2, s-1, s, s+1 of u and of
    // planes s-1, s of f; barrier address and phase parity of plane s+1 of u and of plane s of f
    int e_m2 = 0, e_m = 0, e_s = U_SLOT, e_p = 2 * U_SLOT, e_f1 = 0, e_f = 0;
    uint32_t b_u = bar_a + 16, b_f = bar_a + 8 * RB3_NS;
    uint32_t par_u = 0, par_f = 0;
    for (int s = zb - 1; s <= ze; ++s) {
        mbar_wait(b_u, par_u);
        mbar_wait(b_f, par_f);
        double* const us = su + e_s;
        const double* const um = su + e_m;
        const double* const up = su + e_p;
        // ---- red(s): tile rows 1 .. RY+2 (y = yb-1 .. yb+RY) ----
        if (s >= 1 && s <= n - 2) {
            const double* const fs = sf + e_f;
            const int c0 = (par_red + s) & 1;
#pragma unroll
            for (int i = 0; i < RW; ++i) {
                const int o = o_red + i * RB3_W;
                const double2 c2 = *reinterpret_cast<const double2*>(us + o);
                if ((c0 ^ (i & 1)) == 0) {
                    const double v = rb3_relax<0>(st, w, um, us, up, fs, o, c2);
                    if (ok & (1u << (2 * i))) us[o] = v;
                } else {
                    const double v = rb3_relax<1>(st, w, um, us, up, fs, o, c2);
                    if (ok & (1u << (2 * i + 1))) us[o + 1] = v;
                }
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
        }
        // ---- black(s-1) + store of the finished plane: tile rows 2 .. RY+1 (y = yb .. yb+RY-1) ----
        const int q = s - 1;
        if (q >= zb && q < ze && black_warp) {
            const double* const uqm = su + e_m2;
            const double* const fq = sf + e_f1;
            const bool q_in = q >= 1 && q <= n - 2;
            const int c0 = (par_red + s + 1) & 1;       // black row i lies one row above red row i, in plane s-1: its
                                                        // black node has the other component
#pragma unroll
            for (int i = 0; i < RW; ++i) {
                const int o = o_blk + i * RB3_W;
                double2 c2 = *reinterpret_cast<const double2*>(um + o);
                if ((c0 ^ (i & 1)) == 0) {
                    const double v = rb3_relax<0>(st, w, uqm, um, us, fq, o, c2);
                    if (q_in && (ok & (1u << (8 + 2 * i)))) c2.x = v;
                } else {
                    const double v = rb3_relax<1>(st, w, uqm, um, us, fq, o, c2);
                    if (q_in && (ok & (1u << (9 + 2 * i)))) c2.y = v;
                }
                double* o_g = optr + i * pitch;
                if (ok & (1u << (16 + i))) st2(o_g, c2);
                else if (ok & (1u << (20 + i))) *o_g = c2.x;
            }
        }
        optr += plane;
        e_m2 = e_m; e_m = e_s; e_s = e_p;
        e_p += U_SLOT; b_u += 8;
        if (e_p == RB3_NS * U_SLOT) { e_p = 0; b_u = bar_a; par_u ^= 1; }
        e_f1 = e_f;
        e_f += F_SLOT; b_f += 8;
        if (e_f == RB3_NF * F_SLOT) { e_f = 0; b_f = bar_a + 8 * RB3_NS; par_f ^= 1; }
    }
}

}  // namespace mg
