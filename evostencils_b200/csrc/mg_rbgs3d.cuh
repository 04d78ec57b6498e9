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
#include "mg_fast3d.cuh"

namespace mg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
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
constexpr int RB3_NS = 6;      // ring of u planes
constexpr int RB3_NF = 5;      // ring of f planes

template <int RY>
constexpr size_t rbgs3d_smem_bytes() {
    return (size_t)(RB3_NS * (RY + 4) + RB3_NF * (RY + 2)) * RB3_W * sizeof(double) + 16 * sizeof(unsigned long long);
}

// out = RB-GS sweep of u (red = (x+y+z) even first, then black), whole grid (g.nz == g.n, g.z0 == 0).
// grid = (ceil(n / 60), ceil(n / RY), ceil(n / zchunk)), block = 32 * (RY + 2).
template <int RY>
__global__ void __launch_bounds__(32 * (RY + 2), 3) k_rbgs3d(Geom g, Star3D st, const double* __restrict__ u,
                                                          const double* __restrict__ f, double* __restrict__ out, double w,
                                                          int zchunk) {
    constexpr int UR = RY + 4, FR = RY + 2;
    constexpr uint32_t U_BYTES = UR * RB3_W * 8, F_BYTES = FR * RB3_W * 8;
    extern __shared__ __align__(128) unsigned char rb3_smem[];
    double* su = reinterpret_cast<double*>(rb3_smem);                       // [NS][UR][64]
    double* sf = su + RB3_NS * UR * RB3_W;                                  // [NF][FR][64]
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sf + RB3_NF * FR * RB3_W);   // [NS] u, [NF] f
    pdl_sync();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = g.n;
    const long long pitch = g.pitch, plane = g.plane;
    const int xb = blockIdx.x * RB3_TW, yb = blockIdx.y * RY, zb = blockIdx.z * zchunk;
    const int ze = imin(zb + zchunk, n);                                    // output planes zb .. ze-1
    const int p0 = zb - 2, fp0 = zb - 1;                                    // first staged plane of u / f
    const uint32_t su_a = smem_u32(su), sf_a = smem_u32(sf), bar_a = smem_u32(bars);

    // stage plane p of u (rows yb-2 .. yb+RY+1) / of f (rows yb-1 .. yb+RY); plane and row indices are
    // clamped into the grid (clamped copies are only ever read for nodes that are not updated)
    auto stage_u = [&](int p) {
        const int slot = (p - p0) % RB3_NS;
        const uint32_t bar = bar_a + 8 * slot;
        if (lane == 0) mbar_arrive_expect_tx(bar, U_BYTES);
        __syncwarp();
        if (lane < UR) {
            const int y = imin(imax(yb - 2 + lane, 0), n - 1);
            bulk_g2s(su_a + (uint32_t)((slot * UR + lane) * RB3_W * 8),
                     u + (long long)imin(imax(p, 0), n - 1) * plane + (long long)y * pitch + (xb - 2), RB3_W * 8, bar);
        }
    };
    auto stage_f = [&](int p) {
        const int slot = (p - fp0) % RB3_NF;
        const uint32_t bar = bar_a + 8 * (RB3_NS + slot);
        if (lane == 0) mbar_arrive_expect_tx(bar, F_BYTES);
        __syncwarp();
        if (lane < FR) {
            const int y = imin(imax(yb - 1 + lane, 0), n - 1);
            bulk_g2s(sf_a + (uint32_t)((slot * FR + lane) * RB3_W * 8),
                     f + (long long)imin(imax(p, 0), n - 1) * plane + (long long)y * pitch + (xb - 2), RB3_W * 8, bar);
        }
    };
    auto wait_u = [&](int p) { mbar_wait(bar_a + 8 * ((p - p0) % RB3_NS), ((p - p0) / RB3_NS) & 1); };
    auto wait_f = [&](int p) { mbar_wait(bar_a + 8 * (RB3_NS + (p - fp0) % RB3_NF), ((p - fp0) / RB3_NF) & 1); };
    auto uslot = [&](int p) { return su + ((p - p0) % RB3_NS) * (UR * RB3_W); };
    auto fslot = [&](int p) { return sf + ((p - fp0) % RB3_NF) * (FR * RB3_W); };

    if (threadIdx.x == 0) {
        for (int i = 0; i < RB3_NS + RB3_NF; ++i) mbar_init(bar_a + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == RY + 1) {
        for (int i = 0; i < RB3_NS; ++i)
            if (p0 + i <= ze + 1) stage_u(p0 + i);
        for (int i = 0; i < RB3_NF; ++i)
            if (fp0 + i <= ze) stage_f(fp0 + i);
    }
    wait_u(p0);
    wait_u(p0 + 1);

    // u + w (f - A u) at component c of this lane's pair in tile row `lr` of a plane (um / uq / up: the
    // staged planes below, at and above it; fq: its right-hand side)
    auto relax = [&](const double* um, const double* uq, const double* up, const double* fq, int lr, int c,
                     const double2 c2) -> double {
        const double left = __shfl_up_sync(0xffffffffu, c2.y, 1), right = __shfl_down_sync(0xffffffffu, c2.x, 1);
        const double cen = c ? c2.y : c2.x;
        const double wv = c ? c2.x : left, ev = c ? right : c2.y;
        const int o = lr * RB3_W + 2 * lane + c;
        const double sv = uq[o - RB3_W], nv = uq[o + RB3_W];
        const double zm = um[o], zp = up[o];
        const double fv = fq[o - RB3_W];
        const double r = fv - star3_at(st, zm, sv, wv, cen, ev, nv, zp);
        return cen + w * r;
    };

    for (int s = zb - 1; s <= ze; ++s) {
        wait_u(s + 1);
        wait_f(s);
        // ---- red(s): tile rows 1 .. RY+2 (y = yb-1 .. yb+RY), one warp each ----
        if (s >= 1 && s <= n - 2) {
            const int lr = warp + 1, y = yb - 2 + lr;
            const int c = (y + s) & 1;                                       // xb is even: red = (col + y + s) even
            double* us = uslot(s);
            double2* row = reinterpret_cast<double2*>(us + lr * RB3_W) + lane;
            const double2 c2 = *row;
            const double v = relax(uslot(s - 1), us, uslot(s + 1), fslot(s), lr, c, c2);
            const int x = xb - 2 + 2 * lane + c;
            if (y >= 1 && y <= n - 2 && x >= imax(1, xb - 1) && x <= imin(n - 2, xb + RB3_TW))
                reinterpret_cast<double*>(row)[c] = v;
        }
        fence_proxy_async();
        __syncthreads();
        // planes <= s-3 of u and <= s-2 of f are dead now: refill their slots
        if (warp == RY + 1) {
            if (s - 3 >= p0 && s - 3 + RB3_NS <= ze + 1) stage_u(s - 3 + RB3_NS);
            if (s - 2 >= fp0 && s - 2 + RB3_NF <= ze) stage_f(s - 2 + RB3_NF);
        }
        // ---- black(s-1) + store of the finished plane: tile rows 2 .. RY+1 (y = yb .. yb+RY-1) ----
        const int q = s - 1;
        if (q >= zb && q < ze && warp < RY) {
            const int lr = warp + 2, y = yb + warp;
            const int c = (y + q + 1) & 1;
            const double* uq = uslot(q);
            double2 c2 = *(reinterpret_cast<const double2*>(uq + lr * RB3_W) + lane);
            const double v = relax(uslot(q - 1), uq, uslot(q + 1), fslot(q), lr, c, c2);
            const int x0 = xb - 2 + 2 * lane, x = x0 + c;
            if (q >= 1 && q <= n - 2 && y >= 1 && y <= n - 2 && x >= 1 && x <= n - 2) {
                if (c) c2.y = v; else c2.x = v;
            }
            if (lane >= 1 && lane <= RB3_TW / 2 && y <= n - 1) {
                double* o = out + (long long)q * plane + (long long)y * pitch + x0;
                if (x0 + 1 <= n - 1) st2(o, c2);
                else if (x0 <= n - 1) *o = c2.x;
            }
        }
    }
}

}  // namespace mg
