// mg_fast3d.cuh -- specialised 3D kernels for scalar real 7-point star stencils (3D Poisson), whole
// grids and z-slabs (Geom::nz, z0).
//
// 2.5-D blocking: a thread owns two adjacent x columns (one aligned double2 per row), a CTA of NT
// threads a strip of 2*NT columns x RY rows, and marches over a chunk of z planes keeping the planes
// z-1, z, z+1 in registers.  Every plane is loaded once per CTA (RY+2 rows: the y halo is served by
// L2), the plane z+2 and the right-hand side of plane z+1 are requested one step ahead (volatile
// loads + bar.sync as a scheduling fence, see mg_fast2d.cuh), x neighbours come from the adjacent
// lane by warp shuffle.  Operand order z-, y-, x-, c, x+, y+, z+ with separate multiply and add:
// bit-identical to the generic kernels and the oracle.
#pragma once
#include "mg_fast2d.cuh"

namespace mg {

struct Star3D {
    double cz, cy, cx, cc;
};

__device__ __forceinline__ double star3_at(const Star3D& s, double Zm, double S, double W, double Cc, double E, double N,
                                           double Zp) {
    double a = s.cz * Zm;
    a = a + s.cy * S;
    a = a + s.cx * W;
    a = a + s.cc * Cc;
    a = a + s.cx * E;
    a = a + s.cy * N;
    a = a + s.cz * Zp;
    return a;
}

enum { STAR3_SMOOTH = 0, STAR3_RESIDUAL = 1, STAR3_NORM = 2 };

// MODE 0: out = u + w (f - A u) on the interior nodes of the colour (colour < 0: all), out = u elsewhere
//         (boundary rows / columns / planes included: the sweep is out of place)
// MODE 1: out = f - A u on interior nodes
// MODE 2: partial[cta] = sum (f - A u)^2
template <int MODE, int NT, int RY>
__global__ void __launch_bounds__(NT) k_star3d(Geom g, Star3D s, const double* __restrict__ u,
                                                const double* __restrict__ f, double* __restrict__ out, double w,
                                                int colour, int zchunk, double* __restrict__ partial) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // 0 <= x0 <= n-3
    const int n = g.n, nz = g.nz;
    const long long pitch = g.pitch, plane = g.plane;
    const int y0 = 1 + blockIdx.y * RY;
    const int zb = 1 + blockIdx.z * zchunk;                    // local planes zb .. ze-1
    const int ze = imin(zb + zchunk, nz - 1);
    const bool in0 = x0 >= 1;
    const bool last = x0 == n - 3;
    const bool edge = (lane == 0 && x0 > 0) || lane == 31;
    const int xe = lane == 0 ? x0 - 1 : x0 + 2;
    const bool first_band = blockIdx.y == 0, last_band = y0 + RY >= n - 1;

    // rows y0-1 .. y0+RY of plane z (row index clamped to n = first row of the next plane / guard row;
    // plane index clamped to the last plane: such values are never used)
    auto load_plane = [&](int z, double2 (&p)[RY + 2], double (&e)[RY]) {
        const double* base = u + (long long)imin(z, nz - 1) * plane + x0;
#pragma unroll
        for (int i = 0; i < RY + 2; ++i) p[i] = ldf2(base + (long long)imin(y0 - 1 + i, n) * pitch);
#pragma unroll
        for (int i = 0; i < RY; ++i) {
            e[i] = 0.0;
            if (edge) e[i] = ldf1(u + (long long)imin(z, nz - 1) * plane + (long long)imin(y0 + i, n) * pitch + xe);
        }
    };
    auto load_f = [&](int z, double2 (&p)[RY]) {
        const double* base = f + (long long)imin(z, nz - 1) * plane + x0;
#pragma unroll
        for (int i = 0; i < RY; ++i) p[i] = ldf2(base + (long long)imin(y0 + i, n) * pitch);
    };
    // out-of-place sweep: copy a whole boundary / ghost plane segment owned by this CTA
    auto copy_plane = [&](int z, const double2 (&p)[RY + 2], const double (&e)[RY]) {
        double* o = out + (long long)z * plane + x0;
#pragma unroll
        for (int i = 0; i < RY + 2; ++i) {
            const int y = y0 - 1 + i;
            const bool mine = (i >= 1 && i <= RY && y <= n - 2) || (i == 0 && first_band) || (y == n - 1);
            if (mine && y <= n - 1) st2(o + (long long)y * pitch, p[i]);
        }
        if (last) {
#pragma unroll
            for (int i = 0; i < RY; ++i)
                if (y0 + i <= n - 2) o[(long long)(y0 + i) * pitch + 2] = e[i];
            if (first_band) o[2] = u[(long long)z * plane + n - 1];
            if (last_band) o[(long long)(n - 1) * pitch + 2] = u[(long long)z * plane + (long long)(n - 1) * pitch + n - 1];
        }
    };

    double2 pm[RY], pc[RY + 2], pn[RY + 2], pnn[RY + 2], fc[RY], fn[RY];
    double ec[RY], en[RY], enn[RY];
    {
        double2 t[RY + 2];
        double et[RY];
        load_plane(zb - 1, t, et);
        load_plane(zb, pc, ec);
        load_plane(zb + 1, pn, en);
        load_f(zb, fc);
        __syncthreads();
        if (MODE == STAR3_SMOOTH && blockIdx.z == 0) copy_plane(0, t, et);        // local plane 0: boundary or ghost
#pragma unroll
        for (int i = 0; i < RY; ++i) pm[i] = t[i + 1];
    }
    double acc = 0.0;
    for (int z = zb; z < ze; ++z) {
        load_plane(z + 2, pnn, enn);
        load_f(z + 1, fn);
        __syncthreads();                                       // scheduling fence: prefetch stays ahead of the compute
        const int zg = z + g.z0;
        const bool zin = zg >= 1 && zg <= n - 2;               // ghost planes are never the centre (z in [1, nz-2])
#pragma unroll
        for (int i = 0; i < RY; ++i) {
            const int y = y0 + i;
            const double2 uc = pc[i + 1];
            double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
            double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
            if (lane == 0) left = ec[i];
            if (lane == 31) right = ec[i];
            if (y <= n - 2) {
                const long long o = (long long)z * plane + (long long)y * pitch + x0;
                const double r0 = fc[i].x - star3_at(s, pm[i].x, pc[i].x, left, uc.x, uc.y, pc[i + 2].x, pn[i + 1].x);
                const double r1 = fc[i].y - star3_at(s, pm[i].y, pc[i].y, uc.x, uc.y, right, pc[i + 2].y, pn[i + 1].y);
                if (MODE == STAR3_SMOOTH) {
                    const bool a0 = zin && in0 && (colour < 0 || ((x0 + y + zg) & 1) == colour);
                    const bool a1 = zin && (colour < 0 || ((x0 + 1 + y + zg) & 1) == colour);
                    double2 v;
                    v.x = a0 ? uc.x + w * r0 : uc.x;
                    v.y = a1 ? uc.y + w * r1 : uc.y;
                    st2(out + o, v);
                    if (last) out[o + 2] = right;              // Dirichlet column n-1
                } else if (MODE == STAR3_RESIDUAL) {
                    if (zin) {
                        if (in0) {
                            double2 v;
                            v.x = r0; v.y = r1;
                            st2(out + o, v);
                        } else {
                            out[o + 1] = r1;
                        }
                    }
                } else if (zin) {
                    if (in0) acc = acc + r0 * r0;
                    acc = acc + r1 * r1;
                }
            }
        }
        if (MODE == STAR3_SMOOTH && (first_band || last_band)) {   // Dirichlet rows 0 and n-1 of this plane
            double* o = out + (long long)z * plane + x0;
            if (first_band) {
                st2(o, pc[0]);
                if (last) o[2] = u[(long long)z * plane + n - 1];
            }
            if (last_band) {
#pragma unroll
                for (int i = 1; i < RY + 2; ++i)
                    if (y0 - 1 + i == n - 1) st2(o + (long long)(n - 1) * pitch, pc[i]);
                if (last) o[(long long)(n - 1) * pitch + 2] = u[(long long)z * plane + (long long)(n - 1) * pitch + n - 1];
            }
        }
#pragma unroll
        for (int i = 0; i < RY; ++i) { pm[i] = pc[i + 1]; fc[i] = fn[i]; ec[i] = en[i]; en[i] = enn[i]; }
#pragma unroll
        for (int i = 0; i < RY + 2; ++i) { pc[i] = pn[i]; pn[i] = pnn[i]; }
    }
    if (MODE == STAR3_SMOOTH && ze == nz - 1) copy_plane(nz - 1, pc, ec);       // local plane nz-1 (pc/ec hold it now)
    if (MODE == STAR3_NORM) {
        __shared__ double s32[32];
        double v = warp_sum<double>(acc);
        if (lane == 0) s32[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = s32[0];
            for (int i = 1; i < NT / 32; ++i) t = t + s32[i];
            partial[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = t;
        }
    }
}

// ---- fused prolongation + correction: u += omega * (P e), in place --------------------------------------
// Thread = two adjacent fine columns (x0 even, x0+1 odd) of RY consecutive fine rows of one plane.
// Coarse values at column X = x0/2 are loaded by the thread, those at X+1 come from the next lane
// (lane 31 loads them); every coarse value is shared by 8 fine nodes and served by L1/L2.
// Summation order of prolong_at: coarse planes outer, rows, columns inner; then the power-of-two scale.
template <int NT, int RY>
__global__ void __launch_bounds__(NT) k_prolong3d(Geom gf, Geom gc, const double* __restrict__ e, double* __restrict__ u,
                                                   double omega) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // even fine column, <= n-3
    const int X = x0 >> 1;
    const int n = gf.n;
    const int y0 = 1 + blockIdx.y * RY;                         // odd
    const int z = 1 + blockIdx.z;                               // local fine plane 1 .. nz-2
    const int zg = z + gf.z0;
    if (zg < 1 || zg > n - 2) return;
    const int pz = zg & 1;
    const int Zl = (zg >> 1) - gc.z0;                           // local coarse plane
    const int Ylo = y0 >> 1;
    double a[2][RY / 2 + 1], b[2][RY / 2 + 1];
    double2 uu[RY];
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int j = 0; j < RY / 2 + 1; ++j) {
            const int Y = imin(Ylo + j, gc.n - 1);
            const double* p = e + (long long)(Zl + (k & pz)) * gc.plane + (long long)Y * gc.pitch + X;
            a[k][j] = ldf1(p);
            b[k][j] = 0.0;
            if (lane == 31) b[k][j] = ldf1(p + 1);
        }
#pragma unroll
    for (int i = 0; i < RY; ++i)
        uu[i] = ldf2(u + (long long)z * gf.plane + (long long)imin(y0 + i, n - 1) * gf.pitch + x0);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int j = 0; j < RY / 2 + 1; ++j) {
            const double nb = __shfl_down_sync(0xffffffffu, a[k][j], 1);
            if (lane != 31) b[k][j] = nb;
        }
#pragma unroll
    for (int i = 0; i < RY; ++i) {
        const int y = y0 + i;
        if (y > n - 2) continue;
        const int py = y & 1;
        const int j = (y >> 1) - Ylo;
        // even column (px = 0): a only ; odd column (px = 1): a, b alternating
        double s0 = a[0][j], s1 = a[0][j] + b[0][j];
        if (py) { s0 = s0 + a[0][j + 1]; s1 = s1 + a[0][j + 1]; s1 = s1 + b[0][j + 1]; }
        if (pz) {
            s0 = s0 + a[1][j]; s1 = s1 + a[1][j]; s1 = s1 + b[1][j];
            if (py) { s0 = s0 + a[1][j + 1]; s1 = s1 + a[1][j + 1]; s1 = s1 + b[1][j + 1]; }
        }
        const int k0 = py + pz, k1 = 1 + py + pz;
        if (k0 > 0) s0 = (k0 == 1 ? 0.5 : 0.25) * s0;
        s1 = (k1 == 1 ? 0.5 : (k1 == 2 ? 0.25 : 0.125)) * s1;
        double2 v = uu[i];
        if (x0 >= 1) v.x = v.x + omega * s0;
        v.y = v.y + omega * s1;
        st2(u + (long long)z * gf.plane + (long long)y * gf.pitch + x0, v);
    }
}

// ---- the same stage, two fine planes per thread (round 2) -----------------------------------------------------
// ncu of k_prolong3d<128,8> at 513^3 (profiles/r2/ncu_3d_513.md): DRAM bytes = algorithmic bytes, but the L1 pipe
// is 84 % busy and 65 % of the stall samples wait on it: 26 load/store instructions per 16 nodes, ten of them
// 8-byte loads of coarse values, half of which are duplicates on even planes.  The fine planes 2Z+1 and 2Z+2
// need the coarse planes Z and Z+1 only, so a thread that owns its column pair in BOTH planes loads the coarse
// values once: 2 (RY/2+1) coarse loads per 4 RY nodes instead of per 2 RY.  Summation order of prolong_at
// (coarse planes outer, rows, columns inner, then the power-of-two scale): unchanged, bit-identical.
// grid.z = pairs of fine planes: pair p = global planes 2p+1 (odd) and 2p+2 (even), p0 = first pair of the slab.
template <int NT, int RY>
__global__ void __launch_bounds__(NT) k_prolong3d_zz(Geom gf, Geom gc, const double* __restrict__ e, double* __restrict__ u,
                                                      double omega, int p0) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // even fine column, <= n-3
    const int X = x0 >> 1;
    const int n = gf.n;
    const int y0 = 1 + blockIdx.y * RY;                         // odd
    const int Z = p0 + blockIdx.z;                              // global coarse plane below the pair
    const int zo = 2 * Z + 1 - gf.z0, ze = zo + 1;              // local fine planes of the pair (odd, even global index)
    // a plane of the pair is updated if it is an owned plane of the slab and an interior plane of the grid
    const bool do_o = zo >= 1 && zo <= gf.nz - 2 && 2 * Z + 1 <= n - 2;
    const bool do_e = ze >= 1 && ze <= gf.nz - 2 && 2 * Z + 2 <= n - 2;
    const int Zl = Z - gc.z0;                                   // local coarse plane
    const int Ylo = y0 >> 1;
    double a[2][RY / 2 + 1], b[2][RY / 2 + 1];
    double2 uo[RY], ue[RY];
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int j = 0; j < RY / 2 + 1; ++j) {
            const int Y = imin(Ylo + j, gc.n - 1);
            const int Zk = (k == 0 && !do_o) ? Zl + 1 : Zl + k;   // plane Z is not needed (and may not exist) without the odd plane
            const double* p = e + (long long)Zk * gc.plane + (long long)Y * gc.pitch + X;
            a[k][j] = ldf1(p);
            b[k][j] = 0.0;
            if (lane == 31) b[k][j] = ldf1(p + 1);
        }
#pragma unroll
    for (int i = 0; i < RY; ++i) {
        const long long row = (long long)imin(y0 + i, n - 1) * gf.pitch + x0;
        uo[i] = make_double2(0.0, 0.0);
        ue[i] = make_double2(0.0, 0.0);
        if (do_o) uo[i] = ldf2(u + (long long)zo * gf.plane + row);
        if (do_e) ue[i] = ldf2(u + (long long)ze * gf.plane + row);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int j = 0; j < RY / 2 + 1; ++j) {
            const double nb = __shfl_down_sync(0xffffffffu, a[k][j], 1);
            if (lane != 31) b[k][j] = nb;
        }
#pragma unroll
    for (int i = 0; i < RY; ++i) {
        const int y = y0 + i;
        if (y > n - 2) continue;
        const int py = y & 1;
        const int j = (y >> 1) - Ylo;
        if (do_o) {                                             // pz = 1: coarse planes Z, Z+1
            double s0 = a[0][j], s1 = a[0][j] + b[0][j];
            if (py) { s0 = s0 + a[0][j + 1]; s1 = s1 + a[0][j + 1]; s1 = s1 + b[0][j + 1]; }
            s0 = s0 + a[1][j]; s1 = s1 + a[1][j]; s1 = s1 + b[1][j];
            if (py) { s0 = s0 + a[1][j + 1]; s1 = s1 + a[1][j + 1]; s1 = s1 + b[1][j + 1]; }
            s0 = (py ? 0.25 : 0.5) * s0;
            s1 = (py ? 0.125 : 0.25) * s1;
            double2 v = uo[i];
            if (x0 >= 1) v.x = v.x + omega * s0;
            v.y = v.y + omega * s1;
            st2(u + (long long)zo * gf.plane + (long long)y * gf.pitch + x0, v);
        }
        if (do_e) {                                             // pz = 0: coarse plane Z+1 alone
            double s0 = a[1][j], s1 = a[1][j] + b[1][j];
            if (py) { s0 = s0 + a[1][j + 1]; s1 = s1 + a[1][j + 1]; s1 = s1 + b[1][j + 1]; }
            if (py) s0 = 0.5 * s0;
            s1 = (py ? 0.25 : 0.5) * s1;
            double2 v = ue[i];
            if (x0 >= 1) v.x = v.x + omega * s0;
            v.y = v.y + omega * s1;
            st2(u + (long long)ze * gf.plane + (long long)y * gf.pitch + x0, v);
        }
    }
}

}  // namespace mg
