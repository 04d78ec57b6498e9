// mg_varcoef.cuh -- specialised 2D kernel for the variable-coefficient Poisson operator
// -div(c grad u), c = exp(kappa * (x - x^2) * (y - y^2))   (stencils/gallery.py:87-131).
//
// Same shape as k_star2d (mg_fast2d.cuh): a thread owns two adjacent columns, a CTA a strip of 2*NT
// columns x a band of R rows, every load of the band is issued before the first use.  Nothing but u and f
// is read (24 B/DoF per sweep as for constant coefficients): the face coefficients are recomputed,
// and a face is shared by the two nodes it separates, so a band of R rows costs a thread 3 R x-face
// and 2 (R + 1) y-face evaluations of exp() for its 2 R nodes (2.75 per node at R = 4 instead of 4).
// The face positions (i +- 1/2) h are exact in fp64, so "east face of node i" and "west face of node
// i + 1" are the same argument and give the same value: the arithmetic per node is that of vc_apply
// (mg_kernels_generic.cuh) and of the oracle, operand for operand.
//   MODE STAR_JACOBI, colour < 0 : out = u + (w / a00) (f - A u), out of place (boundary copied)
//   MODE STAR_JACOBI, colour >= 0: the same on the nodes (x + y) % 2 == colour, IN PLACE (out == u): a node of
//                                  one colour reads only nodes of the other, which the pass leaves unchanged
//   MODE STAR_RESIDUAL           : out = f - A u
//   MODE STAR_NORM               : partial[cta] = sum (f - A u)^2
#pragma once
#include "mg_fast2d.cuh"

namespace mg {

struct VC2D {
    double kappa, h, ihh;      // ihh = 1 / (h * h), a power of two: x * ihh == x / (h * h) bit for bit
};

__device__ __forceinline__ double vc_node(const VC2D& s, double cS, double cW, double cE, double cN, double S, double W,
                                          double Cc, double E, double N, double& diag) {
    diag = ((cE + cW) * s.ihh) + ((cN + cS) * s.ihh);
    double a = ((-cS) * s.ihh) * S;
    a = a + ((-cW) * s.ihh) * W;
    a = a + diag * Cc;
    a = a + ((-cE) * s.ihh) * E;
    a = a + ((-cN) * s.ihh) * N;
    return a;
}

#ifndef MG_VC_MIN_CTAS
#define MG_VC_MIN_CTAS 5
#endif
template <int MODE, int NT, int R>
__global__ void __launch_bounds__(NT, MG_VC_MIN_CTAS) k_vcstar2d(Geom g, VC2D s, const double* u, const double* __restrict__ f,
                                                  double* out, double w, int colour, double* __restrict__ partial) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int x0 = (blockIdx.x * NT + threadIdx.x) * 2;       // 0 <= x0 <= n-3, even
    const int n = g.n;
    const long long pitch = g.pitch;
    const int y0 = 1 + blockIdx.y * R;
    const bool in0 = x0 >= 1;
    const bool last = x0 == n - 3;
    const bool edge = (lane == 0 && x0 > 0) || lane == 31;
    const int xe = lane == 0 ? x0 - 1 : x0 + 2;
    double2 uu[R + 2], ff[R];
    double ue[R];
#pragma unroll
    for (int i = 0; i < R + 2; ++i) uu[i] = ldf2(u + (long long)imin(y0 - 1 + i, n) * pitch + x0);
#pragma unroll
    for (int i = 0; i < R; ++i) ff[i] = ldf2(f + (long long)imin(y0 + i, n) * pitch + x0);
#pragma unroll
    for (int i = 0; i < R; ++i) {
        ue[i] = 0.0;
        if (edge) ue[i] = ldf1(u + (long long)imin(y0 + i, n) * pitch + xe);
    }
    __syncthreads();      // scheduling fence (see k_star2d)
    const bool all = colour < 0;
    if (MODE == STAR_JACOBI && all && blockIdx.y == 0) {       // Dirichlet row 0
        st2(out + x0, uu[0]);
        if (last) out[n - 1] = u[n - 1];
    }
    const double h = s.h, hf = 0.5 * s.h, kap = s.kappa;
    const double px0 = x0 * h, px1 = (x0 + 1) * h;
    const double ax0 = vc_axis(px0), ax1 = vc_axis(px1);
    const double axw = vc_axis(px0 - hf), axm = vc_axis(px0 + hf), axe = vc_axis(px1 + hf);
    double acc = 0.0;
    if (all) {
        // face between the rows y-1 and y, carried from row to row
        const double ayf0 = vc_axis((y0 - 1) * h + hf);
        double cS0 = exp(kap * (ax0 * ayf0)), cS1 = exp(kap * (ax1 * ayf0));
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int y = y0 + i;
            const double2 um = uu[i], uc = uu[i + 1], un = uu[i + 2];
            double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
            double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
            if (lane == 0) left = ue[i];
            if (lane == 31) right = ue[i];
            if (y <= n - 2) {
                const double py = y * h;
                const double ayc = vc_axis(py), ayn = vc_axis(py + hf);
                const double cW = exp(kap * (axw * ayc)), cM = exp(kap * (axm * ayc)), cE = exp(kap * (axe * ayc));
                const double cN0 = exp(kap * (ax0 * ayn)), cN1 = exp(kap * (ax1 * ayn));
                double d0, d1;
                const double r0 = ff[i].x - vc_node(s, cS0, cW, cM, cN0, um.x, left, uc.x, uc.y, un.x, d0);
                const double r1 = ff[i].y - vc_node(s, cS1, cM, cE, cN1, um.y, uc.x, uc.y, right, un.y, d1);
                cS0 = cN0;
                cS1 = cN1;
                const long long base = (long long)y * pitch + x0;
                if (MODE == STAR_JACOBI) {
                    double2 o;
                    o.x = in0 ? uc.x + (w / d0) * r0 : uc.x;
                    o.y = uc.y + (w / d1) * r1;
                    st2(out + base, o);
                    if (last) out[base + 2] = right;               // Dirichlet column n-1
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
        if (MODE == STAR_JACOBI && y0 + R >= n - 1) {          // this band reaches the Dirichlet row n-1
#pragma unroll
            for (int i = 1; i < R + 2; ++i)
                if (y0 - 1 + i == n - 1) st2(out + (long long)(n - 1) * pitch + x0, uu[i]);
            if (last) out[(long long)(n - 1) * pitch + n - 1] = u[(long long)(n - 1) * pitch + n - 1];
        }
    } else if (MODE == STAR_JACOBI) {
        // one colour, in place: per row exactly one of the thread's two columns is active
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int y = y0 + i;
            const double2 um = uu[i], uc = uu[i + 1], un = uu[i + 2];
            double left = __shfl_up_sync(0xffffffffu, uc.y, 1);
            double right = __shfl_down_sync(0xffffffffu, uc.x, 1);
            if (lane == 0) left = ue[i];
            if (lane == 31) right = ue[i];
            const bool first = (y & 1) == colour;                 // column x0 (even) has the parity of y
            if (y <= n - 2 && (in0 || !first)) {
                const double py = y * h;
                const double ayc = vc_axis(py), ayn = vc_axis(py + hf), ays = vc_axis(py - hf);
                const double axc = first ? ax0 : ax1;
                const double cW = exp(kap * ((first ? axw : axm) * ayc)), cE = exp(kap * ((first ? axm : axe) * ayc));
                const double cN = exp(kap * (axc * ayn)), cS = exp(kap * (axc * ays));
                double d;
                const double S = first ? um.x : um.y, N = first ? un.x : un.y, Cc = first ? uc.x : uc.y;
                const double W = first ? left : uc.x, E = first ? uc.y : right;
                const double r = (first ? ff[i].x : ff[i].y) - vc_node(s, cS, cW, cE, cN, S, W, Cc, E, N, d);
                out[(long long)y * pitch + x0 + (first ? 0 : 1)] = Cc + (w / d) * r;
            }
        }
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

}  // namespace mg
