// mg_fas.cuh -- stage kernels of the nonlinear full-approximation-scheme example
// (/root/reference/example_problems/FAS_2D_Basic/FAS_2D_Basic.exa4):
// -Laplace(u) + gam u e^u = f in 2D, scalar real fp64, Dirichlet 0.
//
// The operator of the level is the 5-point star the workspace holds (Level::star_*): centre, E, W, N, S
// in the order the .exa4 declares the stencil (:18-24), then the nonlinear term gam e^u u (:25-27),
// separate multiply and add.  One thread per node, x fastest; the grids of this example (65^2 ..
// 1025^2) are latency-bound, so these are plain node kernels driven op by op (no graph).
#pragma once
#include "mg_exec.cuh"

namespace mg {

struct FasStar {
    double cc, cx, cy, gam;
};

// (L u + gam e^u u) at an interior node
__device__ __forceinline__ double fas_apply(const FasStar& s, const double* __restrict__ u, long long i, long long pitch) {
    const double c = u[i];
    double a = s.cc * c;
    a = a + s.cx * u[i + 1];
    a = a + s.cx * u[i - 1];
    a = a + s.cy * u[i + pitch];
    a = a + s.cy * u[i - pitch];
    return a + (s.gam * exp(c)) * c;
}

// One sweep of the smoother `traverse_graph` emits (exastencils_FAS.py:186-246) or the template's
// Smoother function (.exa4:65-75 = one Newton step, colour < 0):
//   newton_steps >= 1: Jacobi-Newton, that many in-place updates of the centre value per node
//                      u += w (f - (L u + gam e^u u)) / (diag(L) + gam (u + 1) e^u), neighbours fixed;
//   newton_steps == 0: Jacobi-Picard, one update with the denominator diag(L);
//   colour < 0: Jacobi (reads u, writes out, boundary copied); colour 0 / 1: in-place pass (out == u) over
//   the interior nodes with (x + y) % 2 == colour, whose neighbours all have the other colour.
__global__ void k_fas_smooth(Geom g, FasStar s, const double* u, const double* __restrict__ f, double* out, double w,
                             int newton_steps, int colour) {
    int x, y, z;
    if (!node_coords<2>(g, x, y, z)) return;
    const long long i = node_index<2>(g, x, y, 0);
    const bool interior = is_interior<2>(g, x, y, 0);
    if (colour >= 0 && (!interior || ((x + y) & 1) != colour)) return;
    double v = u[i];
    if (interior) {
        const double e = s.cx * u[i + 1], wt = s.cx * u[i - 1], nt = s.cy * u[i + g.pitch], st = s.cy * u[i - g.pitch];
        const double fi = f[i];
        const int steps = newton_steps > 0 ? newton_steps : 1;
        for (int k = 0; k < steps; ++k) {
            double a = s.cc * v;
            a = a + e;
            a = a + wt;
            a = a + nt;
            a = a + st;
            const double ev = exp(v);
            const double num = fi - (a + (s.gam * ev) * v);
            const double den = newton_steps > 0 ? s.cc + s.gam * (v + 1.0) * ev : s.cc;
            v = v + w * (num / den);
        }
    }
    out[i] = v;
}

// Residual = RHS - (L u + gam e^u u), boundary 0   (.exa4:77-82)
__global__ void k_fas_residual(Geom g, FasStar s, const double* __restrict__ u, const double* __restrict__ f,
                               double* __restrict__ r) {
    int x, y, z;
    if (!node_coords<2>(g, x, y, z)) return;
    const long long i = node_index<2>(g, x, y, 0);
    r[i] = is_interior<2>(g, x, y, 0) ? f[i] - fas_apply(s, u, i, g.pitch) : 0.0;
}

// Approximation@coarse = R Solution@fine ; Solution@coarse = Approximation   (.exa4:219-222); boundary 0
__global__ void k_fas_restrict_solution(Geom gc, Geom gf, const double* __restrict__ fine, double* __restrict__ approx,
                                        double* __restrict__ sol) {
    int x, y, z;
    if (!node_coords<2>(gc, x, y, z)) return;
    const long long i = node_index<2>(gc, x, y, 0);
    const double v = is_interior<2>(gc, x, y, 0) ? restrict_at<double, 2>(fine, gf, x, y, 0) : 0.0;
    approx[i] = v;
    sol[i] = v;
}

// RHS@coarse = R Residual@fine + (L approx + gam e^approx approx)   (.exa4:225-227); interior only
__global__ void k_fas_coarse_rhs(Geom gc, Geom gf, FasStar sc, const double* __restrict__ res_fine,
                                 const double* __restrict__ approx, double* __restrict__ rhs) {
    int x, y, z;
    if (!node_coords<2>(gc, x, y, z)) return;
    if (!is_interior<2>(gc, x, y, 0)) return;
    const long long i = node_index<2>(gc, x, y, 0);
    rhs[i] = restrict_at<double, 2>(res_fine, gf, x, y, 0) + fas_apply(sc, approx, i, gc.pitch);
}

// Solution -= Approximation   (.exa4:247-249)
__global__ void k_fas_subtract(Geom g, double* __restrict__ u, const double* __restrict__ approx) {
    int x, y, z;
    if (!node_coords<2>(g, x, y, z)) return;
    const long long i = node_index<2>(g, x, y, 0);
    u[i] = u[i] - approx[i];
}

// row_sums[y] = sum over the interior nodes of row y of (f - (L u + gam e^u u))^2: 128 lane-strided partial
// sums in x order, then a fixed butterfly (deterministic); the host adds the rows in order
__global__ void __launch_bounds__(128) k_fas_norm_rows(Geom g, FasStar s, const double* __restrict__ u,
                                                       const double* __restrict__ f, double* __restrict__ row_sums) {
    const int y = blockIdx.x + 1;
    double acc = 0.0;
    for (int x = 1 + threadIdx.x; x <= g.n - 2; x += 128) {
        const long long i = node_index<2>(g, x, y, 0);
        const double r = f[i] - fas_apply(s, u, i, g.pitch);
        acc = acc + r * r;
    }
    __shared__ double part[128];
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int k = 64; k >= 1; k >>= 1) {
        if (threadIdx.x < k) part[threadIdx.x] = part[threadIdx.x] + part[threadIdx.x + k];
        __syncthreads();
    }
    if (threadIdx.x == 0) row_sums[blockIdx.x] = part[0];
}

}  // namespace mg
