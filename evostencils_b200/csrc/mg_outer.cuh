// mg_outer.cuh -- preconditioned BiCGStab outer solver (Helmholtz configuration), statement by
// statement after example_problems/Helmholtz/2D_FD_Helmholtz_fromL3.exa3:144-200: the evolved
// multigrid cycle is the preconditioner (`u = 0; f = p; gen_mgCycle()`), A is the unshifted operator,
// `dot` has no conjugation, the stopping test is |curRes| < tol |initRes|.
// All Krylov scalars live on the device and are updated by a one-thread kernel, so one whole
// iteration (2 cycles, 2 operator applications, 5 dot products, 3 vector updates) is a single
// captured CUDA graph; the host reads back |curRes| once per iteration.
#pragma once
#include "mg_kernels_generic.cuh"

namespace mg {

enum { OV_X = 0, OV_R, OV_RHAT, OV_P, OV_AP, OV_S, OV_T, OV_H, OV_RHS, OV_COUNT };

template <typename T>
struct OuterScalars {
    T alpha, beta, rho, rho_new, omega, cur, init, d0, d1;
};

// out = A u on interior nodes (boundary of out stays 0)
template <typename T, int D>
__global__ void k_outer_apply(Geom g, const OperatorTable<T>* __restrict__ op, const T* u, T* out) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z) || !is_interior<D>(g, x, y, z)) return;
    long long idx = node_index<D>(g, x, y, z);
    FieldSet<T> fs;
    for (int i = 0; i < MG_MAX_FIELDS; ++i) fs.p[i] = const_cast<T*>(u);
    out[idx] = apply_row<T, 1>(op, fs, 0, idx);
}

// out = rhs - A u on interior nodes
template <typename T, int D>
__global__ void k_outer_residual(Geom g, const OperatorTable<T>* __restrict__ op, const T* u, const T* rhs, T* out) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z) || !is_interior<D>(g, x, y, z)) return;
    long long idx = node_index<D>(g, x, y, z);
    FieldSet<T> fs;
    for (int i = 0; i < MG_MAX_FIELDS; ++i) fs.p[i] = const_cast<T*>(u);
    out[idx] = t_sub(rhs[idx], apply_row<T, 1>(op, fs, 0, idx));
}

// partial[b] = sum over the interior rows owned by block b of a * b (no conjugation)
template <typename T, int D>
__global__ void __launch_bounds__(256) k_outer_dot(Geom g, const T* a, const T* b, T* __restrict__ partial) {
    pdl_sync();
    __shared__ T s8[8];
    const int rows = (D == 3) ? g.n * g.nz : g.n;
    T acc = t_zero(T());
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        int z = (D == 3) ? row / g.n : 0;
        int y = (D == 3) ? row - z * g.n : row;
        if (!is_interior<D>(g, 1, y, z)) continue;
        for (int x = 1 + threadIdx.x; x <= g.n - 2; x += 256) {
            long long idx = node_index<D>(g, x, y, z);
            acc = t_add(acc, t_mul(a[idx], b[idx]));
        }
    }
    T total = block_sum_256<T>(acc, s8);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

template <typename T>
__global__ void __launch_bounds__(256) k_outer_dot_final(const T* __restrict__ partial, int n, T* out) {
    pdl_sync();
    __shared__ T s8[8];
    T acc = t_zero(T());
    for (int i = threadIdx.x; i < n; i += 256) acc = t_add(acc, partial[i]);
    T total = block_sum_256<T>(acc, s8);
    if (threadIdx.x == 0) *out = total;
}

__device__ __forceinline__ double t_modulus(double a) { return fabs(a); }
__device__ __forceinline__ double t_modulus(cplx a) { return t_hypot(a.re, a.im); }
__device__ __forceinline__ double t_root(double a) { return sqrt(a); }
__device__ __forceinline__ cplx t_root(cplx z) {
    double r = t_hypot(z.re, z.im);
    double sr = sqrt(0.5 * (r + z.re));
    double si = sqrt(0.5 * (r - z.re));
    return make_cplx(sr, z.im < 0 ? -si : si);
}
template <typename T>
__device__ __forceinline__ T t_one() {
    T one = t_zero(T());
    *reinterpret_cast<double*>(&one) = 1.0;
    return one;
}

// stage 0: init (cur = init = sqrt(d0), scalars = 1); 1: rho/beta; 2: alpha; 3: omega; 4: cur = sqrt(d0)
template <typename T>
__global__ void k_outer_scalars(OuterScalars<T>* s, int stage, double* norm_out) {
    pdl_sync();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    switch (stage) {
        case 0:
            s->cur = t_root(s->d0);
            s->init = s->cur;
            s->alpha = s->beta = s->rho_new = s->omega = t_one<T>();
            *norm_out = t_modulus(s->cur);
            break;
        case 1:
            s->rho = s->rho_new;
            s->rho_new = s->d0;
            s->beta = t_mul(t_div(s->rho_new, s->rho), t_div(s->alpha, s->omega));
            break;
        case 2: s->alpha = t_div(s->rho_new, s->d0); break;
        case 3: s->omega = t_div(s->d0, s->d1); break;
        case 4:
            s->cur = t_root(s->d0);
            *norm_out = t_modulus(s->cur);
            break;
        default: break;
    }
}

// whole-array vector updates (all n^D nodes; the pitch padding is never touched)
// kind 1: p = r + beta (p - omega ap)
// kind 2: h = x + alpha u ; s = r - alpha ap
// kind 3: x = h + omega u ; r = s - omega t
// kind 4: dst = src (copy)
template <typename T, int D>
__global__ void k_outer_update(Geom g, int kind, const OuterScalars<T>* __restrict__ sc, T* v0, T* v1, const T* v2,
                               const T* v3, const T* v4, const T* v5) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z)) return;
    long long i = node_index<D>(g, x, y, z);
    if (kind == 1) {                   // v0 = p, v2 = r, v3 = ap
        v0[i] = t_add(v2[i], t_mul(sc->beta, t_sub(v0[i], t_mul(sc->omega, v3[i]))));
    } else if (kind == 2) {            // v0 = h, v1 = s, v2 = x, v3 = u, v4 = r, v5 = ap
        v0[i] = t_add(v2[i], t_mul(sc->alpha, v3[i]));
        v1[i] = t_sub(v4[i], t_mul(sc->alpha, v5[i]));
    } else if (kind == 3) {            // v0 = x, v1 = r, v2 = h, v3 = u, v4 = s, v5 = t
        v0[i] = t_add(v2[i], t_mul(sc->omega, v3[i]));
        v1[i] = t_sub(v4[i], t_mul(sc->omega, v5[i]));
    } else {                           // v0 = dst, v2 = src
        v0[i] = v2[i];
    }
}

}  // namespace mg
