// mg_coarse.cuh -- coarse-grid solvers: one CTA of 256 threads, all Krylov vectors resident in
// shared memory (global scratch only when the level does not fit into 227 KB).
//
// Follows the solver the reference's `.exa3` requests for the coarsest level
// (`solver_cgs = "CG"`, maxNumIts 1000, targetResReduction 1e-12,
// example_problems/Poisson/2D_FD_Poisson_fromL2.exa3:12-14) and, for the Helmholtz problem, the
// hand-written BiCGStab of example_problems/Helmholtz/2D_FD_Helmholtz_fromL3.exa3:396-433.
// Dot products: lane-strided partial sums over the flat interior index (field-major, then
// row-major), then the xor-butterfly of block_sum_256 -- the oracle mirrors this order exactly, so
// the real-valued coarse solve is bit-reproducible against it.
#pragma once
#include <type_traits>

#include "mg_kernels_generic.cuh"

namespace mg {

// thread index inside the 256-thread Krylov team; threads beyond it (the tail kernel runs 1024) idle
// through the loops but take part in every barrier
__device__ __forceinline__ int krylov_tid() { return threadIdx.x < 256 ? (int)threadIdx.x : (1 << 30); }

template <int D>
struct InteriorIter {
    int m, per_field, total;
    __device__ InteriorIter(const Geom& g, int F) {
        m = g.n - 2;
        per_field = (D == 3) ? m * m * m : m * m;
        total = F * per_field;
    }
    __device__ __forceinline__ void decode(int e, const Geom& g, int& f, long long& idx) const {
        f = e / per_field;
        int q = e - f * per_field;
        int x = q % m;
        int t = q / m;
        int y = t % m;
        int z = (D == 3) ? t / m : 0;
        idx = (long long)f * g.total + node_index<D>(g, x + 1, y + 1, (D == 3) ? z + 1 : 0);
    }
};

template <typename T, int F>
__device__ __forceinline__ FieldSet<T> as_fields(T* base, const Geom& g) {
    FieldSet<T> fs;
#pragma unroll
    for (int i = 0; i < MG_MAX_FIELDS; ++i) fs.p[i] = base + (long long)(i < F ? i : 0) * g.total;
    return fs;
}

// Element loop of the Krylov team with the first element of every thread decoded once (on the
// coarsest levels a thread owns at most one unknown, so no iteration pays the index arithmetic).
template <int D>
struct ElemCache {
    // The first element of a thread is decoded once; for a single field the following ones (e + 256, e + 512, ...)
    // are reached by stepping the node coordinates -- no integer division per element and loop, and no arrays of cached
    // indices (four of them cost the tail kernel, 512 threads, its registers: 60 -> 128 + spills).
    int f0, x, y, z, dx, dy, dz, m;
    long long idx0;
    bool has0, single;
    __device__ ElemCache(const Geom& g, const InteriorIter<D>& it)
        : f0(0), x(0), y(0), z(0), dx(0), dy(0), dz(0), m(it.m), idx0(0), has0(krylov_tid() < it.total),
          single(it.total == it.per_field) {
        if (has0) {
            it.decode(krylov_tid(), g, f0, idx0);
            const int q = krylov_tid() - f0 * it.per_field;
            x = q % m;
            const int t = q / m;
            y = t % m;
            z = (D == 3) ? t / m : 0;
        }
        dx = 256 % m;
        const int t = 256 / m;
        dy = t % m;
        dz = (D == 3) ? t / m : 0;
    }
    template <class Fn>
    __device__ __forceinline__ void each(const Geom& g, const InteriorIter<D>& it, Fn fn) const {
        if (!has0) return;
        fn(f0, idx0);
        if (single) {
            // three elements per round, indices first: their loads overlap (33^2: a thread owns at most four)
            int xx = x, yy = y, zz = z;
            for (int e = krylov_tid() + 256; e < it.total; e += 768) {
                int id[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    xx += dx; yy += dy; zz += dz;
                    if (xx >= m) { xx -= m; ++yy; }
                    if (yy >= m) { yy -= m; ++zz; }
                    id[k] = (int)node_index<D>(g, xx + 1, yy + 1, (D == 3) ? zz + 1 : 0);
                }
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    if (e + 256 * k < it.total) fn(0, (long long)id[k]);
            }
            return;
        }
        for (int e = krylov_tid() + 256; e < it.total; e += 256) {
            int ff; long long i;
            it.decode(e, g, ff, i);
            fn(ff, i);
        }
    }
};

template <typename T, int D>
__device__ T coarse_dot(const T* a, const T* b, const Geom& g, const InteriorIter<D>& it, T* s8) {
    T acc = t_zero(T());
    for (int e = krylov_tid(); e < it.total; e += 256) {
        int f; long long idx;
        it.decode(e, g, f, idx);
        acc = t_add(acc, t_mul(a[idx], b[idx]));
    }
    return block_sum_256<T>(acc, s8);
}

template <typename T, int D>
__device__ T coarse_dot(const T* a, const T* b, const Geom& g, const InteriorIter<D>& it, const ElemCache<D>& ec, T* s8) {
    T acc = t_zero(T());
    ec.each(g, it, [&](int, long long idx) { acc = t_add(acc, t_mul(a[idx], b[idx])); });
    return block_sum_256<T>(acc, s8);
}

// slaved Robin columns + zero rows on a full-array vector (2D, single field)
__device__ inline void coarse_bc(cplx* v, const Geom& g, cplx c) {
    __syncthreads();
    for (int t = krylov_tid(); t < g.n; t += 256) {
        long long row = (long long)t * g.pitch;
        v[row] = t_mul(v[row + 1], c);
        v[row + g.n - 1] = t_mul(v[row + g.n - 2], c);
    }
    __syncthreads();
    for (int t = krylov_tid(); t < g.n; t += 256) {
        v[t] = make_cplx(0.0, 0.0);
        v[(long long)(g.n - 1) * g.pitch + t] = make_cplx(0.0, 0.0);
    }
    __syncthreads();
}
__device__ inline void coarse_bc(double*, const Geom&, cplx) {}

__device__ inline double t_abs(double a) { return fabs(a); }
__device__ inline double t_abs(cplx a) { return t_hypot(a.re, a.im); }
__device__ inline double t_sqrt(double a) { return sqrt(a); }
__device__ inline cplx t_sqrt(cplx z) {      // principal square root
    double r = t_hypot(z.re, z.im);
    double sr = sqrt(0.5 * (r + z.re));
    double si = sqrt(0.5 * (r - z.re));
    return make_cplx(sr, z.im < 0 ? -si : si);
}
__device__ inline bool t_is_zero(double a) { return a == 0.0; }
__device__ inline bool t_is_zero(cplx a) { return a.re == 0.0 && a.im == 0.0; }

// ---- CG --------------------------------------------------------------------------------------------
// work: 4 vectors of F * g.total elements (x, r, p, ap), shared or global.
// CACHED_ONLY: the caller guarantees a scalar (2 D + 1)-entry constant operator (the lean tail interpreter): the generic
// loop -- and with it apply_row and its out-of-line variable-coefficient call, whose ABI costs the whole kernel its
// register allocation -- is not instantiated
template <typename T, int D, int F, bool CACHED_ONLY = false>
__device__ void coarse_cg_body(const Geom& g, const OperatorTable<T>* __restrict__ op, const FieldSet<T>& sol,
                               const FieldSet<T>& rhs, T* work, int max_its, double reduction, T* s8) {
    const long long vec = (long long)F * g.total;
    T* x = work;
    T* r = work + vec;
    T* p = work + 2 * vec;
    T* ap = work + 3 * vec;
    InteriorIter<D> it(g, F);
    const ElemCache<D> ec(g, it);
    for (long long i = krylov_tid(); i < 4 * vec; i += 256) work[i] = t_zero(T());
    __syncthreads();
    ec.each(g, it, [&](int f, long long idx) {
        T v = rhs.p[f][idx - (long long)f * g.total];
        r[idx] = v;
        p[idx] = v;
    });
    __syncthreads();
    T rs = coarse_dot<T, D>(r, r, g, it, ec, s8);
    const double r0 = t_abs(t_sqrt(rs));
    if (r0 != 0.0) {
        FieldSet<T> pf = as_fields<T, F>(p, g);
        // the loop, with (A p)(idx) supplied by the caller: the operator entries of a scalar constant stencil are
        // held in registers (the table lives in global memory: five dependent loads per node and iteration
        // otherwise -- the coarse solve on 33^2 is most of a configs[0] cycle)
        auto iterate = [&](auto&& apply_p) {
        for (int iter = 1; iter <= max_its; ++iter) {
            // ap, x and r are only ever touched by the thread that owns the element (the dot products read a
            // thread's own elements too), so no barrier is needed until p changes
            ec.each(g, it, [&](int f, long long idx) { ap[idx] = apply_p(f, idx); });
            T pap = coarse_dot<T, D>(p, ap, g, it, ec, s8);
            T alpha = t_div(rs, pap);
            ec.each(g, it, [&](int, long long idx) {
                x[idx] = t_add(x[idx], t_mul(alpha, p[idx]));
                r[idx] = t_sub(r[idx], t_mul(alpha, ap[idx]));
            });
            T rs_new = coarse_dot<T, D>(r, r, g, it, ec, s8);
            T beta = t_div(rs_new, rs);                  // issued before the test: overlaps the square root
            if (t_abs(t_sqrt(rs_new)) <= reduction * r0) break;
            ec.each(g, it, [&](int, long long idx) { p[idx] = t_add(r[idx], t_mul(beta, p[idx])); });
            rs = rs_new;
            __syncthreads();
        }
        };
        const int ne = F == 1 ? op->n_entries[0][0] : 0;
        auto cached = [&](auto ne_tag) {
            constexpr int NE = decltype(ne_tag)::value;
            int off[NE];
            T val[NE];
#pragma unroll
            for (int k = 0; k < NE; ++k) {
                off[k] = op->off[0][0][k];
                val[k] = op->val[0][0][k];
            }
            iterate([&](int, long long idx) { return apply_cached<T, NE>(off, val, p, idx); });
        };
        // one cached instantiation per dimension (the star: 2 D + 1 entries): every copy of the loop is code the tail
        // kernel has to fetch
        if constexpr (CACHED_ONLY) {
            cached(std::integral_constant<int, 2 * D + 1>{});
        } else {
            if (F == 1 && !is_variable<T, F>(op) && ne == 2 * D + 1) cached(std::integral_constant<int, 2 * D + 1>{});
            else iterate([&](int f, long long idx) { return apply_row<T, F>(op, pf, f, idx - (long long)f * g.total); });
        }
    }
    __syncthreads();
    // solution field: whole array (boundary = 0 for the error equation)
    for (int f = 0; f < F; ++f)
        for (long long i = krylov_tid(); i < g.total; i += 256) sol.p[f][i] = x[(long long)f * g.total + i];
}

// ---- BiCGStab (Helmholtz/...exa3:396-433), dot without conjugation -----------------------------------
// work: 7 vectors (x, res, res_hat, p, nu, s, t)
template <typename T, int D, int F>
__device__ void coarse_bicgstab_body(const Geom& g, const OperatorTable<T>* __restrict__ op, const FieldSet<T>& sol,
                                     const FieldSet<T>& rhs, T* work, int max_its, double reduction, int robin,
                                     cplx bc_factor, T* s8) {
    const long long vec = (long long)F * g.total;
    T* x = work;
    T* res = work + vec;
    T* rhat = work + 2 * vec;
    T* p = work + 3 * vec;
    T* nu = work + 4 * vec;
    T* s = work + 5 * vec;
    T* tt = work + 6 * vec;
    InteriorIter<D> it(g, F);
    for (long long i = krylov_tid(); i < 7 * vec; i += 256) work[i] = t_zero(T());
    __syncthreads();
    for (int e = krylov_tid(); e < it.total; e += 256) {
        int f; long long idx;
        it.decode(e, g, f, idx);
        T v = rhs.p[f][idx - (long long)f * g.total];   // x = 0  =>  res = f
        res[idx] = v;
        rhat[idx] = v;
    }
    __syncthreads();
    T cur = t_sqrt(coarse_dot<T, D>(res, res, g, it, s8));
    const double init = t_abs(cur);
    if (!t_is_zero(cur)) {
        T one = t_zero(T());
        *reinterpret_cast<double*>(&one) = 1.0;
        T alpha = one, rho_new = one, omega = one;
        FieldSet<T> pf = as_fields<T, F>(p, g);
        FieldSet<T> sf = as_fields<T, F>(s, g);
        for (int iter = 1; iter <= max_its; ++iter) {
            T rho = rho_new;
            rho_new = coarse_dot<T, D>(rhat, res, g, it, s8);
            T beta = t_mul(t_div(rho_new, rho), t_div(alpha, omega));
            for (int e = krylov_tid(); e < it.total; e += 256) {
                int f; long long idx;
                it.decode(e, g, f, idx);
                p[idx] = t_add(res[idx], t_mul(beta, t_sub(p[idx], t_mul(omega, nu[idx]))));
            }
            if (robin) coarse_bc(p, g, bc_factor); else __syncthreads();
            for (int e = krylov_tid(); e < it.total; e += 256) {
                int f; long long idx;
                it.decode(e, g, f, idx);
                nu[idx] = apply_row<T, F>(op, pf, f, idx - (long long)f * g.total);
            }
            __syncthreads();
            alpha = t_div(rho_new, coarse_dot<T, D>(rhat, nu, g, it, s8));
            for (int e = krylov_tid(); e < it.total; e += 256) {
                int f; long long idx;
                it.decode(e, g, f, idx);
                x[idx] = t_add(x[idx], t_mul(alpha, p[idx]));          // h = x + alpha p
                s[idx] = t_sub(res[idx], t_mul(alpha, nu[idx]));
            }
            if (robin) coarse_bc(s, g, bc_factor); else __syncthreads();
            for (int e = krylov_tid(); e < it.total; e += 256) {
                int f; long long idx;
                it.decode(e, g, f, idx);
                tt[idx] = apply_row<T, F>(op, sf, f, idx - (long long)f * g.total);
            }
            __syncthreads();
            T ts = coarse_dot<T, D>(tt, s, g, it, s8);
            T tdt = coarse_dot<T, D>(tt, tt, g, it, s8);
            omega = t_div(ts, tdt);
            for (int e = krylov_tid(); e < it.total; e += 256) {
                int f; long long idx;
                it.decode(e, g, f, idx);
                x[idx] = t_add(x[idx], t_mul(omega, s[idx]));          // x = h + omega s
                res[idx] = t_sub(s[idx], t_mul(omega, tt[idx]));
            }
            __syncthreads();
            cur = t_sqrt(coarse_dot<T, D>(res, res, g, it, s8));
            if (t_abs(cur) < reduction * init) break;
        }
    }
    __syncthreads();
    if (robin) coarse_bc(x, g, bc_factor);
    for (int f = 0; f < F; ++f)
        for (long long i = krylov_tid(); i < g.total; i += 256) sol.p[f][i] = x[(long long)f * g.total + i];
}

// ---- kernel wrappers: one CTA of 256 threads, Krylov vectors in dynamic shared memory when they fit ----
template <typename T, int D, int F, bool SMEM>
__global__ void __launch_bounds__(256) k_coarse_cg(Geom g, const OperatorTable<T>* __restrict__ op, FieldSet<T> sol,
                                                   FieldSet<T> rhs, T* gwork, int max_its, double reduction) {
    pdl_sync();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T s8[8];
    coarse_cg_body<T, D, F>(g, op, sol, rhs, SMEM ? reinterpret_cast<T*>(smem_raw) : gwork, max_its, reduction, s8);
}

template <typename T, int D, int F, bool SMEM>
__global__ void __launch_bounds__(256) k_coarse_bicgstab(Geom g, const OperatorTable<T>* __restrict__ op,
                                                         FieldSet<T> sol, FieldSet<T> rhs, T* gwork, int max_its,
                                                         double reduction, int robin, cplx bc_factor) {
    pdl_sync();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T s8[8];
    coarse_bicgstab_body<T, D, F>(g, op, sol, rhs, SMEM ? reinterpret_cast<T*>(smem_raw) : gwork, max_its, reduction,
                                  robin, bc_factor, s8);
}

}  // namespace mg
