// mg_kernels_generic.cuh -- generic stage kernels: any dimension (2/3), 1-2 fields, real or
// complex, arbitrary constant stencils.  One thread per node, x fastest (coalesced).  These are
// the reference-semantics kernels every configuration can fall back on; the hot 2D/3D scalar
// Poisson stages have specialised versions in mg_kernels_fast.cuh.
#pragma once
#include "mg_common.cuh"

namespace mg {

template <int D>
__device__ __forceinline__ bool node_coords(const Geom& g, int& x, int& y, int& z) {
    x = blockIdx.x * blockDim.x + threadIdx.x;
    y = blockIdx.y;
    z = (D == 3) ? blockIdx.z : 0;
    return x < g.n;
}

template <int D>
__device__ __forceinline__ bool is_interior(const Geom& g, int x, int y, int z) {
    bool in = x >= 1 && x <= g.n - 2 && y >= 1 && y <= g.n - 2;
    if (D == 3) in = in && z >= 1 && z <= g.nz - 2 && z + g.z0 >= 1 && z + g.z0 <= g.n - 2;
    return in;
}

template <int D>
__device__ __forceinline__ long long node_index(const Geom& g, int x, int y, int z) {
    return (D == 3 ? (long long)z * g.plane : 0ll) + (long long)y * g.pitch + x;
}

// ---- variable-coefficient Poisson: -div(c grad u), c = exp(kappa * prod_d (x_d - x_d^2)) ----------------
// stencils/gallery.py:87-185 (get_coefficient_2D / get_coefficient_3D and the `Operator A from Stencil` text
// of Poisson{2,3}DVariableCoefficients.generate_exa3): the coefficient is evaluated at the half-grid
// positions around the node, an off-centre value is (-1.0 * c_face) / (h * h), the centre is
// ((cE + cW) / hh) + ((cN + cS) / hh) [+ ((cT + cB) / hh)].  Node positions i * h, i * h +- 0.5 * h and
// h * h are exact in fp64 (h = 2^-level), so the oracle forms the same arguments bit for bit; only exp()
// differs (device vs libm, <= 1 ulp).  A u is accumulated in memory order like every other operator.
__device__ __forceinline__ double vc_axis(double p) { return p - p * p; }

template <int VD>
__device__ __forceinline__ double vc_apply_d(const OperatorTable<double>* __restrict__ op, const double* __restrict__ u,
                                             long long idx, double& diag) {
    const double h = op->vc_h, hh = h * h, kappa = op->vc_kappa, hf = 0.5 * h;
    const int pitch = op->vc_pitch;
    int z = 0;
    long long r = idx;
    if (VD == 3) {
        z = (int)(idx / op->vc_plane);
        r = idx - (long long)z * op->vc_plane;
    }
    const int y = (int)(r / pitch);
    const int x = (int)(r - (long long)y * pitch);
    const double px = x * h, py = y * h;
    const double axc = vc_axis(px), axe = vc_axis(px + hf), axw = vc_axis(px - hf);
    const double ayc = vc_axis(py), ayn = vc_axis(py + hf), ays = vc_axis(py - hf);
    if (VD == 2) {
        const double cE = exp(kappa * (axe * ayc)), cW = exp(kappa * (axw * ayc));
        const double cN = exp(kappa * (axc * ayn)), cS = exp(kappa * (axc * ays));
        diag = ((cE + cW) / hh) + ((cN + cS) / hh);
        double acc = ((-1.0 * cS) / hh) * u[idx - pitch];
        acc = acc + ((-1.0 * cW) / hh) * u[idx - 1];
        acc = acc + diag * u[idx];
        acc = acc + ((-1.0 * cE) / hh) * u[idx + 1];
        acc = acc + ((-1.0 * cN) / hh) * u[idx + pitch];
        return acc;
    } else {
        const double pz = z * h;
        const double azc = vc_axis(pz), azt = vc_axis(pz + hf), azb = vc_axis(pz - hf);
        const double cE = exp(kappa * ((axe * ayc) * azc)), cW = exp(kappa * ((axw * ayc) * azc));
        const double cN = exp(kappa * ((axc * ayn) * azc)), cS = exp(kappa * ((axc * ays) * azc));
        const double cT = exp(kappa * ((axc * ayc) * azt)), cB = exp(kappa * ((axc * ayc) * azb));
        diag = (((cE + cW) / hh) + ((cN + cS) / hh)) + ((cT + cB) / hh);
        double acc = ((-1.0 * cB) / hh) * u[idx - op->vc_plane];
        acc = acc + ((-1.0 * cS) / hh) * u[idx - pitch];
        acc = acc + ((-1.0 * cW) / hh) * u[idx - 1];
        acc = acc + diag * u[idx];
        acc = acc + ((-1.0 * cE) / hh) * u[idx + 1];
        acc = acc + ((-1.0 * cN) / hh) * u[idx + pitch];
        acc = acc + ((-1.0 * cT) / hh) * u[idx + op->vc_plane];
        return acc;
    }
}

// out of line: the call sits behind a run-time flag in every kernel that applies an operator (apply_row), and the
// 4-6 inlined exp() bodies would otherwise be copied into all of them (code size is what bounds the tail kernel)
__device__ __noinline__ double vc_apply(const OperatorTable<double>* __restrict__ op, const double* __restrict__ u,
                                        long long idx, double& diag) {
    return op->vc_dim == 3 ? vc_apply_d<3>(op, u, idx, diag) : vc_apply_d<2>(op, u, idx, diag);
}

template <typename T, int F>
__device__ __forceinline__ bool is_variable(const OperatorTable<T>* __restrict__ op) {
    if constexpr (sizeof(T) == sizeof(double) && F == 1) return op->vc_kind != 0;
    return false;
}

// (A u)_row at element idx, canonical accumulation order (block column ascending, entries as given)
template <typename T, int F>
__device__ __forceinline__ T apply_row(const OperatorTable<T>* __restrict__ op, const FieldSet<T>& u, int row,
                                       long long idx) {
    if constexpr (sizeof(T) == sizeof(double) && F == 1) {
        if (op->vc_kind != 0) {
            double diag;
            return vc_apply(op, u.p[0], idx, diag);
        }
    }
    T acc = t_zero(T());
    bool first = true;
#pragma unroll
    for (int j = 0; j < F; ++j) {
        const int ne = op->n_entries[row][j];
        for (int k = 0; k < ne; ++k) {
            T term = t_mul(op->val[row][j][k], u.p[j][idx + op->off[row][j][k]]);
            acc = first ? term : t_add(acc, term);
            first = false;
        }
    }
    return acc;
}

// (A u)(idx) for a scalar operator whose NE entries sit in registers (same entry order as apply_row: results are
// bit-identical); the table itself lives in global memory, which costs NE dependent loads per node otherwise
template <typename T, int NE>
__device__ __forceinline__ T apply_cached(const int (&off)[NE], const T (&val)[NE], const T* __restrict__ u, long long idx) {
    T acc = t_mul(val[0], u[idx + off[0]]);
#pragma unroll
    for (int k = 1; k < NE; ++k) acc = t_add(acc, t_mul(val[k], u[idx + off[k]]));
    return acc;
}
template <typename T, int NE>
__device__ __forceinline__ void load_row_entries(const OperatorTable<T>* __restrict__ op, int (&off)[NE], T (&val)[NE]) {
#pragma unroll
    for (int k = 0; k < NE; ++k) {
        off[k] = op->off[0][0][k];
        val[k] = op->val[0][0][k];
    }
}

// res = rhs - A sol
template <typename T, int D, int F>
__device__ __forceinline__ void residual_node(const Geom& g, const OperatorTable<T>* __restrict__ op,
                                              const FieldSet<T>& sol, const FieldSet<T>& rhs, const FieldSet<T>& res,
                                              int x, int y, int z) {
    if (!is_interior<D>(g, x, y, z)) return;
    long long idx = node_index<D>(g, x, y, z);
#pragma unroll
    for (int i = 0; i < F; ++i) res.p[i][idx] = t_sub(rhs.p[i][idx], apply_row<T, F>(op, sol, i, idx));
}

template <typename T, int D, int F>
__global__ void k_residual(Geom g, const OperatorTable<T>* __restrict__ op, FieldSet<T> sol, FieldSet<T> rhs,
                           FieldSet<T> res) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z)) return;
    residual_node<T, D, F>(g, op, sol, rhs, res, x, y, z);
}

// Point smoother, out of place: dst_s = src_s + sum_t W[s][t] r_t on the selected colour
// (colour < 0: all interior nodes), dst_s = src_s elsewhere (including the boundary).
template <typename T, int D, int F>
__device__ __forceinline__ void smooth_point_node(const Geom& g, const OperatorTable<T>* __restrict__ op,
                                                  const SmootherTable<T>* __restrict__ sm, const T* __restrict__ w,
                                                  const FieldSet<T>& src, const FieldSet<T>& rhs, const FieldSet<T>& dst,
                                                  int colour, int x, int y, int z) {
    long long idx = node_index<D>(g, x, y, z);
    const int m = sm->n_slots;
    bool active = is_interior<D>(g, x, y, z) && (colour < 0 || ((x + y + z + g.z0) & 1) == colour);
    if (!active) {
        for (int s = 0; s < m; ++s) {
            int f = sm->slot_field[s];
            dst.p[f][idx] = src.p[f][idx];
        }
        return;
    }
    const T* wk = w + (long long)sm->mask_index[(1 << m) - 1] * m * m;
    if constexpr (sizeof(T) == sizeof(double) && F == 1) {
        if (op->vc_kind != 0) {        // W = omega / a00(node): the table holds omega (mg_add_smoother)
            double diag;
            const double r0 = rhs.p[0][idx] - vc_apply(op, src.p[0], idx, diag);
            dst.p[0][idx] = src.p[0][idx] + (wk[0] / diag) * r0;
            return;
        }
    }
    T r[MG_MAX_FIELDS];
    for (int s = 0; s < m; ++s) {
        int f = sm->slot_field[s];
        r[s] = t_sub(rhs.p[f][idx], apply_row<T, F>(op, src, f, idx));
    }
    for (int s = 0; s < m; ++s) {
        T delta = t_mul(wk[s * m + 0], r[0]);
        for (int t = 1; t < m; ++t) delta = t_add(delta, t_mul(wk[s * m + t], r[t]));
        int f = sm->slot_field[s];
        dst.p[f][idx] = t_add(src.p[f][idx], delta);
    }
}

template <typename T, int D, int F>
__global__ void k_smooth_point(Geom g, const OperatorTable<T>* __restrict__ op,
                               const SmootherTable<T>* __restrict__ sm, const T* __restrict__ w, FieldSet<T> src,
                               FieldSet<T> rhs, FieldSet<T> dst, int colour) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z)) return;
    smooth_point_node<T, D, F>(g, op, sm, w, src, rhs, dst, colour, x, y, z);
}

// Tile smoother (Jacobi), one thread per tile, out of place.
template <typename T, int D, int F>
__device__ __forceinline__ void smooth_tile_node(const Geom& g, const OperatorTable<T>* __restrict__ op,
                                                 const SmootherTable<T>* __restrict__ sm, const T* __restrict__ w,
                                                 const FieldSet<T>& src, const FieldSet<T>& rhs, const FieldSet<T>& dst,
                                                 int tx, int ty, int tz) {
    const int m = sm->n_slots;
    T r[MG_MAX_SLOTS];
    long long idx[MG_MAX_SLOTS];
    unsigned mask = 0;
    for (int s = 0; s < m; ++s) {
        int x = tx * sm->period[0] + sm->slot_pos[s][0];
        int y = ty * sm->period[1] + sm->slot_pos[s][1];
        int z = (D == 3) ? tz * sm->period[2] + sm->slot_pos[s][2] : 0;
        bool exists = x < g.n && y < g.n && (D < 3 || z < g.nz);
        idx[s] = exists ? node_index<D>(g, x, y, z) : -1;
        r[s] = t_zero(T());
        if (exists && is_interior<D>(g, x, y, z)) {
            mask |= 1u << s;
            int f = sm->slot_field[s];
            r[s] = t_sub(rhs.p[f][idx[s]], apply_row<T, F>(op, src, f, idx[s]));
        }
    }
    int k = sm->mask_index[mask];
    const T* wk = w + (long long)(k < 0 ? 0 : k) * m * m;
    for (int s = 0; s < m; ++s) {
        if (idx[s] < 0) continue;
        int f = sm->slot_field[s];
        T old = src.p[f][idx[s]];
        if ((mask >> s) & 1u) {
            T delta = t_zero(T());
            bool first = true;
            for (int t = 0; t < m; ++t) {
                if (!((mask >> t) & 1u)) continue;
                T term = t_mul(wk[s * m + t], r[t]);
                delta = first ? term : t_add(delta, term);
                first = false;
            }
            dst.p[f][idx[s]] = t_add(old, delta);
        } else {
            dst.p[f][idx[s]] = old;
        }
    }
}

template <typename T, int D, int F>
__global__ void k_smooth_tile(Geom g, const OperatorTable<T>* __restrict__ op,
                              const SmootherTable<T>* __restrict__ sm, const T* __restrict__ w, FieldSet<T> src,
                              FieldSet<T> rhs, FieldSet<T> dst, int tiles_x) {
    pdl_sync();
    int tx = blockIdx.x * blockDim.x + threadIdx.x;
    int ty = blockIdx.y;
    int tz = (D == 3) ? blockIdx.z : 0;
    if (tx >= tiles_x) return;
    smooth_tile_node<T, D, F>(g, op, sm, w, src, rhs, dst, tx, ty, tz);
}

// full weighting: coarse(I) = sum_o w_o fine(2I + o), offsets in memory order
template <typename T, int D>
__device__ __forceinline__ T restrict_at(const T* fine, const Geom& gf, int X, int Y, int Z /* fine-local 2Z */) {
    T acc = t_zero(T());
    bool first = true;
    const double w1[3] = {0.25, 0.5, 0.25};
    for (int oz = (D == 3 ? -1 : 0); oz <= (D == 3 ? 1 : 0); ++oz)
        for (int oy = -1; oy <= 1; ++oy)
            for (int ox = -1; ox <= 1; ++ox) {
                double wgt = w1[ox + 1] * w1[oy + 1];   // products of powers of two: exact
                if (D == 3) wgt = w1[oz + 1] * wgt;
                T term = t_scale(wgt, fine[node_index<D>(gf, 2 * X + ox, 2 * Y + oy, Z + oz)]);
                acc = first ? term : t_add(acc, term);
                first = false;
            }
    return acc;
}

template <typename T, int D, int F>
__device__ __forceinline__ void restrict_node(const Geom& gc, const Geom& gf, const FieldSet<T>& fine,
                                              const FieldSet<T>& coarse_rhs, int x, int y, int z) {
    if (!is_interior<D>(gc, x, y, z)) return;
    long long idx = node_index<D>(gc, x, y, z);
    const int zf = (D == 3) ? 2 * (z + gc.z0) - gf.z0 : 0;      // fine-local plane of the coarse node
#pragma unroll
    for (int i = 0; i < F; ++i) coarse_rhs.p[i][idx] = restrict_at<T, D>(fine.p[i], gf, x, y, zf);
}

template <typename T, int D, int F>
__global__ void k_restrict(Geom gc, Geom gf, FieldSet<T> fine, FieldSet<T> coarse_rhs) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(gc, x, y, z)) return;
    restrict_node<T, D, F>(gc, gf, fine, coarse_rhs, x, y, z);
}

// fused residual + restriction (generic: each coarse thread recomputes its 3^D fine residuals)
template <typename T, int D, int F>
__device__ __forceinline__ void residual_restrict_node(const Geom& gc, const Geom& gf,
                                                       const OperatorTable<T>* __restrict__ op, const FieldSet<T>& sol,
                                                       const FieldSet<T>& rhs, const FieldSet<T>& coarse_rhs, int X,
                                                       int Y, int Z) {
    if (!is_interior<D>(gc, X, Y, Z)) return;
    long long cidx = node_index<D>(gc, X, Y, Z);
    const int zf = (D == 3) ? 2 * (Z + gc.z0) - gf.z0 : 0;
    const double w1[3] = {0.25, 0.5, 0.25};
#pragma unroll
    for (int i = 0; i < F; ++i) {
        T acc = t_zero(T());
        bool first = true;
        for (int oz = (D == 3 ? -1 : 0); oz <= (D == 3 ? 1 : 0); ++oz)
            for (int oy = -1; oy <= 1; ++oy)
                for (int ox = -1; ox <= 1; ++ox) {
                    double wgt = w1[ox + 1] * w1[oy + 1];
                    if (D == 3) wgt = w1[oz + 1] * wgt;
                    long long fidx = node_index<D>(gf, 2 * X + ox, 2 * Y + oy, zf + oz);
                    T r = t_sub(rhs.p[i][fidx], apply_row<T, F>(op, sol, i, fidx));
                    T term = t_scale(wgt, r);
                    acc = first ? term : t_add(acc, term);
                    first = false;
                }
        coarse_rhs.p[i][cidx] = acc;
    }
}

template <typename T, int D, int F>
__global__ void k_residual_restrict(Geom gc, Geom gf, const OperatorTable<T>* __restrict__ op, FieldSet<T> sol,
                                    FieldSet<T> rhs, FieldSet<T> coarse_rhs) {
    pdl_sync();
    int X, Y, Z;
    if (!node_coords<D>(gc, X, Y, Z)) return;
    residual_restrict_node<T, D, F>(gc, gf, op, sol, rhs, coarse_rhs, X, Y, Z);
}

// (P e)(x): 2^-k * (sum of the 2^k enclosing coarse nodes in memory order), k = #odd coordinates
template <typename T, int D>
__device__ __forceinline__ T prolong_at(const T* coarse, const Geom& gc, int x, int y, int z) {
    // z is the GLOBAL fine plane index; the coarse array is addressed slab-locally
    int px = x & 1, py = y & 1, pz = (D == 3) ? (z & 1) : 0;
    int X = x >> 1, Y = y >> 1, Z = (z >> 1) - gc.z0;
    T acc = t_zero(T());
    bool first = true;
    for (int cz = 0; cz <= pz; ++cz)
        for (int cy = 0; cy <= py; ++cy)
            for (int cx = 0; cx <= px; ++cx) {
                T term = coarse[node_index<D>(gc, X + cx, Y + cy, Z + cz)];
                acc = first ? term : t_add(acc, term);
                first = false;
            }
    int k = px + py + pz;
    if (k == 0) return acc;
    const double scale = (k == 1) ? 0.5 : (k == 2 ? 0.25 : 0.125);
    return t_scale(scale, acc);
}

// sol += omega * P coarse   (add != 0)   |   dst = P coarse   (add == 0)
template <typename T, int D, int F>
__device__ __forceinline__ void prolong_node(const Geom& gf, const Geom& gc, const FieldSet<T>& coarse,
                                             const FieldSet<T>& dst, double omega, int add, int x, int y, int z) {
    if (!is_interior<D>(gf, x, y, z)) return;
    long long idx = node_index<D>(gf, x, y, z);
#pragma unroll
    for (int i = 0; i < F; ++i) {
        T pe = prolong_at<T, D>(coarse.p[i], gc, x, y, z + gf.z0);
        dst.p[i][idx] = add ? t_add(dst.p[i][idx], t_scale(omega, pe)) : pe;
    }
}

template <typename T, int D, int F>
__global__ void k_prolong(Geom gf, Geom gc, FieldSet<T> coarse, FieldSet<T> dst, double omega, int add) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(gf, x, y, z)) return;
    prolong_node<T, D, F>(gf, gc, coarse, dst, omega, add, x, y, z);
}

// zero-fill (16 bytes per thread and step)
__global__ void k_zero(double2* __restrict__ p, long long n16) {
    pdl_sync();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += (long long)gridDim.x * blockDim.x)
        p[i] = make_double2(0.0, 0.0);
}

// Robin columns of the Helmholtz problem (2D): u(0,y) = c u(1,y), u(n-1,y) = c u(n-2,y); rows y=0,n-1 -> 0
__global__ void k_bc_helmholtz(Geom g, cplx* u, cplx c) {
    pdl_sync();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.n) return;
    long long row = (long long)t * g.pitch;
    u[row] = t_mul(u[row + 1], c);
    u[row + g.n - 1] = t_mul(u[row + g.n - 2], c);
}
__global__ void k_bc_helmholtz_rows(Geom g, cplx* u) {
    pdl_sync();
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.n) return;
    u[t] = make_cplx(0.0, 0.0);
    u[(long long)(g.n - 1) * g.pitch + t] = make_cplx(0.0, 0.0);
}

// ---- residual norm: deterministic two-stage reduction -------------------------------------------
__device__ __forceinline__ double sq_acc(double r) { return r * r; }
__device__ __forceinline__ cplx sq_acc(cplx r) { return t_mul(r, r); }   // dot without conjugation

template <typename T>
__device__ __forceinline__ T warp_sum(T v);
template <>
__device__ __forceinline__ double warp_sum<double>(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <>
__device__ __forceinline__ cplx warp_sum<cplx>(cplx v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v.re += __shfl_xor_sync(0xffffffffu, v.re, o);
        v.im += __shfl_xor_sync(0xffffffffu, v.im, o);
    }
    return v;
}

// block of 256 threads; result valid in every thread
template <typename T>
__device__ __forceinline__ T block_sum_256(T v, T* smem8) {
    v = warp_sum<T>(v);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0 && warp < 8) smem8[warp] = v;      // CTAs with more than 256 threads: only the first 8 warps contribute
    __syncthreads();
    T a0 = t_add(smem8[0], smem8[4]), a1 = t_add(smem8[1], smem8[5]);
    T a2 = t_add(smem8[2], smem8[6]), a3 = t_add(smem8[3], smem8[7]);
    return t_add(t_add(a0, a2), t_add(a1, a3));
}

// partial[b] = sum over the rows owned by block b of (rhs - A sol)^2 ; rows = F * n^(D-1)
template <typename T, int D, int F>
__global__ void __launch_bounds__(256) k_norm_partial(Geom g, const OperatorTable<T>* __restrict__ op,
                                                      FieldSet<T> sol, FieldSet<T> rhs, T* __restrict__ partial) {
    pdl_sync();
    __shared__ T s8[8];
    const int rows_per_field = (D == 3) ? g.n * g.nz : g.n;
    const int rows = F * rows_per_field;
    T acc = t_zero(T());
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        int f = row / rows_per_field;
        int rr = row - f * rows_per_field;
        int z = (D == 3) ? rr / g.n : 0;
        int y = (D == 3) ? rr - z * g.n : rr;
        if (!is_interior<D>(g, 1, y, z)) continue;
        for (int x = 1 + threadIdx.x; x <= g.n - 2; x += 256) {
            long long idx = node_index<D>(g, x, y, z);
            T r = t_sub(rhs.p[f][idx], apply_row<T, F>(op, sol, f, idx));
            acc = t_add(acc, sq_acc(r));
        }
    }
    T total = block_sum_256<T>(acc, s8);
    if (threadIdx.x == 0) partial[blockIdx.x] = total;
}

// out[0] = |sqrt(sum partial)|  (real: sqrt; complex: modulus of the complex square root);
// squares != 0 (real data): the sum itself -- a rank's share of a domain-decomposed norm (mg_dd.cuh)
template <typename T>
__global__ void __launch_bounds__(256) k_norm_final(const T* __restrict__ partial, int n, double* __restrict__ out,
                                                    int squares) {
    pdl_sync();
    __shared__ T s8[8];
    T acc = t_zero(T());
    for (int i = threadIdx.x; i < n; i += 256) acc = t_add(acc, partial[i]);
    T total = block_sum_256<T>(acc, s8);
    if (threadIdx.x == 0) {
        if constexpr (sizeof(T) == sizeof(double)) {
            const double t = *reinterpret_cast<double*>(&total);
            out[0] = squares ? t : sqrt(t);
        } else {
            cplx c = *reinterpret_cast<cplx*>(&total);
            out[0] = sqrt(t_hypot(c.re, c.im));       // |sqrt(z)| = sqrt(|z|)
        }
    }
}

}  // namespace mg
