// mg_common.cuh -- shared types of the B200 multigrid engine (sm_100a only).
//
// Numerical contract: every kernel evaluates stencils in the entry order it
// is given, with separate multiply and add -- this translation unit is compiled with
// --fmad=false so that results are bit-identical to the numpy oracle for real fp64.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define MG_MAX_DIM 3
#define MG_MAX_FIELDS 2
#define MG_MAX_ENTRIES 27          // full 3x3x3 stencil
#define MG_MAX_SLOTS 8             // maximum_local_system_size (optimization/program.py:774)

namespace mg {

struct cplx {
    double re, im;
};

__host__ __device__ inline cplx make_cplx(double a, double b) { cplx c; c.re = a; c.im = b; return c; }

// ---- arithmetic traits: T = double or cplx ------------------------------------------------
__host__ __device__ inline double t_zero(double) { return 0.0; }
__host__ __device__ inline cplx t_zero(cplx) { return make_cplx(0.0, 0.0); }
__host__ __device__ inline double t_add(double a, double b) { return a + b; }
__host__ __device__ inline cplx t_add(cplx a, cplx b) { return make_cplx(a.re + b.re, a.im + b.im); }
__host__ __device__ inline double t_sub(double a, double b) { return a - b; }
__host__ __device__ inline cplx t_sub(cplx a, cplx b) { return make_cplx(a.re - b.re, a.im - b.im); }
__host__ __device__ inline double t_mul(double a, double b) { return a * b; }
__host__ __device__ inline cplx t_mul(cplx a, cplx b) {
    return make_cplx(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re);
}
__host__ __device__ inline double t_scale(double s, double a) { return s * a; }
__host__ __device__ inline cplx t_scale(double s, cplx a) { return make_cplx(s * a.re, s * a.im); }
__host__ __device__ inline cplx t_div(cplx a, cplx b) {
    double d = b.re * b.re + b.im * b.im;
    return make_cplx((a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d);
}
__host__ __device__ inline double t_div(double a, double b) { return a / b; }
// |z| in plain IEEE operations (not hypot, whose last bit differs between libm and the device): the
// oracle states the same formula, so complex Krylov scalars are bit-reproducible against it
__host__ __device__ inline double t_hypot(double re, double im) { return sqrt(re * re + im * im); }

// ---- geometry of one level -------------------------------------------------------------------
struct Geom {
    int dim;
    int n;                 // nodes per dimension = 2^level + 1
    int pitch;             // elements per x-row (n rounded up to 16)
    long long plane;       // pitch * n (3D) ; total elements = plane * nz (3D) or pitch * n (2D)
    long long total;       // allocated elements of one field array
    // 3D slab (domain decomposition along z): the array holds planes z0 .. z0+nz-1 of the global grid;
    // local planes 0 and nz-1 are ghost planes (a neighbour's data) or the global Dirichlet planes.
    // A whole grid is the slab nz = n, z0 = 0.  2D: nz = 1, z0 = 0.
    int nz;
    int z0;
};

// Block operator of one level: F x F constant stencils, linear element offsets precomputed.
template <typename T>
struct OperatorTable {
    int n_entries[MG_MAX_FIELDS][MG_MAX_FIELDS];
    int off[MG_MAX_FIELDS][MG_MAX_FIELDS][MG_MAX_ENTRIES];     // linear offset in elements
    T val[MG_MAX_FIELDS][MG_MAX_FIELDS][MG_MAX_ENTRIES];
    // Variable-coefficient operator (mg_set_variable_coefficient; scalar real problems): when vc_kind != 0
    // the values above are placeholders and every kernel evaluates the (2d+1)-point stencil of
    // -div(c grad u) at the node from c at the half-grid positions (mg_kernels_generic.cuh vc_apply);
    // the geometry is repeated here because the operator is applied from a linear element index.
    int vc_kind;               // mg_vc_kind
    int vc_dim, vc_n, vc_pitch;
    long long vc_plane;
    double vc_kappa, vc_h;
};

// One `solve locally` statement.
template <typename T>
struct SmootherTable {
    int n_slots;
    int period[MG_MAX_DIM];
    int slot_field[MG_MAX_SLOTS];
    int slot_pos[MG_MAX_SLOTS][MG_MAX_DIM];
    int mask_index[1 << MG_MAX_SLOTS];      // validity mask -> index into w (or -1)
    // w[k][s][t] follows in device memory (n_masks * n_slots * n_slots values of T)
};

// Programmatic dependent launch: every kernel lets its successor start launching at once
// (launch_dependents) and then waits until its predecessor has completed and flushed (wait), before
// it touches global memory.  Hides the launch latency between the ~70 small kernels of a cycle; a
// no-op when the kernel was launched without the attribute.
__device__ __forceinline__ void pdl_sync() {
#if defined(__CUDA_ARCH__)
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}

template <typename T>
struct FieldSet {
    T* p[MG_MAX_FIELDS];
};

}  // namespace mg
