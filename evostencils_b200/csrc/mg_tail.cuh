// mg_tail.cuh -- "coarse tail": every op of a cycle that touches only small levels (<= tail_max_level,
// default 6: at most 65^D nodes) is executed by ONE persistent CTA of 256 threads that interprets a
// command list passed by value in the kernel parameters, with __syncthreads() between commands.
// A V-cycle spends ~10 launches per level on those levels; each is a few microseconds of launch and
// drain latency for nanoseconds of work, so the tail removes ~50 kernel nodes per cycle from the
// captured graph.  The commands call the same per-node device functions as the grid kernels
// (mg_kernels_generic.cuh) and the same Krylov bodies (mg_coarse.cuh): results are bit-identical.
// Fields of these levels are L1/L2 resident (65^2 doubles = 34 KB per array).
//
// The program lives in device memory (uploaded once when the cycle is captured): its header (level
// table, Krylov settings) and the commands in use are copied to shared memory in one pass at kernel
// start (measured with clock64: reading each command from the kernel parameter bank, or from L2
// one step ahead, put ~1000 cycles under every step -- more than the work of a step on levels <= 4).
// Cluster mode (tail_ctas > 1): the kernel is launched as ONE thread-block cluster.  Commands on
// levels above `solo_level` are "wide": every CTA takes a slice of the nodes and the commands are
// separated by barrier.cluster (release/acquire at cluster scope, ~0.2 us, makes the other CTAs'
// stores visible to plain loads).  Commands on the smallest levels and the Krylov solve are run by
// CTA 0 alone and separated by __syncthreads() only.  Every CTA walks the same command list, so
// all of them execute the same sequence of cluster barriers.
// Coloured star sweeps run in place and visit only the nodes of their colour (a node of one
// colour reads only nodes of the other), which halves the work of a Gauss-Seidel step.
#pragma once
#include <type_traits>
#include "../../include/mgengine.h"
#include "mg_coarse.cuh"

namespace mg {

constexpr int TAIL_MAX_CMDS = 224;
constexpr int TAIL_MAX_LEVELS = 10;
constexpr int TAIL_NT = 512;            // threads of the tail CTA (the Krylov bodies use the first 256)

enum TailKind {
    T_SMOOTH_POINT = 0, T_SMOOTH_TILE, T_RESIDUAL, T_RESTRICT, T_RESIDUAL_RESTRICT, T_ZERO, T_PROLONG, T_CGS, T_BC
};

struct TailCmd {
    int kind;
    int lv;                        // index (level - min_level) of the level that is written
    int colour;                    // SMOOTH_POINT: -1 all, 0/1 colour
    int add;                       // PROLONG: 1 = sol += omega P e, 0 = dst = P e
    int star;                      // scalar real star stencil: coefficients from TailLevel::st, weight = omega
    int inplace;                   // coloured star sweep in place (c == a): only nodes of `colour` are visited
    int wide;                      // executed by every CTA of the cluster (else by CTA 0 alone)
    int next_wide;                 // `wide` of the following command (1 after the last): selects the barrier
    int bc;                        // SMOOTH_POINT (scalar complex, Robin): the sweep also writes the slaved columns x = 0, n-1
    int scalar1;                   // SMOOTH_POINT: single-slot point smoother (the register-cached scalar path applies)
    double omega;
    const void* sm;                // SmootherTable<T>
    const void* w;                 // W tables
    void* a[MG_MAX_FIELDS];        // source solution (smooth, residual*) / fine source (restrict) / coarse source (prolong)
    void* b[MG_MAX_FIELDS];        // right-hand side of the level the residual is formed on
    void* c[MG_MAX_FIELDS];        // destination
};

static_assert(sizeof(TailCmd) % 16 == 0, "commands are fetched in 16-byte words");
constexpr int TAIL_CMD_WORDS = sizeof(TailCmd) / 16;

struct StarCoef {                   // (2d+1)-point star, canonical order z-, y-, x-, c, x+, y+, z+
    double c, x, y, z;
};

struct TailLevel {
    Geom g;
    const void* op;                // OperatorTable<T>
    cplx bc_factor;                // Helmholtz Robin factor 1 / (1 - i k h)
    StarCoef st;
};

template <int D>
__device__ __forceinline__ double star_apply(const StarCoef& s, const double* u, long long idx, const Geom& g) {
    double a;
    if (D == 3) {
        a = s.z * u[idx - g.plane];
        a = a + s.y * u[idx - g.pitch];
    } else {
        a = s.y * u[idx - g.pitch];
    }
    a = a + s.x * u[idx - 1];
    a = a + s.c * u[idx];
    a = a + s.x * u[idx + 1];
    a = a + s.y * u[idx + g.pitch];
    if (D == 3) a = a + s.z * u[idx + g.plane];
    return a;
}

struct TailHeader {
    int n_cmds;
    int cgs_kind, cgs_max_its, robin;
    double cgs_reduction;
    void* gwork;                   // global Krylov scratch when the vectors do not fit into shared memory
    int cgs_in_smem;
    int pad_[3];
    TailLevel lv[TAIL_MAX_LEVELS];
};
static_assert(sizeof(TailHeader) % 16 == 0, "the header is copied to shared memory in 16-byte words");

struct TailProgram {
    TailHeader h;
    TailCmd cmd[TAIL_MAX_CMDS];    // only the first h.n_cmds are uploaded
};

__device__ __forceinline__ uint4 ld_cmd_word(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

template <typename T>
__device__ __forceinline__ FieldSet<T> tail_fields(void* const* p) {
    FieldSet<T> fs;
#pragma unroll
    for (int i = 0; i < MG_MAX_FIELDS; ++i) fs.p[i] = static_cast<T*>(p[i]);
    return fs;
}

template <int D>
__device__ __forceinline__ void tail_decode(int i, int n, int& x, int& y, int& z) {
    x = i % n;
    int t = i / n;
    if (D == 3) {
        y = t % n;
        z = t / n;
    } else {
        y = t;
        z = 0;
    }
}

__device__ __forceinline__ unsigned cluster_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ unsigned cluster_size() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ inline void tail_bc(const Geom&, double*, cplx) {}
__device__ inline void tail_bc(const Geom& g, cplx* u, cplx c) {
    for (int t = threadIdx.x; t < g.n; t += TAIL_NT) {
        long long row = (long long)t * g.pitch;
        u[row] = t_mul(u[row + 1], c);
        u[row + g.n - 1] = t_mul(u[row + g.n - 2], c);
    }
    __syncthreads();
    for (int t = threadIdx.x; t < g.n; t += TAIL_NT) {
        u[t] = make_cplx(0.0, 0.0);
        u[(long long)(g.n - 1) * g.pitch + t] = make_cplx(0.0, 0.0);
    }
}

// LEAN: only the commands of a scalar real star-stencil cycle (star sweeps, star residual /
// restriction, zero, prolongation, CG).  The full interpreter is ~200 KB of code, more than the
// instruction cache holds, and since every step runs a different part of it each step paid
// ~1000 cycles of instruction fetch (clock64, even for a zero fill); the lean variant is what a
// Poisson cycle needs and stays cache resident.  The host picks it when the program qualifies
// (tail_is_lean in mg_exec.cuh).
template <typename T, int D, int F, bool LEAN>
__global__ void __launch_bounds__(TAIL_NT) k_tail(const TailProgram* __restrict__ gprog) {
    pdl_sync();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ T s8[8];
    __shared__ __align__(16) unsigned char s_head[sizeof(TailHeader)];
    __shared__ __align__(16) uint4 s_cmd[TAIL_MAX_CMDS * TAIL_CMD_WORDS];
#ifdef MG_TAIL_TIMING
    __shared__ long long t_step[TAIL_MAX_CMDS + 1];
#endif
    {
        // the whole program (header + commands in use) in one coalesced pass: one L2 latency in total
        // (fetching command i+1 during step i put an L2 round trip, ~1200 cycles, under every step)
        const uint4* src = reinterpret_cast<const uint4*>(gprog);
        uint4* dst = reinterpret_cast<uint4*>(s_head);
        const int n = reinterpret_cast<const int*>(gprog)[0];           // TailHeader::n_cmds
        for (int k = threadIdx.x; k < (int)(sizeof(TailHeader) / 16); k += TAIL_NT) dst[k] = src[k];
        const uint4* csrc = reinterpret_cast<const uint4*>(&gprog->cmd[0]);
        for (int k = threadIdx.x; k < n * TAIL_CMD_WORDS; k += TAIL_NT) s_cmd[k] = csrc[k];
    }
    __syncthreads();
    // LEAN selects a small interpreter (instruction fetch is what a command costs: the full one is ~400 KB of code):
    // real data: star sweeps / star residual + restriction / CG only; complex data: the register-cached scalar
    // paths and BiCGStab only (the Helmholtz cycle)
    constexpr bool LEAN_STAR = LEAN && sizeof(T) == sizeof(double) && F == 1;
    // (a lean interpreter for two-field real cycles -- generic node functions + CG only -- was measured: no gain for the
    // elasticity configuration, 5.87 against 5.79 ms per solve, so those cycles use the full one)
    constexpr bool GENERIC = !LEAN;
    const TailHeader& prog = *reinterpret_cast<const TailHeader*>(s_head);
    const int n_cmds = prog.n_cmds;
    const int n_ctas = (int)cluster_size(), rank = (int)cluster_rank();
    for (int ci = 0; ci < n_cmds; ++ci) {
#ifdef MG_TAIL_TIMING
        if (threadIdx.x == 0) t_step[ci] = clock64();
#endif
        const TailCmd& cmd = reinterpret_cast<const TailCmd*>(s_cmd)[ci];
        const bool wide = cmd.wide && n_ctas > 1;
        const bool next_wide = cmd.next_wide && n_ctas > 1;
        if (!wide && rank != 0) {                               // CTA 0 runs this command alone
            if (next_wide) cluster_barrier();
            continue;
        }
        const int first = (wide ? rank * TAIL_NT : 0) + (int)threadIdx.x;
        const int stride = wide ? n_ctas * TAIL_NT : TAIL_NT;
        const Geom g = prog.lv[cmd.lv].g;
        const OperatorTable<T>* op = static_cast<const OperatorTable<T>*>(prog.lv[cmd.lv].op);
        const int nodes = (D == 3) ? g.n * g.n * g.nz : g.n * g.n;
        const FieldSet<T> A = tail_fields<T>(cmd.a), B = tail_fields<T>(cmd.b), C = tail_fields<T>(cmd.c);
        int x, y, z;
        switch (cmd.kind) {
            case T_SMOOTH_POINT: {
                if constexpr (F == 1 && sizeof(T) == sizeof(double)) {
                    if (cmd.star && cmd.inplace) {                  // u += w (f - A u) on the nodes of the colour only
                        const StarCoef st = prog.lv[cmd.lv].st;
                        const double* rhs = static_cast<const double*>(cmd.b[0]);
                        double* u = static_cast<double*>(cmd.c[0]);
                        // n - 1 = 2^L: slots of (n-1)/2 colour nodes per interior row (y, z), decoded by shifts
                        const int half = (g.n - 1) >> 1, sh = 31 - __clz(half);
                        const int rows = (D == 3) ? (g.n - 2) * (g.nz - 2) : g.n - 2;
#pragma unroll 4
                        for (int i = first; i < rows * half; i += stride) {
                            const int r = i >> sh, k = i & (half - 1);
                            y = 1 + ((D == 3) ? r % (g.n - 2) : r);
                            z = (D == 3) ? 1 + r / (g.n - 2) : 0;
                            x = 1 + 2 * k + ((1 + y + z + g.z0 + cmd.colour) & 1);
                            if (x > g.n - 2 || (D == 3 && (z + g.z0 < 1 || z + g.z0 > g.n - 2))) continue;
                            const long long idx = node_index<D>(g, x, y, z);
                            u[idx] = u[idx] + cmd.omega * (rhs[idx] - star_apply<D>(st, u, idx, g));
                        }
                        break;
                    }
                    if (cmd.star) {                                 // dst = src + w (f - A src) on the colour
                        const StarCoef st = prog.lv[cmd.lv].st;
                        const double* src = static_cast<const double*>(cmd.a[0]);
                        const double* rhs = static_cast<const double*>(cmd.b[0]);
                        double* dst = static_cast<double*>(cmd.c[0]);
#pragma unroll 4
                        for (int i = first; i < nodes; i += stride) {
                            tail_decode<D>(i, g.n, x, y, z);
                            const long long idx = node_index<D>(g, x, y, z);
                            double v = src[idx];
                            if (is_interior<D>(g, x, y, z) && (cmd.colour < 0 || ((x + y + z + g.z0) & 1) == cmd.colour))
                                v = v + cmd.omega * (rhs[idx] - star_apply<D>(st, src, idx, g));
                            dst[idx] = v;
                        }
                        break;
                    }
                }
                if constexpr (!LEAN_STAR) {
                    const SmootherTable<T>* sm = static_cast<const SmootherTable<T>*>(cmd.sm);
                    const T* w = static_cast<const T*>(cmd.w);
                    if constexpr (F == 1) {
                        // scalar operator with 5 entries (the complex Helmholtz star): entries and
                        // weight in registers, the arithmetic of smooth_point_node for one slot.  bc: the slaved Robin
                        // columns are written by the thread that updates their interior neighbour (the rows y = 0, n-1
                        // are zero and are never written), which replaces the T_BC command after the sweep
                        const int ne = op->n_entries[0][0];
                        if (!is_variable<T, F>(op) && ne == 5 && D == 2) {
                            const T w0 = w[(long long)sm->mask_index[1]];
                            const T* src = A.p[0];                      // == dst for an in-place colour pass
                            const T* __restrict__ rhs = B.p[0];
                            T* dst = C.p[0];
                            const bool bc = cmd.bc != 0;
                            const cplx bcf = prog.lv[cmd.lv].bc_factor;
                            auto sweep = [&](auto ne_tag) {
                                constexpr int NE = decltype(ne_tag)::value;
                                int off[NE];
                                T val[NE];
                                load_row_entries<T, NE>(op, off, val);
                                if (cmd.inplace) {
                                    // colour pass in place (dst == src, set by the host for the 5-point star only): the
                                    // nodes of the colour alone, decoded by shifts (n - 1 = 2^L) -- half the nodes of the
                                    // out-of-place pass, no copy of the others
                                    const int half = (g.n - 1) >> 1, sh = 31 - __clz(half);
                                    for (int i = first; i < (g.n - 2) * half; i += stride) {
                                        y = 1 + (i >> sh);
                                        x = 1 + 2 * (i & (half - 1)) + ((1 + y + g.z0 + cmd.colour) & 1);
                                        if (x > g.n - 2) continue;
                                        const long long idx = node_index<D>(g, x, y, 0);
                                        const T r = t_sub(rhs[idx], apply_cached<T, NE>(off, val, dst, idx));
                                        const T v = t_add(dst[idx], t_mul(w0, r));
                                        dst[idx] = v;
                                        if constexpr (sizeof(T) == sizeof(cplx)) {
                                            if (bc && x == 1) dst[idx - 1] = t_mul(v, bcf);
                                            if (bc && x == g.n - 2) dst[idx + 1] = t_mul(v, bcf);
                                        }
                                    }
                                    return;
                                }
                                for (int i = first; i < nodes; i += stride) {
                                    tail_decode<D>(i, g.n, x, y, z);
                                    const long long idx = node_index<D>(g, x, y, z);
                                    const bool active = is_interior<D>(g, x, y, z) &&
                                                        (cmd.colour < 0 || ((x + y + z + g.z0) & 1) == cmd.colour);
                                    if (!active) {
                                        // a slaved column node is written by the thread of its interior neighbour
                                        // when that one is updated in this pass
                                        bool slaved = false;
                                        if (bc && (x == 0 || x == g.n - 1) && y >= 1 && y <= g.n - 2) {
                                            const int xn = x == 0 ? 1 : g.n - 2;
                                            slaved = cmd.colour < 0 || ((xn + y + g.z0) & 1) == cmd.colour;
                                        }
                                        if (!slaved) dst[idx] = src[idx];
                                        continue;
                                    }
                                    const T r = t_sub(rhs[idx], apply_cached<T, NE>(off, val, src, idx));
                                    const T v = t_add(src[idx], t_mul(w0, r));
                                    dst[idx] = v;
                                    if constexpr (sizeof(T) == sizeof(cplx)) {
                                        if (bc && x == 1) dst[idx - 1] = t_mul(v, bcf);
                                        if (bc && x == g.n - 2) dst[idx + 1] = t_mul(v, bcf);
                                    }
                                }
                            };
                            sweep(std::integral_constant<int, 5>{});
                            break;
                        }
                    }
                    if constexpr (GENERIC)
                        for (int i = first; i < nodes; i += stride) {
                            tail_decode<D>(i, g.n, x, y, z);
                            smooth_point_node<T, D, F>(g, op, sm, w, A, B, C, cmd.colour, x, y, z);
                        }
                }
                break;
            }
            case T_SMOOTH_TILE: if constexpr (!LEAN) {
                const SmootherTable<T>* sm = static_cast<const SmootherTable<T>*>(cmd.sm);
                const T* w = static_cast<const T*>(cmd.w);
                const int t0 = (g.n - 1) / sm->period[0] + 1, t1 = (g.n - 1) / sm->period[1] + 1;
                const int t2 = (D == 3) ? (g.n - 1) / sm->period[2] + 1 : 1;
                for (int i = first; i < t0 * t1 * t2; i += stride) {
                    int tx = i % t0, r = i / t0;
                    smooth_tile_node<T, D, F>(g, op, sm, w, A, B, C, tx, r % t1, r / t1);
                }
                break;
            } else break;
            case T_RESIDUAL:
                if constexpr (F == 1 && sizeof(T) == sizeof(double)) {
                    if (cmd.star) {
                        const StarCoef st = prog.lv[cmd.lv].st;
                        const double* src = static_cast<const double*>(cmd.a[0]);
                        const double* rhs = static_cast<const double*>(cmd.b[0]);
                        double* dst = static_cast<double*>(cmd.c[0]);
#pragma unroll 4
                        for (int i = first; i < nodes; i += stride) {
                            tail_decode<D>(i, g.n, x, y, z);
                            if (!is_interior<D>(g, x, y, z)) continue;
                            const long long idx = node_index<D>(g, x, y, z);
                            dst[idx] = rhs[idx] - star_apply<D>(st, src, idx, g);
                        }
                        break;
                    }
                }
                if constexpr (!LEAN_STAR) {
                    if constexpr (F == 1) {
                        const int ne = op->n_entries[0][0];
                        if (!is_variable<T, F>(op) && ne == 5) {
                            auto pass = [&](auto ne_tag) {
                                constexpr int NE = decltype(ne_tag)::value;
                                int off[NE];
                                T val[NE];
                                load_row_entries<T, NE>(op, off, val);
                                for (int i = first; i < nodes; i += stride) {
                                    tail_decode<D>(i, g.n, x, y, z);
                                    if (!is_interior<D>(g, x, y, z)) continue;
                                    const long long idx = node_index<D>(g, x, y, z);
                                    C.p[0][idx] = t_sub(B.p[0][idx], apply_cached<T, NE>(off, val, A.p[0], idx));
                                }
                            };
                            pass(std::integral_constant<int, 5>{});
                            break;
                        }
                    }
                    if constexpr (GENERIC)
                        for (int i = first; i < nodes; i += stride) {
                            tail_decode<D>(i, g.n, x, y, z);
                            residual_node<T, D, F>(g, op, A, B, C, x, y, z);
                        }
                }
                break;
            case T_RESTRICT:
            case T_RESIDUAL_RESTRICT: {
                const Geom gf = prog.lv[cmd.lv + 1].g;
                const OperatorTable<T>* opf = static_cast<const OperatorTable<T>*>(prog.lv[cmd.lv + 1].op);
                if constexpr (F == 1 && sizeof(T) == sizeof(double)) {
                    if (cmd.star && cmd.kind == T_RESIDUAL_RESTRICT) {      // same loop order as residual_restrict_node
                        const StarCoef st = prog.lv[cmd.lv + 1].st;
                        const double* src = static_cast<const double*>(cmd.a[0]);
                        const double* rhs = static_cast<const double*>(cmd.b[0]);
                        double* dst = static_cast<double*>(cmd.c[0]);
                        const double w1[3] = {0.25, 0.5, 0.25};
#pragma unroll 2
                        for (int i = first; i < nodes; i += stride) {
                            tail_decode<D>(i, g.n, x, y, z);
                            if (!is_interior<D>(g, x, y, z)) continue;
                            const int zf = (D == 3) ? 2 * (z + g.z0) - gf.z0 : 0;
                            double acc = 0.0;
                            bool first = true;
#pragma unroll
                            for (int oz = (D == 3 ? -1 : 0); oz <= (D == 3 ? 1 : 0); ++oz)
#pragma unroll
                                for (int oy = -1; oy <= 1; ++oy)
#pragma unroll
                                    for (int ox = -1; ox <= 1; ++ox) {
                                        double wgt = w1[ox + 1] * w1[oy + 1];
                                        if (D == 3) wgt = w1[oz + 1] * wgt;
                                        const long long fidx = node_index<D>(gf, 2 * x + ox, 2 * y + oy, zf + oz);
                                        const double term = wgt * (rhs[fidx] - star_apply<D>(st, src, fidx, gf));
                                        acc = first ? term : acc + term;
                                        first = false;
                                    }
                            dst[node_index<D>(g, x, y, z)] = acc;
                        }
                        break;
                    }
                }
                if constexpr (!LEAN_STAR && F == 1 && D == 2) {
                    const int ne = opf->n_entries[0][0];
                    if (cmd.kind == T_RESIDUAL_RESTRICT && !is_variable<T, F>(opf) && ne == 5) {
                        // residual_restrict_node with the fine operator's entries in registers (same order of the
                        // nine terms, same accumulation inside each)
                        auto pass = [&](auto ne_tag) {
                            constexpr int NE = decltype(ne_tag)::value;
                            int off[NE];
                            T val[NE];
                            load_row_entries<T, NE>(opf, off, val);
                            const double w1[3] = {0.25, 0.5, 0.25};
                            for (int i = first; i < nodes; i += stride) {
                                tail_decode<D>(i, g.n, x, y, z);
                                if (!is_interior<D>(g, x, y, z)) continue;
                                T acc = t_zero(T());
                                bool first_term = true;
#pragma unroll
                                for (int oy = -1; oy <= 1; ++oy)
#pragma unroll
                                    for (int ox = -1; ox <= 1; ++ox) {
                                        const double wgt = w1[ox + 1] * w1[oy + 1];
                                        const long long fidx = node_index<D>(gf, 2 * x + ox, 2 * y + oy, 0);
                                        const T r = t_sub(B.p[0][fidx], apply_cached<T, NE>(off, val, A.p[0], fidx));
                                        const T term = t_scale(wgt, r);
                                        acc = first_term ? term : t_add(acc, term);
                                        first_term = false;
                                    }
                                C.p[0][node_index<D>(g, x, y, z)] = acc;
                            }
                        };
                        pass(std::integral_constant<int, 5>{});
                        break;
                    }
                }
                for (int i = first; i < nodes; i += stride) {
                    tail_decode<D>(i, g.n, x, y, z);
                    if (cmd.kind == T_RESTRICT) restrict_node<T, D, F>(g, gf, A, C, x, y, z);
                    else if constexpr (GENERIC) residual_restrict_node<T, D, F>(g, gf, opf, A, B, C, x, y, z);
                }
                break;
            }
            case T_ZERO:
                for (int f = 0; f < F; ++f)
                    for (long long i = first; i < g.total; i += stride) C.p[f][i] = t_zero(T());
                break;
            case T_PROLONG: {
                const Geom gc = prog.lv[cmd.lv - 1].g;
                if constexpr (F == 1 && sizeof(T) == sizeof(double)) {
                    // branch-free form of prolong_node: all corner candidates are loaded (a coarse node is
                    // its own candidate where the fine index is even) and added in the order (cz, cy, cx)
                    const double* e = static_cast<const double*>(cmd.a[0]);
                    double* u = static_cast<double*>(cmd.c[0]);
                    const int sh = 31 - __clz(g.n - 1);
                    const int rows = (D == 3) ? (g.n - 2) * (g.nz - 2) : g.n - 2;
#pragma unroll 2
                    for (int i = first; i < (rows << sh); i += stride) {
                        const int r = i >> sh;
                        x = 1 + (i & (g.n - 2));
                        y = 1 + ((D == 3) ? r % (g.n - 2) : r);
                        z = (D == 3) ? 1 + r / (g.n - 2) : 0;
                        if (x > g.n - 2 || (D == 3 && (z + g.z0 < 1 || z + g.z0 > g.n - 2))) continue;
                        const int zg = z + g.z0;
                        const int px = x & 1, py = y & 1, pz = (D == 3) ? (zg & 1) : 0;
                        const long long c = node_index<D>(gc, x >> 1, y >> 1, (D == 3) ? (zg >> 1) - gc.z0 : 0);
                        const long long dy = py ? gc.pitch : 0, dz = pz ? gc.plane : 0;
                        double acc = e[c];
                        const double c100 = e[c + px], c010 = e[c + dy], c110 = e[c + dy + px];
                        if (px) acc = acc + c100;
                        if (py) {
                            acc = acc + c010;
                            if (px) acc = acc + c110;
                        }
                        if (D == 3) {
                            const double c001 = e[c + dz], c101 = e[c + dz + px], c011 = e[c + dz + dy],
                                         c111 = e[c + dz + dy + px];
                            if (pz) {
                                acc = acc + c001;
                                if (px) acc = acc + c101;
                                if (py) {
                                    acc = acc + c011;
                                    if (px) acc = acc + c111;
                                }
                            }
                        }
                        const int k = px + py + pz;
                        const double pe = k == 0 ? acc : (k == 1 ? 0.5 : (k == 2 ? 0.25 : 0.125)) * acc;
                        const long long idx = node_index<D>(g, x, y, z);
                        u[idx] = cmd.add ? u[idx] + cmd.omega * pe : pe;
                    }
                    break;
                }
                for (int i = first; i < nodes; i += stride) {
                    tail_decode<D>(i, g.n, x, y, z);
                    prolong_node<T, D, F>(g, gc, A, C, cmd.omega, cmd.add, x, y, z);
                }
                break;
            }
            case T_CGS: {
                T* work = prog.cgs_in_smem ? reinterpret_cast<T*>(smem_raw) : static_cast<T*>(prog.gwork);
                if constexpr (LEAN_STAR) {
                    coarse_cg_body<T, D, F, true>(g, op, C, B, work, prog.cgs_max_its, prog.cgs_reduction, s8);
                } else if constexpr (LEAN) {
                    coarse_bicgstab_body<T, D, F>(g, op, C, B, work, prog.cgs_max_its, prog.cgs_reduction, prog.robin,
                                                  prog.lv[cmd.lv].bc_factor, s8);
                } else {
                    if (prog.cgs_kind == MG_CGS_CG)
                        coarse_cg_body<T, D, F>(g, op, C, B, work, prog.cgs_max_its, prog.cgs_reduction, s8);
                    else
                        coarse_bicgstab_body<T, D, F>(g, op, C, B, work, prog.cgs_max_its, prog.cgs_reduction, prog.robin,
                                                      prog.lv[cmd.lv].bc_factor, s8);
                }
                break;
            }
            case T_BC:
                if constexpr (!LEAN_STAR)
                    for (int f = 0; f < F; ++f) tail_bc(g, C.p[f], prog.lv[cmd.lv].bc_factor);
                break;
            default: break;
        }
        if (n_ctas > 1 && (wide || next_wide)) cluster_barrier();
        else __syncthreads();
    }
#ifdef MG_TAIL_TIMING
    if (threadIdx.x == 0) {
        t_step[n_cmds] = clock64();
        for (int ci = 0; ci < n_cmds; ++ci)
            printf("step %d kind %d lv %d colour %d star %d cycles %lld\n", ci, gprog->cmd[ci].kind, gprog->cmd[ci].lv,
                   gprog->cmd[ci].colour, gprog->cmd[ci].star, t_step[ci + 1] - t_step[ci]);
    }
#endif
}

}  // namespace mg
