// mg_kernels_fast.cuh -- dispatch of the specialised stage kernels (scalar real star stencils:
// 2D/3D Poisson).  Each fast_* hook returns true when it issued the stage; false falls through to
// the generic kernel.  Results are bit-identical to the generic kernels (same operand order, no
// FMA contraction), which tests/test_gpu_parity.py checks against the oracle on both paths.
#pragma once
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "mg_fast2d.cuh"
#include "mg_fast3d.cuh"
#include "mg_rbgs3d.cuh"
#include "mg_rbgs3d2.cuh"
#include "mg_restrict3d.cuh"
#include "mg_varcoef.cuh"
#include "mg_workspace.cuh"

namespace mg {

constexpr int FAST_NT = 128;            // threads per CTA of the marching kernels
constexpr int MAX_PARTIALS = 16384;     // capacity of Level::partial

inline int imax_h(int a, int b) { return a > b ? a : b; }
inline int imin_h(int a, int b) { return a < b ? a : b; }

inline int tunable(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : dflt;
}

inline void fast_count(mg_workspace* ws, int k = 1) { ws->pending_count += k; }

template <int D>
inline bool star_level_ok(const mg_workspace* ws, const Level& lv, int cols_per_thread) {
    return D == 2 && ws->dtype == MG_F64 && ws->n_fields == 1 && lv.star &&
           (lv.g.n - 1) % (cols_per_thread * FAST_NT) == 0;
}

inline Star2D star2d(const Level& lv) {
    Star2D s;
    s.cs = lv.star_y; s.cw = lv.star_x; s.cc = lv.star_c;
    return s;
}

inline dim3 strip_grid(const Level& lv, int cols_per_thread, int rows, int n_rows) {
    return dim3((lv.g.n - 1) / (cols_per_thread * FAST_NT), (n_rows + rows - 1) / rows, 1);
}

// rows per CTA of the tile kernels: 4 (measured best at 4096^2: R = 8 costs occupancy, R = 2 halo
// traffic) unless the level is too small to give ~7 CTAs per SM, then 2
inline int pick_rows(const Level& lv, int cols_per_thread, int max_ctas = 1 << 30) {
    const int strips = (lv.g.n - 1) / (cols_per_thread * FAST_NT);
    const int n_rows = lv.g.n - 2;
    for (int r = 4; r >= 2; r >>= 1) {
        const int ctas = strips * ((n_rows + r - 1) / r);
        if ((ctas >= 1024 || r == 2) && ctas <= max_ctas) return r;
    }
    return 8;   // only when max_ctas forces it
}

// rows per CTA of the marching (restriction / prolongation) kernels: 16 when the level is large
// enough to fill the machine, fewer below
inline int march_rows(int n_rows, int strips) {
    int r = 16;
    while (r > 2 && strips * ((n_rows + r - 1) / r) < 1024) r >>= 1;
    return r;
}

template <int MODE>
inline void launch_star2d(mg_workspace* ws, Level& lv, int r, const double* u, const double* f, double* out, double w,
                          double* partial) {
    dim3 grid = strip_grid(lv, 2, r, lv.g.n - 2);
    if (r == 8) launch_k(ws, k_star2d<MODE, FAST_NT, 8>, grid, FAST_NT, 0, lv.g, star2d(lv), u, f, out, w, partial);
    else if (r == 4) launch_k(ws, k_star2d<MODE, FAST_NT, 4>, grid, FAST_NT, 0, lv.g, star2d(lv), u, f, out, w, partial);
    else launch_k(ws, k_star2d<MODE, FAST_NT, 2>, grid, FAST_NT, 0, lv.g, star2d(lv), u, f, out, w, partial);
}

inline int valid_rows(int r, int dflt) { return (r == 2 || r == 4 || r == 8) ? r : dflt; }

// ---- 2D variable-coefficient Poisson (mg_varcoef.cuh): levels whose columns tile into strips -----------
template <int D>
inline bool vc_level_ok(const mg_workspace* ws, const Level& lv) {
    static const int enabled = tunable("MG_FAST_VC", 1);
    return enabled && D == 2 && ws->vc_kind == MG_VC_EXP_KAPPA && ws->dtype == MG_F64 && ws->n_fields == 1 &&
           (lv.g.n - 1) % (2 * FAST_NT) == 0;
}

template <int MODE>
inline dim3 launch_vcstar2d(mg_workspace* ws, Level& lv, int r, const double* u, const double* f, double* out, double w,
                            int colour, double* partial) {
    VC2D s;
    s.kappa = ws->vc_kappa;
    s.h = 1.0 / (double)(lv.g.n - 1);
    s.ihh = (double)(lv.g.n - 1) * (double)(lv.g.n - 1);
    dim3 grid = strip_grid(lv, 2, r, lv.g.n - 2);
    if (r == 4) launch_k(ws, k_vcstar2d<MODE, FAST_NT, 4>, grid, FAST_NT, 0, lv.g, s, u, f, out, w, colour, partial);
    else launch_k(ws, k_vcstar2d<MODE, FAST_NT, 2>, grid, FAST_NT, 0, lv.g, s, u, f, out, w, colour, partial);
    return grid;
}

// block-Jacobi tiles (px x py <= 4 unknowns) on a scalar star stencil: any level size
template <typename T, int D, int F>
inline bool fast_tile_smooth(mg_workspace* ws, Level& lv, Smoother& sm) {
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        static const int enabled = tunable("MG_FAST_TILE", 1);
        if (!enabled || !(ws->dtype == MG_F64 && ws->n_fields == 1 && lv.star) || sm.point || !sm.jacobi || sm.colored)
            return false;
        const int px = sm.period[0], py = sm.period[1];
        if (sm.m != px * py || sm.m > 4 || sm.full_index < 0 || !sm.canonical_slots) return false;
        Tile2D t;
        t.px = px; t.py = py; t.m = sm.m;
        t.cw = (2 * FAST_NT) / (px == 3 ? 6 : 1) * (px == 3 ? 6 : 1);   // even multiple of px, <= 2*NT
        t.rb = py == 3 ? 3 : 4;
        t.full_index = sm.full_index;
        static const int pf_env = tunable("MG_TILE_PF", 20);
        t.pf = lv.g.n > 2049 ? pf_env : 0;
        dim3 grid((lv.g.n - 1 + t.cw - 1) / t.cw, (lv.g.n + t.rb - 1) / t.rb, 1);
        const SmootherTable<double>* tab = static_cast<const SmootherTable<double>*>(sm.d_table);
        const double* w = static_cast<const double*>(sm.d_w);
        const double* src = static_cast<const double*>(lv.sol[lv.cur[0]][0]);
        const double* rhs = static_cast<const double*>(lv.rhs[0]);
        double* dst = static_cast<double*>(lv.sol[lv.cur[0] ^ 1][0]);
#define MG_TILE(PX_, PY_) launch_k(ws, k_tile2d<FAST_NT, PX_, PY_>, grid, FAST_NT, 0, lv.g, star2d(lv), t, tab, w, src, rhs, dst)
        if (px == 1 && py == 2) MG_TILE(1, 2);
        else if (px == 2 && py == 1) MG_TILE(2, 1);
        else if (px == 1 && py == 3) MG_TILE(1, 3);
        else if (px == 3 && py == 1) MG_TILE(3, 1);
        else if (px == 2 && py == 2) MG_TILE(2, 2);
        else if (px == 1 && py == 4) MG_TILE(1, 4);
        else if (px == 4 && py == 1) MG_TILE(4, 1);
        else return false;
#undef MG_TILE
        lv.cur[0] ^= 1;
        fast_count(ws);
        return true;
    }
    return false;
}

// ---- 3D: 7-point star, whole grids and z-slabs ------------------------------------------------------
inline Star3D star3d(const Level& lv) {
    Star3D s;
    s.cz = lv.star_z; s.cy = lv.star_y; s.cx = lv.star_x; s.cc = lv.star_c;
    return s;
}

// threads per CTA of the 3D kernels: the strip of 2*NT columns must tile [0, n-1)
inline int star3d_nt(const mg_workspace* ws, const Level& lv) {
    if (!(ws->dim == 3 && ws->dtype == MG_F64 && ws->n_fields == 1 && lv.star)) return 0;
    static const int enabled = tunable("MG_FAST_3D", 1);
    if (!enabled) return 0;
    for (int nt = 128; nt >= 32; nt >>= 1)
        if ((lv.g.n - 1) % (2 * nt) == 0) return nt;
    return 0;
}

template <int MODE>
inline dim3 launch_star3d(mg_workspace* ws, Level& lv, int nt, const double* u, const double* f, double* out, double w,
                          int colour, double* partial) {
    static const int ry_env = tunable("MG_3D_RY", 0), zc_env = tunable("MG_3D_ZCHUNK", 0);
    const int ry = (ry_env == 2 || ry_env == 4) ? ry_env : 4;
    const int strips = (lv.g.n - 1) / (2 * nt), bands = (lv.g.n - 2 + ry - 1) / ry, planes = lv.g.nz - 2;
    int zchunk = zc_env > 0 ? zc_env : 32;
    while (zchunk > 4 && strips * bands * ((planes + zchunk - 1) / zchunk) < 1024) zchunk >>= 1;
    dim3 grid(strips, bands, (planes + zchunk - 1) / zchunk);
    const Star3D s = star3d(lv);
#define MG_L3(NT_, RY_) launch_k(ws, k_star3d<MODE, NT_, RY_>, grid, NT_, 0, lv.g, s, u, f, out, w, colour, zchunk, partial)
    if (ry == 4) {
        if (nt == 128) MG_L3(128, 4); else if (nt == 64) MG_L3(64, 4); else MG_L3(32, 4);
    } else {
        if (nt == 128) MG_L3(128, 2); else if (nt == 64) MG_L3(64, 2); else MG_L3(32, 2);
    }
#undef MG_L3
    return grid;
}

// ---- TMA tensor maps of whole 3D field arrays (x fastest; box = 64 columns x `rows` rows x 1 plane) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// descriptors are cached by (array, nodes per dimension, planes, box rows): the sweeps of a solve reuse three of them.
// `base` / `planes`: first plane and plane count of the view (a z-slab view starts one plane before the array:
// the second ghost plane, mg_engine.cu create_impl).
inline const CUtensorMap* plane_tensor_map(const double* base, const Geom& g, int rows, int planes = -1) {
    static std::mutex lock;
    static std::map<std::tuple<const void*, int, int, int>, CUtensorMap> cache;
    std::lock_guard<std::mutex> guard(lock);
    if (planes < 0) planes = g.nz;
    const auto key = std::make_tuple(static_cast<const void*>(base), g.n, planes, rows);
    auto it = cache.find(key);
    if (it != cache.end()) return &it->second;
    EncodeTiledFn encode = encode_tiled_fn();
    if (!encode) return nullptr;
    CUtensorMap tm;
    const cuuint64_t dims[3] = {(cuuint64_t)g.n, (cuuint64_t)g.n, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)g.pitch * sizeof(double), (cuuint64_t)g.plane * sizeof(double)};
    const cuuint32_t box[3] = {(cuuint32_t)RB3_W, (cuuint32_t)rows, 1u};
    const cuuint32_t elem[3] = {1u, 1u, 1u};
    if (encode(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, elem,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return nullptr;
    return &cache.emplace(key, tm).first->second;
}

// How the TMA-staged kernels see a 3D level: whole grids as they are; a z-slab with two ghost planes
// (Level::wide_ghost) from one plane before the array, plane 0 of the view = global plane zoff.  zlo .. zhi =
// global planes an out-of-place sweep must write: the owned ones, plus the global Dirichlet plane that the
// first / last slab holds as its local plane 0 / nz-1.
struct SlabView {
    long long shift;        // elements between the view's first plane and the array's
    int planes, zoff, zlo, zhi;
};
inline SlabView slab_view(const Level& lv) {
    SlabView v;
    const Geom& g = lv.g;
    if (!lv.wide_ghost) {
        v.shift = 0; v.planes = g.nz; v.zoff = g.z0; v.zlo = g.z0; v.zhi = g.z0 + g.nz - 1;
        return v;
    }
    v.shift = g.plane; v.planes = g.nz + 2; v.zoff = g.z0 - 1;
    v.zlo = g.z0 == 0 ? 0 : g.z0 + 1;
    v.zhi = g.z0 + g.nz - 1 == g.n - 1 ? g.n - 1 : g.z0 + g.nz - 2;
    return v;
}
// the one-pass kernels apply to whole grids and to slabs with two ghost planes that own at least two planes
inline bool one_pass_geometry_ok(const Level& lv) {
    if (lv.g.z0 == 0 && lv.g.nz == lv.g.n) return true;
    return lv.wide_ghost && lv.g.nz - 2 >= 2;
}

// one-pass 3D RB-GS sweep (mg_rbgs3d.cuh): whole grids and z-slabs with two ghost planes
template <int RY, int NS, int NF>
inline bool launch_rbgs3d_cfg(mg_workspace* ws, Level& lv, const double* u, const double* f, double* out, double w) {
    static const int zc_env = tunable("MG_RBGS3D_ZCHUNK", 0), pf_env = tunable("MG_RBGS3D_PF", 0);
    static bool opted_in[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !opted_in[dev]) {
        if (cudaFuncSetAttribute(k_rbgs3d<RY, NS, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)rbgs3d_smem_bytes<RY, NS, NF>()) != cudaSuccess)
            return false;
        opted_in[dev] = true;
    }
    // z-slab with two ghost planes (Level::wide_ghost): the view starts one plane before the array
    const SlabView v = slab_view(lv);
    const CUtensorMap* tm_u = plane_tensor_map(u - v.shift, lv.g, RY + 4, v.planes);
    const CUtensorMap* tm_f = plane_tensor_map(f - v.shift, lv.g, RY + 2, v.planes);
    if (!tm_u || !tm_f) return false;
    const int n = lv.g.n, nzo = v.zhi - v.zlo + 1;
    const int gx = (n + RB3_TW - 1) / RB3_TW, gy = (n + RY - 1) / RY;
    int zchunk = zc_env >= 3 ? zc_env : 64;
    while (zc_env < 3 && zchunk > 4 && gx * gy * ((nzo + zchunk - 1) / zchunk) < 888) zchunk >>= 1;
    launch_k(ws, k_rbgs3d<RY, NS, NF>, dim3(gx, gy, (nzo + zchunk - 1) / zchunk), 32 * (RY + 2),
             rbgs3d_smem_bytes<RY, NS, NF>(), lv.g, star3d(lv), *tm_u, *tm_f, out - v.shift, w, zchunk, pf_env,
             v.zoff, v.zlo, v.zhi);
    return true;
}

// the same sweep with two tile rows per warp (mg_rbgs3d2.cuh)
template <int RY, int NS, int NF>
inline bool launch_rbgs3d2_cfg(mg_workspace* ws, Level& lv, const double* u, const double* f, double* out, double w) {
    static const int zc_env = tunable("MG_RBGS3D_ZCHUNK", 0), pf_env = tunable("MG_RBGS3D_PF", 0);
    static bool opted_in[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !opted_in[dev]) {
        if (cudaFuncSetAttribute(k_rbgs3d2<RY, NS, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)rbgs3d_smem_bytes<RY, NS, NF>()) != cudaSuccess)
            return false;
        opted_in[dev] = true;
    }
    const SlabView v = slab_view(lv);
    const CUtensorMap* tm_u = plane_tensor_map(u - v.shift, lv.g, RY + 4, v.planes);
    const CUtensorMap* tm_f = plane_tensor_map(f - v.shift, lv.g, RY + 2, v.planes);
    if (!tm_u || !tm_f) return false;
    const int n = lv.g.n, nzo = v.zhi - v.zlo + 1;
    const int gx = (n + RB3_TW - 1) / RB3_TW, gy = (n + RY - 1) / RY;
    // planes per chunk: as long as possible (every chunk pays the pipeline fill of four planes) with at least
    // two waves of CTAs, chunks of equal length (513 planes: three chunks of 171, 0.531 ms against 0.545 with 64)
    int zchunk = zc_env >= 3 ? zc_env : 256;
    while (zc_env < 3 && zchunk > 4 && gx * gy * ((nzo + zchunk - 1) / zchunk) < 888) zchunk >>= 1;
    if (zc_env < 3) {
        const int chunks = (nzo + zchunk - 1) / zchunk;
        zchunk = (nzo + chunks - 1) / chunks;
    }
    launch_k(ws, k_rbgs3d2<RY, NS, NF>, dim3(gx, gy, (nzo + zchunk - 1) / zchunk), 16 * (RY + 2),
             rbgs3d_smem_bytes<RY, NS, NF>(), lv.g, star3d(lv), *tm_u, *tm_f, out - v.shift, w, zchunk, pf_env,
             v.zoff, v.zlo, v.zhi);
    return true;
}

// would a red-black sweep of this level run as the one-pass kernel?  (the domain-decomposition driver
// exchanges two ghost planes once per sweep then, instead of one plane per colour pass)
inline bool rbgs3d_applicable(const mg_workspace* ws, const Level& lv) {
    static const int enabled = tunable("MG_RBGS3D", 1);
    const int min_n = tunable("MG_RBGS3D_MIN_N", 257);      // not cached: the tests lower it to 65
    return enabled && ws->use_fast && star3d_nt(ws, lv) && one_pass_geometry_ok(lv) && lv.g.n >= 65 && lv.g.n >= min_n;
}

inline bool launch_rbgs3d(mg_workspace* ws, Level& lv, const double* u, const double* f, double* out, double w) {
    static const int cfg = tunable("MG_RBGS3D_CFG", 0);
    static const int version = tunable("MG_RBGS3D_V", 2);
    if (!rbgs3d_applicable(ws, lv)) return false;
    if (version == 2) {
        switch (cfg) {
            case 1: return launch_rbgs3d2_cfg<14, 6, 5>(ws, lv, u, f, out, w);
            case 2: return launch_rbgs3d2_cfg<6, 7, 6>(ws, lv, u, f, out, w);
            case 3: return launch_rbgs3d2_cfg<12, 6, 5>(ws, lv, u, f, out, w);
            case 4: return launch_rbgs3d2_cfg<8, 6, 5>(ws, lv, u, f, out, w);
            case 5: return launch_rbgs3d2_cfg<8, 5, 4>(ws, lv, u, f, out, w);
            case 6: return launch_rbgs3d2_cfg<10, 6, 5>(ws, lv, u, f, out, w);
            case 7: return launch_rbgs3d2_cfg<12, 5, 4>(ws, lv, u, f, out, w);
            case 8: return launch_rbgs3d2_cfg<10, 5, 4>(ws, lv, u, f, out, w);
            default: return launch_rbgs3d2_cfg<8, 7, 6>(ws, lv, u, f, out, w);
        }
    }
    switch (cfg) {
        // rows per tile, ring sizes of u / f (measured at 513^3, profiles/README.md): the default keeps
        // 3 CTAs per SM (74 KB each) with 3 planes in flight
        case 1: return launch_rbgs3d_cfg<8, 6, 5>(ws, lv, u, f, out, w);
        case 2: return launch_rbgs3d_cfg<16, 6, 5>(ws, lv, u, f, out, w);
        case 3: return launch_rbgs3d_cfg<12, 6, 5>(ws, lv, u, f, out, w);
        default: return launch_rbgs3d_cfg<8, 7, 6>(ws, lv, u, f, out, w);
    }
}

template <typename T, int D, int F>
inline bool fast_smooth(mg_workspace* ws, Level& lv, Smoother& sm, int c_begin, int c_end, bool zero_in) {
    // zero_in: the current solution is a pending zero fill (Level::zero_pending); true is returned only
    // when the kernel ran without reading it
    if constexpr (D == 3 && F == 1 && sizeof(T) == sizeof(double)) {
        const int nt = star3d_nt(ws, lv);
        if (zero_in || !nt || !sm.point || sm.m != 1 || lv.g.nz < 3) return false;
        const double* rhs = static_cast<const double*>(lv.rhs[0]);
        if (sm.colored && c_begin == 0 && c_end == 2 &&
            launch_rbgs3d(ws, lv, static_cast<const double*>(lv.sol[lv.cur[0]][0]), rhs,
                          static_cast<double*>(lv.sol[lv.cur[0] ^ 1][0]), sm.w00)) {
            lv.cur[0] ^= 1;
            fast_count(ws);
            return true;
        }
        for (int c = c_begin; c < c_end; ++c) {
            const double* src = static_cast<const double*>(lv.sol[lv.cur[0]][0]);
            double* dst = static_cast<double*>(lv.sol[lv.cur[0] ^ 1][0]);
            launch_star3d<STAR3_SMOOTH>(ws, lv, nt, src, rhs, dst, sm.w00, sm.colored ? c : -1, nullptr);
            lv.cur[0] ^= 1;
            fast_count(ws);
        }
        return true;
    }
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        if (ws->vc_kind) {
            if (zero_in || !vc_level_ok<D>(ws, lv) || !sm.point || sm.m != 1) return false;
            static const int rows_vc = tunable("MG_VC_ROWS", 0);
            const int r = (rows_vc == 2 || rows_vc == 4) ? rows_vc : imin_h(4, pick_rows(lv, 2));
            const double* rhs = static_cast<const double*>(lv.rhs[0]);
            if (sm.colored) {
                // two colour passes, each in place on the current buffer (no flip)
                double* cur = static_cast<double*>(lv.sol[lv.cur[0]][0]);
                for (int c = c_begin; c < c_end; ++c) {
                    launch_vcstar2d<STAR_JACOBI>(ws, lv, r, cur, rhs, cur, sm.w00, c, nullptr);
                    fast_count(ws);
                }
            } else {
                launch_vcstar2d<STAR_JACOBI>(ws, lv, r, static_cast<const double*>(lv.sol[lv.cur[0]][0]), rhs,
                                             static_cast<double*>(lv.sol[lv.cur[0] ^ 1][0]), sm.w00, -1, nullptr);
                lv.cur[0] ^= 1;
                fast_count(ws);
            }
            return true;
        }
        if (sm.colored && !(c_begin == 0 && c_end == 2)) return false;        // single colour passes: generic kernel
        if (!star_level_ok<D>(ws, lv, 2) || !sm.point || sm.m != 1) return false;
        static const int rows_j = tunable("MG_JACOBI_ROWS", 0), rows_rb = tunable("MG_RBGS_ROWS", 0);
        const double* src = zero_in ? nullptr : static_cast<const double*>(lv.sol[lv.cur[0]][0]);
        double* dst = static_cast<double*>(lv.sol[lv.cur[0] ^ 1][0]);
        const double* rhs = static_cast<const double*>(lv.rhs[0]);
        if (sm.colored) {
            const int r = valid_rows(rows_rb, pick_rows(lv, 2));
            dim3 grid = strip_grid(lv, 2, r, lv.g.n - 2);
            static const int pf_env = tunable("MG_RBGS_PF", 20);
            const int pf = lv.g.n > 2049 ? pf_env : 0;            // smaller levels are L2 resident anyway
#define MG_RB(R_, Z_) launch_k(ws, k_rbgs2d<FAST_NT, R_, Z_>, grid, FAST_NT, 0, lv.g, star2d(lv), src, rhs, dst, sm.w00, pf)
            if (zero_in) {
                if (r == 8) MG_RB(8, true); else if (r == 4) MG_RB(4, true); else MG_RB(2, true);
            } else {
                if (r == 8) MG_RB(8, false); else if (r == 4) MG_RB(4, false); else MG_RB(2, false);
            }
#undef MG_RB
        } else {
            launch_star2d<STAR_JACOBI>(ws, lv, valid_rows(rows_j, pick_rows(lv, 2)), src, rhs, dst, sm.w00, nullptr);
        }
        lv.cur[0] ^= 1;
        fast_count(ws);
        return true;
    }
    return false;
}

template <typename T, int D, int F>
inline bool fast_residual(mg_workspace* ws, Level& lv) {
    if constexpr (D == 3 && F == 1 && sizeof(T) == sizeof(double)) {
        const int nt = star3d_nt(ws, lv);
        if (!nt || lv.g.nz < 3) return false;
        launch_star3d<STAR3_RESIDUAL>(ws, lv, nt, static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                      static_cast<const double*>(lv.rhs[0]), static_cast<double*>(lv.res[0]), 0.0, -1, nullptr);
        fast_count(ws);
        return true;
    }
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        if (ws->vc_kind) {
            if (!vc_level_ok<D>(ws, lv)) return false;
            launch_vcstar2d<STAR_RESIDUAL>(ws, lv, imin_h(4, pick_rows(lv, 2)), static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                           static_cast<const double*>(lv.rhs[0]), static_cast<double*>(lv.res[0]), 0.0, -1, nullptr);
            fast_count(ws);
            return true;
        }
        if (!star_level_ok<D>(ws, lv, 2)) return false;
        static const int rows = tunable("MG_RESIDUAL_ROWS", 0);
        launch_star2d<STAR_RESIDUAL>(ws, lv, valid_rows(rows, pick_rows(lv, 2)),
                                     static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                     static_cast<const double*>(lv.rhs[0]), static_cast<double*>(lv.res[0]), 0.0, nullptr);
        fast_count(ws);
        return true;
    }
    return false;
}

template <typename T, int D, int F>
inline bool fast_norm(mg_workspace* ws, Level& lv) {
    if constexpr (D == 3 && F == 1 && sizeof(T) == sizeof(double)) {
        const int nt = star3d_nt(ws, lv);
        if (!nt || lv.g.nz < 3) return false;
        dim3 grid = launch_star3d<STAR3_NORM>(ws, lv, nt, static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                              static_cast<const double*>(lv.rhs[0]), nullptr, 0.0, -1,
                                              static_cast<double*>(lv.partial));
        lv.norm_count = grid.x * grid.y * grid.z;
        fast_count(ws);
        return true;
    }
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        if (ws->vc_kind) {
            if (!vc_level_ok<D>(ws, lv)) return false;
            const int r = imin_h(4, pick_rows(lv, 2, MAX_PARTIALS));
            dim3 grid = strip_grid(lv, 2, r, lv.g.n - 2);
            if ((int)(grid.x * grid.y) > MAX_PARTIALS) return false;
            launch_vcstar2d<STAR_NORM>(ws, lv, r, static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                       static_cast<const double*>(lv.rhs[0]), nullptr, 0.0, -1, static_cast<double*>(lv.partial));
            lv.norm_count = grid.x * grid.y;
            fast_count(ws);
            return true;
        }
        if (!star_level_ok<D>(ws, lv, 2)) return false;
        static const int rows = tunable("MG_NORM_ROWS", 0);
        const int r = valid_rows(rows, pick_rows(lv, 2, MAX_PARTIALS));
        dim3 grid = strip_grid(lv, 2, r, lv.g.n - 2);
        if ((int)(grid.x * grid.y) > MAX_PARTIALS) return false;
        launch_star2d<STAR_NORM>(ws, lv, r, static_cast<const double*>(lv.sol[lv.cur[0]][0]),
                                 static_cast<const double*>(lv.rhs[0]), nullptr, 0.0, static_cast<double*>(lv.partial));
        lv.norm_count = grid.x * grid.y;
        fast_count(ws);
        return true;
    }
    return false;
}

// fused 3D residual + restriction (mg_restrict3d.cuh): whole grids, fine level of at least 129^3 nodes
template <int RYC, int NS, int NF>
inline bool launch_resrestrict3d_cfg(mg_workspace* ws, Level& lf, Level& lc, const double* u, const double* f, double* out) {
    static const int zc_env = tunable("MG_RESTRICT3D_ZCHUNK", 0);
    static bool opted_in[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !opted_in[dev]) {
        if (cudaFuncSetAttribute(k_resrestrict3d<RYC, NS, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)resrestrict3d_smem_bytes<RYC, NS, NF>()) != cudaSuccess)
            return false;
        opted_in[dev] = true;
    }
    const SlabView v = slab_view(lf);
    const CUtensorMap* tm_u = plane_tensor_map(u - v.shift, lf.g, 2 * RYC + 3, v.planes);
    const CUtensorMap* tm_f = plane_tensor_map(f - v.shift, lf.g, 2 * RYC + 1, v.planes);
    if (!tm_u || !tm_f) return false;
    const int m = lc.g.n - 2;                                              // interior coarse nodes per dimension
    // owned coarse planes (global indices); `out` is addressed from the coarse array's local plane 0
    const int Zlo = imax_h(1, lc.g.z0 + 1), Zhi = imin_h(lc.g.n - 2, lc.g.z0 + lc.g.nz - 2), mz = Zhi - Zlo + 1;
    if (mz < 1) return true;                                               // nothing owned on the coarse level
    const int gx = (m + RS3_TXC - 1) / RS3_TXC, gy = (m + RYC - 1) / RYC;
    int zchunk = zc_env >= 2 ? zc_env : 32;
    while (zc_env < 2 && zchunk > 2 && gx * gy * ((mz + zchunk - 1) / zchunk) < 888) zchunk >>= 1;
    launch_k(ws, k_resrestrict3d<RYC, NS, NF>, dim3(gx, gy, (mz + zchunk - 1) / zchunk), 32 * (2 * RYC + 1),
             resrestrict3d_smem_bytes<RYC, NS, NF>(), lc.g, lf.g, star3d(lf), *tm_u, *tm_f, out, zchunk,
             v.zoff, lc.g.z0, Zlo, Zhi);
    return true;
}

// the same stage with two fine rows per warp (k_resrestrict3d2)
template <int RYC, int NS, int NF>
inline bool launch_resrestrict3d2_cfg(mg_workspace* ws, Level& lf, Level& lc, const double* u, const double* f, double* out) {
    static const int zc_env = tunable("MG_RESTRICT3D_ZCHUNK", 0);
    static bool opted_in[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !opted_in[dev]) {
        if (cudaFuncSetAttribute(k_resrestrict3d2<RYC, NS, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)resrestrict3d_smem_bytes<RYC, NS, NF>()) != cudaSuccess)
            return false;
        opted_in[dev] = true;
    }
    const SlabView v = slab_view(lf);
    const CUtensorMap* tm_u = plane_tensor_map(u - v.shift, lf.g, 2 * RYC + 3, v.planes);
    const CUtensorMap* tm_f = plane_tensor_map(f - v.shift, lf.g, 2 * RYC + 1, v.planes);
    if (!tm_u || !tm_f) return false;
    const int m = lc.g.n - 2;
    const int Zlo = imax_h(1, lc.g.z0 + 1), Zhi = imin_h(lc.g.n - 2, lc.g.z0 + lc.g.nz - 2), mz = Zhi - Zlo + 1;
    if (mz < 1) return true;
    const int gx = (m + RS3_TXC - 1) / RS3_TXC, gy = (m + RYC - 1) / RYC;
    int zchunk = zc_env >= 2 ? zc_env : 32;
    while (zc_env < 2 && zchunk > 2 && gx * gy * ((mz + zchunk - 1) / zchunk) < 888) zchunk >>= 1;
    launch_k(ws, k_resrestrict3d2<RYC, NS, NF>, dim3(gx, gy, (mz + zchunk - 1) / zchunk), 32 * (RYC + 1),
             resrestrict3d_smem_bytes<RYC, NS, NF>(), lc.g, lf.g, star3d(lf), *tm_u, *tm_f, out, zchunk,
             v.zoff, lc.g.z0, Zlo, Zhi);
    return true;
}

// would residual + restriction from this fine level run as the fused TMA kernel?
inline bool resrestrict3d_applicable(const mg_workspace* ws, const Level& lf) {
    static const int enabled = tunable("MG_RESTRICT3D", 1);
    const int min_n = tunable("MG_RESTRICT3D_MIN_N", 129);     // not cached: the tests lower it to 65
    return enabled && ws->use_fast && star3d_nt(ws, lf) && one_pass_geometry_ok(lf) && lf.g.n >= 65 && lf.g.n >= min_n;
}

template <typename T, int D, int F>
inline bool fast_residual_restrict(mg_workspace* ws, Level& lf, Level& lc) {
    if constexpr (D == 3 && F == 1 && sizeof(T) == sizeof(double)) {
        static const int cfg = tunable("MG_RESTRICT3D_CFG", 0);
        if (!resrestrict3d_applicable(ws, lf)) return false;
        const double* u = static_cast<const double*>(lf.sol[lf.cur[0]][0]);
        const double* f = static_cast<const double*>(lf.rhs[0]);
        double* out = static_cast<double*>(lc.rhs[0]);
        bool done;
        static const int version = tunable("MG_RESTRICT3D_V", 1);   // 2: two rows per warp (k_resrestrict3d2): measured no gain (0.4225 against 0.418 ms at 513^3)
        if (version == 2) {
            switch (cfg) {
                case 1: done = launch_resrestrict3d2_cfg<4, 4, 3>(ws, lf, lc, u, f, out); break;
                case 2: done = launch_resrestrict3d2_cfg<8, 5, 3>(ws, lf, lc, u, f, out); break;
                case 3: done = launch_resrestrict3d2_cfg<6, 5, 4>(ws, lf, lc, u, f, out); break;
                case 4: done = launch_resrestrict3d2_cfg<4, 6, 5>(ws, lf, lc, u, f, out); break;
                default: done = launch_resrestrict3d2_cfg<4, 5, 4>(ws, lf, lc, u, f, out); break;
            }
            if (done) fast_count(ws);
            return done;
        }
        switch (cfg) {
            case 1: done = launch_resrestrict3d_cfg<4, 4, 3>(ws, lf, lc, u, f, out); break;
            case 2: done = launch_resrestrict3d_cfg<8, 5, 3>(ws, lf, lc, u, f, out); break;
            case 3: done = launch_resrestrict3d_cfg<2, 6, 4>(ws, lf, lc, u, f, out); break;
            default: done = launch_resrestrict3d_cfg<4, 5, 4>(ws, lf, lc, u, f, out); break;
        }
        if (done) fast_count(ws);
        return done;
    }
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        static const int tile_rc = tunable("MG_RESTRICT_RC", 3);     // 0: marching kernel
        // the tile kernel clamps its columns itself: any level the 2-column kernels accept (>= 257^2)
        if (!star_level_ok<D>(ws, lf, tile_rc > 0 ? 2 : 4) || (tile_rc == 0 && (lc.g.n - 1) % (2 * FAST_NT) != 0)) return false;
        if (tile_rc > 0) {
            const double* u = static_cast<const double*>(lf.sol[lf.cur[0]][0]);
            const double* f = static_cast<const double*>(lf.rhs[0]);
            double* out = static_cast<double*>(lc.rhs[0]);
            const int warps = (lc.g.n - 2 + 30) / 31, nrows = lc.g.n - 2;
            const int gx = (warps + FAST_NT / 32 - 1) / (FAST_NT / 32);
            static const int pf_env = tunable("MG_RESTRICT_PF", 20);
            const int pf = lf.g.n > 2049 ? pf_env : 0;            // smaller levels are L2 resident anyway
            if (tile_rc == 1)
                launch_k(ws, k_resrestrict2d_t<FAST_NT, 1>, dim3(gx, nrows, 1), FAST_NT, 0, lc.g, lf.g, star2d(lf), u, f, out, pf);
            else if (tile_rc == 2)
                launch_k(ws, k_resrestrict2d_t<FAST_NT, 2>, dim3(gx, (nrows + 1) / 2, 1), FAST_NT, 0, lc.g, lf.g, star2d(lf), u, f, out, pf);
            else if (tile_rc == 3)
                launch_k(ws, k_resrestrict2d_t<FAST_NT, 3>, dim3(gx, (nrows + 2) / 3, 1), FAST_NT, 0, lc.g, lf.g, star2d(lf), u, f, out, pf);
            else
                launch_k(ws, k_resrestrict2d_t<FAST_NT, 4>, dim3(gx, (nrows + 3) / 4, 1), FAST_NT, 0, lc.g, lf.g, star2d(lf), u, f, out, pf);
            fast_count(ws);
            return true;
        }
        static const int rows_env = tunable("MG_RESTRICT_ROWS", 0);
        const int rows = rows_env > 0 ? rows_env : march_rows(lc.g.n - 2, (lc.g.n - 1) / (2 * FAST_NT));
        launch_k(ws, k_resrestrict2d<FAST_NT>, strip_grid(lc, 2, rows, lc.g.n - 2), FAST_NT, 0, 
            lc.g, lf.g, star2d(lf), static_cast<const double*>(lf.sol[lf.cur[0]][0]),
            static_cast<const double*>(lf.rhs[0]), static_cast<double*>(lc.rhs[0]), rows);
        fast_count(ws);
        return true;
    }
    return false;
}

template <typename T, int D, int F>
inline bool fast_prolong_add(mg_workspace* ws, Level& lf, Level& lc, const FieldSet<T>& src, double omega) {
    if constexpr (D == 3 && F == 1 && sizeof(T) == sizeof(double)) {
        const int nt = star3d_nt(ws, lf);
        if (!nt || lf.g.nz < 3) return false;
        // rows per thread: 8 on large levels (513^3: 0.486 ms against 0.510 with 4 and 0.604 with 2), else 4
        static const int ry_tuned = tunable("MG_PROLONG3D_RY", 0);
        const int ry_env = ry_tuned ? ry_tuned : (lf.g.n >= 257 ? 8 : 4);
        const double* e = reinterpret_cast<const double*>(src.p[0]);
        double* u = static_cast<double*>(lf.sol[lf.cur[0]][0]);
        static const int version = tunable("MG_PROLONG3D_V", 2);
        if (version == 2) {
            // pairs of fine planes (2Z+1, 2Z+2) that contain an owned interior plane of this level
            const int zg_first = imax_h(1, lf.g.z0 + 1), zg_last = imin_h(lf.g.n - 2, lf.g.z0 + lf.g.nz - 2);
            if (zg_last < zg_first) return true;
            const int p_lo = (zg_first - 1) >> 1, p_hi = (zg_last - 1) >> 1;
            auto launch2 = [&](auto ry_tag) {
                constexpr int RY = decltype(ry_tag)::value;
                dim3 grid((lf.g.n - 1) / (2 * nt), (lf.g.n - 2 + RY - 1) / RY, p_hi - p_lo + 1);
                if (nt == 128) launch_k(ws, k_prolong3d_zz<128, RY>, grid, 128, 0, lf.g, lc.g, e, u, omega, p_lo);
                else if (nt == 64) launch_k(ws, k_prolong3d_zz<64, RY>, grid, 64, 0, lf.g, lc.g, e, u, omega, p_lo);
                else launch_k(ws, k_prolong3d_zz<32, RY>, grid, 32, 0, lf.g, lc.g, e, u, omega, p_lo);
            };
            static const int ry2 = tunable("MG_PROLONG3D_RY", 0);
            if (ry2 == 8) launch2(std::integral_constant<int, 8>{});
            else if (ry2 == 2) launch2(std::integral_constant<int, 2>{});
            else launch2(std::integral_constant<int, 4>{});
            fast_count(ws);
            return true;
        }
        auto launch = [&](auto ry_tag) {
            constexpr int RY = decltype(ry_tag)::value;
            dim3 grid((lf.g.n - 1) / (2 * nt), (lf.g.n - 2 + RY - 1) / RY, lf.g.nz - 2);
            if (nt == 128) launch_k(ws, k_prolong3d<128, RY>, grid, 128, 0, lf.g, lc.g, e, u, omega);
            else if (nt == 64) launch_k(ws, k_prolong3d<64, RY>, grid, 64, 0, lf.g, lc.g, e, u, omega);
            else launch_k(ws, k_prolong3d<32, RY>, grid, 32, 0, lf.g, lc.g, e, u, omega);
        };
        if (ry_env == 8) launch(std::integral_constant<int, 8>{});
        else if (ry_env == 2) launch(std::integral_constant<int, 2>{});
        else launch(std::integral_constant<int, 4>{});
        fast_count(ws);
        return true;
    }
    if constexpr (D == 2 && F == 1 && sizeof(T) == sizeof(double)) {
        static const int tile_rc = tunable("MG_PROLONG_RC", 4);      // 0: marching kernel
        if (ws->dtype != MG_F64 || (lf.g.n - 1) % ((tile_rc > 0 ? 2 : 4) * FAST_NT) != 0) return false;
        if (tile_rc > 0) {
            const double* e = reinterpret_cast<const double*>(src.p[0]);
            double* u = static_cast<double*>(lf.sol[lf.cur[0]][0]);
            const int gx = (lf.g.n - 1) / (2 * FAST_NT), nrows = lc.g.n - 1;
            if (tile_rc == 1)
                launch_k(ws, k_prolong2d_t<FAST_NT, 1>, dim3(gx, nrows, 1), FAST_NT, 0, lf.g, lc.g, e, u, omega);
            else if (tile_rc == 2)
                launch_k(ws, k_prolong2d_t<FAST_NT, 2>, dim3(gx, (nrows + 1) / 2, 1), FAST_NT, 0, lf.g, lc.g, e, u, omega);
            else
                launch_k(ws, k_prolong2d_t<FAST_NT, 4>, dim3(gx, (nrows + 3) / 4, 1), FAST_NT, 0, lf.g, lc.g, e, u, omega);
            fast_count(ws);
            return true;
        }
        static const int rows_env = tunable("MG_PROLONG_ROWS", 0);
        const int rows = rows_env > 0 ? rows_env : march_rows(lc.g.n - 1, (lf.g.n - 1) / (4 * FAST_NT));
        launch_k(ws, k_prolong2d<FAST_NT>, strip_grid(lf, 4, rows, lc.g.n - 1), FAST_NT, 0, 
            lf.g, lc.g, reinterpret_cast<const double*>(src.p[0]), static_cast<double*>(lf.sol[lf.cur[0]][0]), omega,
            rows);
        fast_count(ws);
        return true;
    }
    return false;
}

}  // namespace mg
