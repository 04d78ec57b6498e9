// mg_engine.cu -- C ABI of the B200 multigrid cycle engine (see include/mgengine.h).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -lineinfo -shared ...
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "mg_exec.cuh"

namespace mg {
thread_local std::string g_last_error;
}
using namespace mg;

// dtype x dimension x fields dispatch; only the combinations the configurations need are instantiated
#define MG_DISPATCH(ws, CALL)                                                                       \
    [&]() -> int {                                                                                  \
        if ((ws)->dtype == MG_F64) {                                                                \
            if ((ws)->dim == 2 && (ws)->n_fields == 1) { using T = double; constexpr int D = 2, F = 1; return CALL; } \
            if ((ws)->dim == 2 && (ws)->n_fields == 2) { using T = double; constexpr int D = 2, F = 2; return CALL; } \
            if ((ws)->dim == 3 && (ws)->n_fields == 1) { using T = double; constexpr int D = 3, F = 1; return CALL; } \
        } else {                                                                                    \
            if ((ws)->dim == 2 && (ws)->n_fields == 1) { using T = cplx; constexpr int D = 2, F = 1; return CALL; }   \
        }                                                                                           \
        return fail(MG_ERR_INVALID, "unsupported dtype/dimension/field combination");              \
    }()

static int check_ws(mg_workspace* ws) {
    if (!ws) return fail(MG_ERR_INVALID, "null workspace");
    cudaError_t e = cudaSetDevice(ws->device);
    if (e != cudaSuccess) return fail(MG_ERR_CUDA, cudaGetErrorString(e));
    return MG_OK;
}

extern "C" const char* mg_last_error(void) { return g_last_error.c_str(); }

static int create_impl(mg_workspace** out, int dimension, int min_level, int max_level, int n_fields, int dtype,
                       int device, const int32_t* z_begin, const int32_t* z_end);

extern "C" int mg_create(mg_workspace** out, int dimension, int min_level, int max_level, int n_fields, int dtype,
                         int device) {
    return create_impl(out, dimension, min_level, max_level, n_fields, dtype, device, nullptr, nullptr);
}

extern "C" int mg_create_slab(mg_workspace** out, int min_level, int max_level, int n_fields, int dtype, int device,
                              const int32_t* z_begin, const int32_t* z_end) {
    if (!z_begin || !z_end) return fail(MG_ERR_INVALID, "slab plane ranges are null");
    for (int level = min_level; level <= max_level; ++level) {
        const int n = (1 << level) + 1, zb = z_begin[level - min_level], ze = z_end[level - min_level];
        if (zb < 1 || ze > n - 2 || ze < zb - 1) return fail(MG_ERR_INVALID, "slab plane range outside the interior");
    }
    return create_impl(out, 3, min_level, max_level, n_fields, dtype, device, z_begin, z_end);
}

static int create_impl(mg_workspace** out, int dimension, int min_level, int max_level, int n_fields, int dtype,
                       int device, const int32_t* z_begin, const int32_t* z_end) {
    if (!out) return fail(MG_ERR_INVALID, "out is null");
    if (dimension < 2 || dimension > 3) return fail(MG_ERR_INVALID, "dimension must be 2 or 3");
    if (min_level < 1 || max_level < min_level || max_level > 14) return fail(MG_ERR_INVALID, "bad level range");
    if (n_fields < 1 || n_fields > MG_MAX_FIELDS) return fail(MG_ERR_INVALID, "n_fields must be 1 or 2");
    if (dtype != MG_F64 && dtype != MG_C128) return fail(MG_ERR_INVALID, "bad dtype");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(MG_ERR_CUDA, "no CUDA device: the engine has no CPU fallback");
    if (device < 0 || device >= count) return fail(MG_ERR_INVALID, "bad device ordinal");
    MG_CUDA(cudaSetDevice(device));
    mg_workspace* ws = new mg_workspace();
    ws->dim = dimension; ws->min_level = min_level; ws->max_level = max_level;
    ws->n_fields = n_fields; ws->dtype = dtype; ws->device = device;
    ws->elem = dtype == MG_F64 ? 8 : 16;
    // small levels run in the tail kernel: a cluster of 16 CTAs up to 129^2 nodes (2D; measured best:
    // level 7, CTA 0 alone up to level 5), a single CTA up to 17^3 nodes (3D)
    // 3D (round 2): one cluster of 16 CTAs up to 33^3 (level 5), CTA 0 alone up to 9^3: a 129^3 V(2,1) cycle -- the
    // replicated part of a decomposed 513^3 solve on 8 GPUs -- 0.270 -> 0.236 ms; with 65^3 in the cluster it is slower
    // (0.308 ms: 8192 threads against full-grid kernels)
    ws->tail_max_level = dimension == 3 ? 5 : 7;
    ws->tail_ctas = 16;
    if (const char* v = std::getenv("MG_TAIL_CTAS")) ws->tail_ctas = std::atoi(v);
    if (ws->tail_ctas < 1 || ws->tail_ctas > 16) ws->tail_ctas = 1;
    ws->tail_solo_level = dimension == 3 ? 3 : 5;
    if (const char* v = std::getenv("MG_TAIL_SOLO")) ws->tail_solo_level = std::atoi(v);
    if (const char* v = std::getenv("MG_TAIL_LEVEL")) ws->tail_max_level = std::atoi(v);
    if (const char* v = std::getenv("MG_TAIL_LEAN")) ws->tail_lean = std::atoi(v) != 0;
    if (const char* v = std::getenv("MG_PDL")) ws->use_pdl = std::atoi(v) != 0;
    if (const char* v = std::getenv("MG_LAZY_ZERO")) ws->lazy_zero = std::atoi(v) != 0;
    MG_CUDA(cudaStreamCreateWithFlags(&ws->stream, cudaStreamNonBlocking));
    MG_CUDA(cudaEventCreate(&ws->ev0));
    MG_CUDA(cudaEventCreate(&ws->ev1));
    MG_CUDA(cudaEventCreate(&ws->mark[0]));
    MG_CUDA(cudaEventCreate(&ws->mark[1]));
    MG_CUDA(cudaMalloc(&ws->d_norm, sizeof(double)));
    MG_CUDA(cudaMallocHost(&ws->h_norm, sizeof(double)));
    ws->levels.resize(max_level - min_level + 1);
    for (int level = min_level; level <= max_level; ++level) {
        Level& lv = ws->L(level);
        lv.level = level;
        Geom& g = lv.g;
        g.dim = dimension;
        g.n = (1 << level) + 1;
        g.pitch = (g.n + 15) / 16 * 16;
        g.plane = (long long)g.pitch * g.n;
        g.nz = dimension == 3 ? g.n : 1;
        g.z0 = 0;
        if (z_begin) {                         // owned planes [zb, ze] + one ghost / boundary plane on either side
            g.z0 = z_begin[level - min_level] - 1;
            g.nz = z_end[level - min_level] - z_begin[level - min_level] + 3;
            ws->slab = true;
        }
        g.total = dimension == 3 ? g.plane * g.nz : g.plane;
        const bool finest = level == max_level;
        const int arrays_per_field = 4 + (finest ? 2 : 0);
        // one guard row before and after every array so that halo reads of the specialised
        // kernels stay inside the slab
        // z-slabs: one whole plane more on either side -- the second ghost plane of the one-pass 3D kernels
        // (local planes -1 and nz; filled by the halo exchange, mg_dd.cuh)
        lv.wide_ghost = z_begin != nullptr;
        const long long guard = 2 * g.pitch + (lv.wide_ghost ? g.plane : 0);
        const long long stride = g.total + 2 * guard;
        size_t bytes = (size_t)stride * arrays_per_field * n_fields * ws->elem;
        e = cudaMalloc(&lv.slab, bytes);
        if (e != cudaSuccess) {
            mg_destroy(ws);
            return fail(MG_ERR_NOMEM, "cudaMalloc of level " + std::to_string(level) + " failed: " + cudaGetErrorString(e));
        }
        lv.slab_bytes = bytes;
        MG_CUDA(cudaMemsetAsync(lv.slab, 0, bytes, ws->stream));
        char* base = static_cast<char*>(lv.slab);
        long long k = 0;
        auto next = [&]() { void* p = base + (size_t)(k * stride + guard) * ws->elem; ++k; return p; };
        for (int f = 0; f < n_fields; ++f) {
            lv.sol[0][f] = next(); lv.sol[1][f] = next(); lv.rhs[f] = next(); lv.res[f] = next();
            if (finest) { lv.cor[f] = next(); lv.init[f] = next(); }
        }
        const int rows = n_fields * (dimension == 3 ? g.n * g.nz : g.n);
        lv.norm_blocks = rows < 1184 ? rows : 1184;          // 8 CTAs of 256 threads per SM x 148 SMs
        MG_CUDA(cudaMalloc(&lv.partial, (size_t)MAX_PARTIALS * ws->elem));
    }
    MG_CUDA(cudaStreamSynchronize(ws->stream));
    *out = ws;
    return MG_OK;
}

static void dd_release(mg_workspace* ws);

static void free_smoothers(mg_workspace* ws) {
    for (auto& s : ws->smoothers) {
        if (s.d_table) cudaFree(s.d_table);
        if (s.d_w) cudaFree(s.d_w);
    }
    ws->smoothers.clear();
}

static void free_captured(mg_workspace* ws) {
    for (auto& c : ws->captured)
        if (c.exec) cudaGraphExecDestroy(c.exec);
    ws->captured.clear();
}

static void free_cycles(mg_workspace* ws) {
    free_captured(ws);
    if (ws->outer_graph) {
        cudaGraphExecDestroy(ws->outer_graph);
        ws->outer_graph = nullptr;
        ws->outer_graph_cycle = -1;
    }
    for (auto& c : ws->cycles)
        for (auto& kv : c.variants)
            if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    ws->cycles.clear();
}

// drop the instantiated graphs (they are bound to a stream's capture) but keep the op lists
static void free_cycles_graphs(mg_workspace* ws) {
    free_captured(ws);           // they may reference tail programs of the pool that is reset below
    for (auto& c : ws->cycles) {
        for (auto& kv : c.variants)
            if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        c.variants.clear();
    }
    if (ws->outer_graph) {
        cudaGraphExecDestroy(ws->outer_graph);
        ws->outer_graph = nullptr;
        ws->outer_graph_cycle = -1;
    }
    ws->tail_pool_used = 0;       // the tail programs of the destroyed graphs
}

extern "C" int mg_destroy(mg_workspace* ws) {
    if (!ws) return MG_OK;
    cudaSetDevice(ws->device);
    if (ws->own_stream && ws->stream != ws->own_stream) {
        // running on a caller's stream (mg_set_stream, mg_dd_init), which may be gone by now: wait for the device
        cudaDeviceSynchronize();
        ws->stream = ws->own_stream;
    }
    if (ws->stream) cudaStreamSynchronize(ws->stream);
    free_cycles(ws);
    free_smoothers(ws);
    for (auto& lv : ws->levels) {
        if (lv.slab) cudaFree(lv.slab);
        if (lv.d_op) cudaFree(lv.d_op);
        if (lv.partial) cudaFree(lv.partial);
    }
    dd_release(ws);
    if (ws->cgs_work) cudaFree(ws->cgs_work);
    if (ws->krylov_work) cudaFree(ws->krylov_work);
    for (void* p : ws->fas_raw) cudaFree(p);
    delete ws->tail;
    if (ws->tail_pool) cudaFree(ws->tail_pool);
    if (ws->tail_stage) cudaFreeHost(ws->tail_stage);
    if (ws->aux_stream) cudaStreamDestroy(ws->aux_stream);
    if (ws->outer_graph) cudaGraphExecDestroy(ws->outer_graph);
    if (ws->outer_slab) cudaFree(ws->outer_slab);
    if (ws->d_outer_scalars) cudaFree(ws->d_outer_scalars);
    if (ws->d_op_outer) cudaFree(ws->d_op_outer);
    if (ws->d_norm) cudaFree(ws->d_norm);
    if (ws->h_norm) cudaFreeHost(ws->h_norm);
    if (ws->ev0) cudaEventDestroy(ws->ev0);
    if (ws->ev1) cudaEventDestroy(ws->ev1);
    if (ws->mark[0]) cudaEventDestroy(ws->mark[0]);
    if (ws->mark[1]) cudaEventDestroy(ws->mark[1]);
    if (ws->flush_buf) cudaFree(ws->flush_buf);
    if (ws->own_stream) ws->stream = ws->own_stream;          // a borrowed stream is not ours to destroy
    if (ws->stream) cudaStreamDestroy(ws->stream);
    delete ws;
    return MG_OK;
}

extern "C" int mg_launch_count(mg_workspace* ws, int64_t* out) {
    if (!ws || !out) return fail(MG_ERR_INVALID, "null argument");
    *out = ws->launches;
    return MG_OK;
}

extern "C" int mg_set_operator(mg_workspace* ws, int level, int row, int col, int n_entries, const int32_t* offsets,
                               const double* values) {
    int rc = check_ws(ws);
    if (rc) return rc;
    if (!ws->has_level(level)) return fail(MG_ERR_INVALID, "level out of range");
    if (row < 0 || row >= ws->n_fields || col < 0 || col >= ws->n_fields) return fail(MG_ERR_INVALID, "bad block index");
    if (n_entries < 0 || n_entries > MG_MAX_ENTRIES) return fail(MG_ERR_INVALID, "too many stencil entries");
    HostOperator& h = ws->L(level).hop;
    const int vs = ws->dtype == MG_C128 ? 2 : 1;
    h.n_entries[row][col] = n_entries;
    for (int k = 0; k < n_entries; ++k) {
        for (int c = 0; c < MG_MAX_DIM; ++c) h.offs[row][col][k][c] = c < ws->dim ? offsets[k * ws->dim + c] : 0;
        for (int c = 0; c < ws->dim; ++c)
            if (offsets[k * ws->dim + c] < -1 || offsets[k * ws->dim + c] > 1)
                return fail(MG_ERR_INVALID, "stencil offsets must lie in [-1, 1]");
        h.val[row][col][k][0] = values[k * vs];
        h.val[row][col][k][1] = vs == 2 ? values[k * vs + 1] : 0.0;
    }
    h.set[row][col] = true;
    ws->L(level).op_dirty = true;
    return MG_OK;
}

extern "C" int mg_set_variable_coefficient(mg_workspace* ws, int kind, double kappa) {
    int rc = check_ws(ws);
    if (rc) return rc;
    if (kind != MG_VC_NONE && kind != MG_VC_EXP_KAPPA) return fail(MG_ERR_INVALID, "bad variable-coefficient kind");
    if (kind != MG_VC_NONE && !(ws->dtype == MG_F64 && ws->n_fields == 1))
        return fail(MG_ERR_INVALID, "variable coefficients exist for scalar real problems only");
    if (kind != MG_VC_NONE && ws->slab) return fail(MG_ERR_INVALID, "variable coefficients are not available on a z-slab workspace");
    if (kind != MG_VC_NONE)
        for (auto& sm : ws->smoothers)
            if (!sm.point) return fail(MG_ERR_STATE, "tile smoothers registered: clear the smoothers first");
    ws->vc_kind = kind;
    ws->vc_kappa = kappa;
    static const int star2[5][3] = {{0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}};
    static const int star3[7][3] = {{0, 0, -1}, {0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (auto& lv : ws->levels) {
        lv.op_dirty = true;
        if (kind == MG_VC_NONE || lv.hop.set[0][0]) continue;
        // structure of the operator (values are evaluated per node on the device): the (2d+1)-point star
        const int ne = 2 * ws->dim + 1;
        lv.hop.n_entries[0][0] = ne;
        for (int k = 0; k < ne; ++k)
            for (int c = 0; c < MG_MAX_DIM; ++c) lv.hop.offs[0][0][k][c] = ws->dim == 2 ? star2[k][c] : star3[k][c];
        lv.hop.set[0][0] = true;
    }
    return MG_OK;
}

extern "C" int mg_set_boundary(mg_workspace* ws, int bc_kind, double k) {
    if (!ws) return fail(MG_ERR_INVALID, "null workspace");
    if (bc_kind == MG_BC_HELMHOLTZ_ROBIN && !(ws->dtype == MG_C128 && ws->dim == 2))
        return fail(MG_ERR_INVALID, "Robin boundary needs a 2D complex workspace");
    if (bc_kind != MG_BC_DIRICHLET && bc_kind != MG_BC_HELMHOLTZ_ROBIN) return fail(MG_ERR_INVALID, "bad bc kind");
    ws->bc_kind = bc_kind;
    ws->bc_k = k;
    return MG_OK;
}

extern "C" int mg_set_coarse_solver(mg_workspace* ws, int kind, int max_iterations, double reduction) {
    if (!ws) return fail(MG_ERR_INVALID, "null workspace");
    if (kind != MG_CGS_CG && kind != MG_CGS_BICGSTAB) return fail(MG_ERR_INVALID, "bad coarse solver kind");
    ws->cgs_kind = kind;
    ws->cgs_max_its = max_iterations;
    ws->cgs_reduction = reduction;
    return MG_OK;
}

template <typename T>
static int add_smoother_t(mg_workspace* ws, Smoother& sm, const int32_t* slot_pos, int n_masks, const uint32_t* masks,
                          const double* w_table) {
    SmootherTable<T> tab;
    memset(&tab, 0, sizeof(tab));
    tab.n_slots = sm.m;
    for (int d = 0; d < MG_MAX_DIM; ++d) tab.period[d] = sm.period[d];
    for (int s = 0; s < sm.m; ++s) {
        tab.slot_field[s] = sm.slot_field[s];
        for (int d = 0; d < MG_MAX_DIM; ++d) tab.slot_pos[s][d] = d < ws->dim ? slot_pos[s * ws->dim + d] : 0;
    }
    for (int i = 0; i < (1 << MG_MAX_SLOTS); ++i) tab.mask_index[i] = -1;
    for (int k = 0; k < n_masks; ++k) {
        if (masks[k] >= (1u << sm.m)) return fail(MG_ERR_INVALID, "smoother mask out of range");
        tab.mask_index[masks[k]] = k;
    }
    if (tab.mask_index[(1 << sm.m) - 1] < 0) return fail(MG_ERR_INVALID, "W table for the all-valid mask is missing");
    MG_CUDA(cudaMalloc(&sm.d_table, sizeof(tab)));
    MG_CUDA(cudaMemcpy(sm.d_table, &tab, sizeof(tab), cudaMemcpyHostToDevice));
    size_t wbytes = (size_t)n_masks * sm.m * sm.m * sizeof(T);
    MG_CUDA(cudaMalloc(&sm.d_w, wbytes));
    MG_CUDA(cudaMemcpy(sm.d_w, w_table, wbytes, cudaMemcpyHostToDevice));   // same memory layout as T
    sm.full_index = tab.mask_index[(1 << sm.m) - 1];
    sm.canonical_slots = ws->dim == 2;
    for (int s = 0; s < sm.m && sm.canonical_slots; ++s)
        sm.canonical_slots = sm.slot_field[s] == 0 && slot_pos[s * ws->dim] == s / sm.period[1] &&
                             slot_pos[s * ws->dim + 1] == s % sm.period[1];
    sm.w00 = w_table[(size_t)tab.mask_index[(1 << sm.m) - 1] * sm.m * sm.m * (sizeof(T) / sizeof(double))];
    return MG_OK;
}

extern "C" int mg_add_smoother(mg_workspace* ws, int level, int jacobi, int colored, const int32_t* period, int n_slots,
                               const int32_t* slot_field, const int32_t* slot_pos, int n_masks, const uint32_t* masks,
                               const double* w_table, int* handle) {
    int rc = check_ws(ws);
    if (rc) return rc;
    if (!ws->has_level(level)) return fail(MG_ERR_INVALID, "level out of range");
    if (n_slots < 1 || n_slots > MG_MAX_SLOTS) return fail(MG_ERR_INVALID, "local system size out of range");
    if (n_masks < 1 || !masks || !w_table || !handle) return fail(MG_ERR_INVALID, "null/empty smoother table");
    Smoother sm;
    sm.level = level;
    sm.jacobi = jacobi != 0;
    sm.colored = colored != 0;
    sm.m = n_slots;
    sm.point = true;
    for (int d = 0; d < MG_MAX_DIM; ++d) {
        sm.period[d] = d < ws->dim ? period[d] : 1;
        if (sm.period[d] < 1 || sm.period[d] > MG_MAX_SLOTS) return fail(MG_ERR_INVALID, "bad tile shape");
        if (sm.period[d] != 1) sm.point = false;
    }
    for (int s = 0; s < n_slots; ++s) {
        if (slot_field[s] < 0 || slot_field[s] >= ws->n_fields) return fail(MG_ERR_INVALID, "bad slot field");
        sm.slot_field[s] = slot_field[s];
        for (int d = 0; d < ws->dim; ++d)
            if (slot_pos[s * ws->dim + d] < 0 || slot_pos[s * ws->dim + d] >= sm.period[d])
                return fail(MG_ERR_INVALID, "slot position outside the tile");
    }
    if (sm.point && n_slots > ws->n_fields) return fail(MG_ERR_INVALID, "point smoother with more slots than fields");
    if (!sm.point && (sm.colored || !sm.jacobi)) return fail(MG_ERR_INVALID, "tile smoothers are Jacobi only");
    if (!sm.point && ws->slab) return fail(MG_ERR_INVALID, "tile smoothers are not available on a z-slab workspace");
    if (!sm.point && ws->vc_kind) return fail(MG_ERR_INVALID, "tile smoothers are not available for variable-coefficient operators");
    if (ws->vc_kind && !(n_masks == 1 && n_slots == 1)) return fail(MG_ERR_INVALID, "variable-coefficient point smoother: one slot, one mask");
    rc = ws->dtype == MG_F64 ? add_smoother_t<double>(ws, sm, slot_pos, n_masks, masks, w_table)
                             : add_smoother_t<cplx>(ws, sm, slot_pos, n_masks, masks, w_table);
    if (rc) {
        if (sm.d_table) cudaFree(sm.d_table);
        if (sm.d_w) cudaFree(sm.d_w);
        return rc;
    }
    ws->smoothers.push_back(sm);
    *handle = (int)ws->smoothers.size() - 1;
    return MG_OK;
}

extern "C" int mg_clear_smoothers(mg_workspace* ws) {
    int rc = check_ws(ws);
    if (rc) return rc;
    MG_CUDA(cudaStreamSynchronize(ws->stream));
    free_cycles(ws);          // cycles reference smoother handles
    free_smoothers(ws);
    return MG_OK;
}

#include "mg_engine_run.inc"
#include "mg_engine_fas.inc"
#include "mg_engine_noalloc.inc"
#include "mg_engine_dd.inc"
