// mg_exec.cuh -- op dispatch: one lowered op -> kernel launches on the workspace stream.
// Buffer (ping-pong) state is tracked per level and field at issue time, so the same code path
// serves direct execution (mg_exec_ops) and CUDA-graph capture (mg_build_cycle).
#pragma once
#include "mg_coarse.cuh"
#include "mg_tail.cuh"
#include "mg_kernels_fast.cuh"
#include "mg_outer.cuh"
#include "mg_dd.cuh"
#include "mg_workspace.cuh"

namespace mg {

template <typename T>
int materialise_zero(mg_workspace* ws, Level& lv);      // deferred zero fill, defined with exec_zero below

constexpr int TAIL_RING_SLOTS = 32;                          // device slots for directly launched tail programs
constexpr size_t TAIL_POOL_BYTES = (size_t)16 << 20;         // device slots for the programs of captured cycles

template <typename T>
inline FieldSet<T> fieldset(void* const* arr) {
    FieldSet<T> fs;
    for (int i = 0; i < MG_MAX_FIELDS; ++i) fs.p[i] = static_cast<T*>(arr[i] ? arr[i] : arr[0]);
    return fs;
}

template <typename T>
inline FieldSet<T> cur_sol(Level& lv) {
    FieldSet<T> fs;
    for (int i = 0; i < MG_MAX_FIELDS; ++i) {
        void* p = lv.sol[lv.cur[i]][i];
        fs.p[i] = static_cast<T*>(p ? p : lv.sol[lv.cur[0]][0]);
    }
    return fs;
}

template <int D>
inline dim3 node_grid(const Geom& g, int bx = 128) {
    return dim3((g.n + bx - 1) / bx, g.n, D == 3 ? g.nz : 1);
}


// ---- operator upload ------------------------------------------------------------------------------
template <typename T>
inline T host_value(const double* v);
template <>
inline double host_value<double>(const double* v) { return v[0]; }
template <>
inline cplx host_value<cplx>(const double* v) { return make_cplx(v[0], v[1]); }

template <typename T>
int upload_operator(mg_workspace* ws, Level& lv) {
    if (!lv.op_dirty) return MG_OK;
    OperatorTable<T> tab;
    memset(&tab, 0, sizeof(tab));
    for (int i = 0; i < ws->n_fields; ++i)
        for (int j = 0; j < ws->n_fields; ++j) {
            if (!lv.hop.set[i][j])
                return fail(MG_ERR_STATE, "operator block not set on level " + std::to_string(lv.level));
            tab.n_entries[i][j] = lv.hop.n_entries[i][j];
            for (int k = 0; k < lv.hop.n_entries[i][j]; ++k) {
                const int* o = lv.hop.offs[i][j][k];
                long long lin = o[0] + (long long)o[1] * lv.g.pitch + (ws->dim == 3 ? (long long)o[2] * lv.g.plane : 0);
                tab.off[i][j][k] = (int)lin;
                tab.val[i][j][k] = host_value<T>(lv.hop.val[i][j][k]);
            }
        }
    if (ws->vc_kind) {
        const int d = ws->dim;
        static const int want2[5][3] = {{0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}};
        static const int want3[7][3] = {{0, 0, -1}, {0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
        bool ok = lv.hop.n_entries[0][0] == 2 * d + 1;
        for (int k = 0; k < 2 * d + 1 && ok; ++k)
            for (int c = 0; c < d; ++c) ok = ok && lv.hop.offs[0][0][k][c] == (d == 2 ? want2[k][c] : want3[k][c]);
        if (!ok) return fail(MG_ERR_STATE, "variable-coefficient operator: the stencil must be the (2d+1)-point star");
        tab.vc_kind = ws->vc_kind;
        tab.vc_dim = d;
        tab.vc_n = lv.g.n;
        tab.vc_pitch = lv.g.pitch;
        tab.vc_plane = lv.g.plane;
        tab.vc_kappa = ws->vc_kappa;
        tab.vc_h = 1.0 / (double)(lv.g.n - 1);
    }
    if (!lv.d_op) MG_CUDA(cudaMalloc(&lv.d_op, sizeof(OperatorTable<T>)));
    MG_CUDA(cudaMemcpyAsync(lv.d_op, &tab, sizeof(tab), cudaMemcpyHostToDevice, ws->stream));
    MG_CUDA(cudaStreamSynchronize(ws->stream));
    // classify: scalar real star stencil in canonical order (z-, y-, x-, c, x+, y+, z+) with
    // symmetric coefficients per axis -> specialised kernels apply
    lv.star = false;
    if (ws->dtype == MG_F64 && ws->n_fields == 1 && !ws->vc_kind) {
        const HostOperator& h = lv.hop;
        const int d = ws->dim, ne = h.n_entries[0][0];
        if (ne == 2 * d + 1) {
            static const int want2[5][3] = {{0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}};
            static const int want3[7][3] = {{0, 0, -1}, {0, -1, 0}, {-1, 0, 0}, {0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
            bool ok = true;
            for (int k = 0; k < ne && ok; ++k)
                for (int c = 0; c < d; ++c)
                    ok = ok && h.offs[0][0][k][c] == (d == 2 ? want2[k][c] : want3[k][c]);
            if (ok) {
                if (d == 2) {
                    ok = h.val[0][0][0][0] == h.val[0][0][4][0] && h.val[0][0][1][0] == h.val[0][0][3][0];
                    lv.star_y = h.val[0][0][0][0]; lv.star_x = h.val[0][0][1][0]; lv.star_c = h.val[0][0][2][0];
                } else {
                    ok = h.val[0][0][0][0] == h.val[0][0][6][0] && h.val[0][0][1][0] == h.val[0][0][5][0] &&
                         h.val[0][0][2][0] == h.val[0][0][4][0];
                    lv.star_z = h.val[0][0][0][0]; lv.star_y = h.val[0][0][1][0]; lv.star_x = h.val[0][0][2][0];
                    lv.star_c = h.val[0][0][3][0];
                }
                lv.star = ok;
            }
        }
    }
    lv.op_dirty = false;
    return MG_OK;
}

// ---- boundary condition on the current solution buffers ----------------------------------------------
template <typename T>
inline int apply_bc(mg_workspace*, Level&) { return MG_OK; }
template <>
inline int apply_bc<cplx>(mg_workspace* ws, Level& lv) {
    if (ws->bc_kind != MG_BC_HELMHOLTZ_ROBIN) return MG_OK;
    double h = 1.0 / (double)(lv.g.n - 1);
    cplx c = t_div(make_cplx(1.0, 0.0), make_cplx(1.0, -ws->bc_k * h));
    int blocks = (lv.g.n + 127) / 128;
    for (int f = 0; f < ws->n_fields; ++f) {
        cplx* u = static_cast<cplx*>(lv.sol[lv.cur[f]][f]);
        launch_k(ws, k_bc_helmholtz, blocks, 128, 0, lv.g, u, c);
        launch_k(ws, k_bc_helmholtz_rows, blocks, 128, 0, lv.g, u);
        count(ws, 2);
    }
    return MG_OK;
}

// ---- stages ------------------------------------------------------------------------------------------
template <typename T, int D, int F>
int exec_smooth(mg_workspace* ws, const mg_op& op) {
    if (op.smoother < 0 || op.smoother >= (int)ws->smoothers.size()) return fail(MG_ERR_INVALID, "bad smoother handle");
    Smoother& sm = ws->smoothers[op.smoother];
    if (sm.level != op.level) return fail(MG_ERR_INVALID, "smoother registered for another level");
    Level& lv = ws->L(op.level);
    int rc = upload_operator<T>(ws, lv);
    if (rc) return rc;
    const OperatorTable<T>* tab = static_cast<const OperatorTable<T>*>(lv.d_op);
    const SmootherTable<T>* st = static_cast<const SmootherTable<T>*>(sm.d_table);
    const T* w = static_cast<const T*>(sm.d_w);
    FieldSet<T> rhs = fieldset<T>(lv.rhs);
    if (sm.point) {
        // op.src selects the colour passes of a coloured smoother: 0 = both (one op = one sweep),
        // 1 = first colour only, 2 = second colour only (domain decomposition exchanges halos in between)
        const int c_begin = (sm.colored && op.src == 2) ? 1 : 0;
        const int c_end = sm.colored ? (op.src == 1 ? 1 : 2) : 1;
        if (lv.zero_pending) {
            if (ws->use_fast && fast_smooth<T, D, F>(ws, lv, sm, c_begin, c_end, true)) {
                lv.zero_pending = false;
                return MG_OK;
            }
            rc = materialise_zero<T>(ws, lv);
            if (rc) return rc;
        }
        if (ws->use_fast && fast_smooth<T, D, F>(ws, lv, sm, c_begin, c_end, false)) return MG_OK;
        for (int c = c_begin; c < c_end; ++c) {
            FieldSet<T> src = cur_sol<T>(lv);
            FieldSet<T> dst = src;
            for (int s = 0; s < sm.m; ++s) {
                int f = sm.slot_field[s];
                dst.p[f] = static_cast<T*>(lv.sol[lv.cur[f] ^ 1][f]);
            }
            launch_k(ws, k_smooth_point<T, D, F>, node_grid<D>(lv.g), 128, 0, lv.g, tab, st, w, src, rhs, dst,
                                                                               sm.colored ? c : -1);
            count(ws);
            for (int s = 0; s < sm.m; ++s) lv.cur[sm.slot_field[s]] ^= 1;
            rc = apply_bc<T>(ws, lv);
            if (rc) return rc;
        }
    } else {
        rc = materialise_zero<T>(ws, lv);
        if (rc) return rc;
        if (ws->use_fast && fast_tile_smooth<T, D, F>(ws, lv, sm)) return MG_OK;
        FieldSet<T> src = cur_sol<T>(lv);
        FieldSet<T> dst = src;
        bool flip[MG_MAX_FIELDS] = {};
        for (int s = 0; s < sm.m; ++s) flip[sm.slot_field[s]] = true;
        for (int f = 0; f < F; ++f)
            if (flip[f]) dst.p[f] = static_cast<T*>(lv.sol[lv.cur[f] ^ 1][f]);
        int tiles[3] = {1, 1, 1};
        for (int d = 0; d < D; ++d) tiles[d] = (lv.g.n - 1) / sm.period[d] + 1;
        dim3 grid((tiles[0] + 63) / 64, tiles[1], tiles[2]);
        launch_k(ws, k_smooth_tile<T, D, F>, grid, 64, 0, lv.g, tab, st, w, src, rhs, dst, tiles[0]);
        count(ws);
        for (int f = 0; f < F; ++f)
            if (flip[f]) lv.cur[f] ^= 1;
        rc = apply_bc<T>(ws, lv);
        if (rc) return rc;
    }
    return MG_OK;
}

template <typename T, int D, int F>
int exec_residual(mg_workspace* ws, int level) {
    Level& lv = ws->L(level);
    int rc = upload_operator<T>(ws, lv);
    if (rc) return rc;
    if (ws->use_fast && fast_residual<T, D, F>(ws, lv)) return MG_OK;
    launch_k(ws, k_residual<T, D, F>, node_grid<D>(lv.g), 128, 0, 
        lv.g, static_cast<const OperatorTable<T>*>(lv.d_op), cur_sol<T>(lv), fieldset<T>(lv.rhs), fieldset<T>(lv.res));
    count(ws);
    return MG_OK;
}

template <typename T>
inline FieldSet<T> field_by_kind(mg_workspace* ws, Level& lv, int kind) {
    switch (kind) {
        case MG_RHS: return fieldset<T>(lv.rhs);
        case MG_RES: return fieldset<T>(lv.res);
        case MG_COR: return lv.level == ws->max_level ? fieldset<T>(lv.cor) : cur_sol<T>(lv);
        default: return cur_sol<T>(lv);
    }
}

template <typename T, int D, int F>
int exec_restrict(mg_workspace* ws, const mg_op& op, bool fused) {
    if (!ws->has_level(op.level) || !ws->has_level(op.level + 1)) return fail(MG_ERR_INVALID, "restrict: level out of range");
    Level& lc = ws->L(op.level);
    Level& lf = ws->L(op.level + 1);
    if (fused) {
        int rc = upload_operator<T>(ws, lf);
        if (rc) return rc;
        if (ws->use_fast && fast_residual_restrict<T, D, F>(ws, lf, lc)) return MG_OK;
        if (ws->vc_kind) {
            // variable coefficients: the fused generic kernel would evaluate the coefficients of 3^D fine nodes
            // per coarse node; residual + plain restriction evaluates each node once and is bit-identical
            launch_k(ws, k_residual<T, D, F>, node_grid<D>(lf.g), 128, 0, lf.g, static_cast<const OperatorTable<T>*>(lf.d_op),
                     cur_sol<T>(lf), fieldset<T>(lf.rhs), fieldset<T>(lf.res));
            launch_k(ws, k_restrict<T, D, F>, node_grid<D>(lc.g), 128, 0, lc.g, lf.g, fieldset<T>(lf.res), fieldset<T>(lc.rhs));
            count(ws, 2);
            return MG_OK;
        }
        if constexpr (D == 3) {
            // 3D star: the fused generic kernel recomputes 27 residuals per coarse node; the specialised
            // residual kernel + plain restriction moves fewer bytes and is bit-identical
            if (ws->use_fast && fast_residual<T, D, F>(ws, lf)) {
                launch_k(ws, k_restrict<T, D, F>, node_grid<D>(lc.g), 128, 0, lc.g, lf.g, fieldset<T>(lf.res), fieldset<T>(lc.rhs));
                count(ws);
                return MG_OK;
            }
        }
        launch_k(ws, k_residual_restrict<T, D, F>, node_grid<D>(lc.g), 128, 0, 
            lc.g, lf.g, static_cast<const OperatorTable<T>*>(lf.d_op), cur_sol<T>(lf), fieldset<T>(lf.rhs),
            fieldset<T>(lc.rhs));
    } else {
        launch_k(ws, k_restrict<T, D, F>, node_grid<D>(lc.g), 128, 0, lc.g, lf.g, field_by_kind<T>(ws, lf, op.src),
                                                                         fieldset<T>(lc.rhs));
    }
    count(ws);
    return MG_OK;
}

template <typename T, int D, int F>
int exec_prolong(mg_workspace* ws, const mg_op& op, bool add) {
    if (!ws->has_level(op.level) || !ws->has_level(op.level - 1)) return fail(MG_ERR_INVALID, "prolong: level out of range");
    Level& lf = ws->L(op.level);
    Level& lc = ws->L(op.level - 1);
    FieldSet<T> src = field_by_kind<T>(ws, lc, op.src);
    FieldSet<T> dst = add ? cur_sol<T>(lf) : field_by_kind<T>(ws, lf, MG_COR);
    if (!(add && ws->use_fast && fast_prolong_add<T, D, F>(ws, lf, lc, src, op.omega))) {
        launch_k(ws, k_prolong<T, D, F>, node_grid<D>(lf.g), 128, 0, lf.g, lc.g, src, dst, op.omega, add ? 1 : 0);
        count(ws);
    }
    if (add) return apply_bc<T>(ws, lf);
    return MG_OK;
}

template <typename T>
int launch_zero(mg_workspace* ws, Level& lv) {
    for (int f = 0; f < ws->n_fields; ++f) {
        // a kernel, not a memset node: keeps the programmatic-dependent-launch chain of the cycle intact
        // z-slabs: the second ghost planes (before plane 0, after plane nz-1) are part of the field
        const long long extra = lv.wide_ghost ? lv.g.plane : 0;
        const long long n16 = (long long)(lv.g.total + 2 * extra) * sizeof(T) / 16;        // pitch is a multiple of 16 elements
        const int blocks = (int)((n16 + 255) / 256 < 1184 ? (n16 + 255) / 256 : 1184);
        launch_k(ws, k_zero, blocks, 256, 0, reinterpret_cast<double2*>(static_cast<T*>(lv.sol[0][f]) - extra), n16);
        count(ws);
    }
    return MG_OK;
}

// `lazy`: only record that sol[0] is zero.  A specialised smoother that follows immediately consumes
// the flag (it substitutes 0.0 for every load of u: same arithmetic, a third less traffic and no
// fill); any other op that touches the level issues the fill first (materialise_zero).
template <typename T>
int exec_zero(mg_workspace* ws, int level, bool lazy = false) {
    Level& lv = ws->L(level);
    for (int f = 0; f < ws->n_fields; ++f) lv.cur[f] = 0;
    lv.zero_pending = lazy;
    return lazy ? MG_OK : launch_zero<T>(ws, lv);
}

template <typename T>
int materialise_zero(mg_workspace* ws, Level& lv) {
    if (!lv.zero_pending) return MG_OK;
    lv.zero_pending = false;
    return launch_zero<T>(ws, lv);
}

template <typename T>
int materialise_all_zeros(mg_workspace* ws) {
    for (auto& lv : ws->levels) {
        int rc = materialise_zero<T>(ws, lv);
        if (rc) return rc;
    }
    return MG_OK;
}

// everything that must not happen during stream capture: operator uploads, scratch allocation,
// dynamic shared memory opt-in
template <typename T, int D, int F>
int prepare(mg_workspace* ws) {
    for (auto& lv : ws->levels) {
        bool any = false;
        for (int i = 0; i < ws->n_fields; ++i)
            for (int j = 0; j < ws->n_fields; ++j) any = any || lv.hop.set[i][j];
        if (!any) continue;
        int rc = upload_operator<T>(ws, lv);
        if (rc) return rc;
    }
    Level& lv = ws->L(ws->min_level);
    const int nvec = ws->cgs_kind == MG_CGS_CG ? 4 : 7;
    size_t bytes = (size_t)nvec * F * lv.g.total * sizeof(T);
    ws->cgs_smem_bytes = bytes <= 200 * 1024 ? bytes : 0;
    if (!ws->tail) ws->tail = new TailProgram();
    ws->tail->h.n_cmds = 0;
    if (!ws->tail_pool) {
        ws->tail_ring_begin = TAIL_POOL_BYTES;
        ws->tail_pool_bytes = TAIL_POOL_BYTES + (size_t)TAIL_RING_SLOTS * sizeof(TailProgram);
        MG_CUDA(cudaMalloc(&ws->tail_pool, ws->tail_pool_bytes));
        MG_CUDA(cudaHostAlloc(&ws->tail_stage, (size_t)(TAIL_RING_SLOTS + 1) * sizeof(TailProgram), cudaHostAllocDefault));
        MG_CUDA(cudaStreamCreateWithFlags(&ws->aux_stream, cudaStreamNonBlocking));
    }
    // The opt-in is a property of the function, not of the workspace: several workspaces live in
    // one process (slab + replicated coarse levels, one per virtual rank), so always allow the
    // maximum instead of this workspace's size.
#ifdef MG_TAIL_TIMING
    const int smem_cap = 196 * 1024;           // the per-command clock array is static shared memory of the tail kernel
#else
    const int smem_cap = 200 * 1024;
#endif
    MG_CUDA(cudaFuncSetAttribute(k_coarse_cg<T, D, F, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    MG_CUDA(cudaFuncSetAttribute(k_coarse_bicgstab<T, D, F, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    MG_CUDA(cudaFuncSetAttribute(k_tail<T, D, F, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    MG_CUDA(cudaFuncSetAttribute(k_tail<T, D, F, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    if (ws->tail_ctas > 8) {
        MG_CUDA(cudaFuncSetAttribute(k_tail<T, D, F, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        MG_CUDA(cudaFuncSetAttribute(k_tail<T, D, F, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
    if (bytes > (size_t)smem_cap && ws->cgs_work_bytes < bytes) {
        if (ws->cgs_work) cudaFree(ws->cgs_work);
        MG_CUDA(cudaMalloc(&ws->cgs_work, bytes));
        ws->cgs_work_bytes = bytes;
    }
    return MG_OK;
}

template <typename T, int D, int F>
int exec_cgs(mg_workspace* ws, int level) {
    if (level != ws->min_level) return fail(MG_ERR_INVALID, "coarse-grid solver op must be on the workspace's min level");
    if (ws->slab) return fail(MG_ERR_INVALID, "the coarse-grid solver runs on whole grids, not on a z-slab");
    Level& lv = ws->L(level);
    if (lv.op_dirty) return fail(MG_ERR_STATE, "operator not uploaded (prepare not run)");
    for (int f = 0; f < F; ++f) lv.cur[f] = 0;
    const int nvec = ws->cgs_kind == MG_CGS_CG ? 4 : 7;
    size_t bytes = (size_t)nvec * F * lv.g.total * sizeof(T);
    const bool smem = bytes <= 200 * 1024;
    const OperatorTable<T>* tab = static_cast<const OperatorTable<T>*>(lv.d_op);
    FieldSet<T> sol = cur_sol<T>(lv), rhs = fieldset<T>(lv.rhs);
    T* gwork = static_cast<T*>(ws->cgs_work);
    double h = 1.0 / (double)(lv.g.n - 1);
    cplx bcf = t_div(make_cplx(1.0, 0.0), make_cplx(1.0, -ws->bc_k * h));
    int robin = ws->bc_kind == MG_BC_HELMHOLTZ_ROBIN ? 1 : 0;
    if (ws->cgs_kind == MG_CGS_CG) {
        if (smem) {
            launch_k(ws, k_coarse_cg<T, D, F, true>, 1, 256, bytes, lv.g, tab, sol, rhs, nullptr, ws->cgs_max_its, ws->cgs_reduction);
        } else {
            launch_k(ws, k_coarse_cg<T, D, F, false>, 1, 256, 0, lv.g, tab, sol, rhs, gwork, ws->cgs_max_its, ws->cgs_reduction);
        }
    } else {
        if (smem) {
            launch_k(ws, k_coarse_bicgstab<T, D, F, true>, 1, 256, bytes, lv.g, tab, sol, rhs, nullptr, ws->cgs_max_its, ws->cgs_reduction, robin, bcf);
        } else {
            launch_k(ws, k_coarse_bicgstab<T, D, F, false>, 1, 256, 0, lv.g, tab, sol, rhs, gwork, ws->cgs_max_its, ws->cgs_reduction, robin, bcf);
        }
    }
    count(ws);
    return MG_OK;
}

// sol += e on interior nodes
template <typename T, int D, int F>
__global__ void k_add_interior(Geom g, FieldSet<T> sol, FieldSet<T> e) {
    pdl_sync();
    int x, y, z;
    if (!node_coords<D>(g, x, y, z) || !is_interior<D>(g, x, y, z)) return;
    const long long idx = node_index<D>(g, x, y, z);
#pragma unroll
    for (int i = 0; i < F; ++i) sol.p[i][idx] = t_add(sol.p[i][idx], e.p[i][idx]);
}

// Krylov method as a smoother (MG_OP_KRYLOV, include/mgengine.h): the residual equation is solved from a
// zero guess into the idle solution buffers by the coarse-grid solver's kernel, then added
template <typename T, int D, int F>
int exec_krylov(mg_workspace* ws, const mg_op& op) {
    if (ws->slab) return fail(MG_ERR_INVALID, "Krylov smoothers run on whole grids, not on a z-slab");
    if (op.src != MG_CGS_CG && op.src != MG_CGS_BICGSTAB) return fail(MG_ERR_INVALID, "bad Krylov method");
    const int n_its = (int)op.omega;
    if (n_its < 1) return fail(MG_ERR_INVALID, "Krylov smoother without iterations");
    Level& lv = ws->L(op.level);
    int rc = materialise_zero<T>(ws, lv);
    if (rc) return rc;
    rc = exec_residual<T, D, F>(ws, op.level);
    if (rc) return rc;
    const OperatorTable<T>* tab = static_cast<const OperatorTable<T>*>(lv.d_op);
    FieldSet<T> scratch;
    for (int i = 0; i < MG_MAX_FIELDS; ++i) {
        const int f = lv.sol[0][i] ? i : 0;
        scratch.p[i] = static_cast<T*>(lv.sol[lv.cur[f] ^ 1][f]);
    }
    FieldSet<T> res = fieldset<T>(lv.res);
    const size_t bytes = (size_t)(op.src == MG_CGS_CG ? 4 : 7) * F * lv.g.total * sizeof(T);
    const bool smem = bytes <= 200 * 1024;
    if (!smem && ws->krylov_work_bytes < bytes) return fail(MG_ERR_STATE, "Krylov scratch not allocated");
    T* gwork = static_cast<T*>(ws->krylov_work);
    const double h = 1.0 / (double)(lv.g.n - 1);
    const cplx bcf = t_div(make_cplx(1.0, 0.0), make_cplx(1.0, -ws->bc_k * h));
    const int robin = ws->bc_kind == MG_BC_HELMHOLTZ_ROBIN ? 1 : 0;
    if (op.src == MG_CGS_CG) {
        if (smem) launch_k(ws, k_coarse_cg<T, D, F, true>, 1, 256, bytes, lv.g, tab, scratch, res, nullptr, n_its, ws->cgs_reduction);
        else launch_k(ws, k_coarse_cg<T, D, F, false>, 1, 256, 0, lv.g, tab, scratch, res, gwork, n_its, ws->cgs_reduction);
    } else {
        if (smem) launch_k(ws, k_coarse_bicgstab<T, D, F, true>, 1, 256, bytes, lv.g, tab, scratch, res, nullptr, n_its, ws->cgs_reduction, robin, bcf);
        else launch_k(ws, k_coarse_bicgstab<T, D, F, false>, 1, 256, 0, lv.g, tab, scratch, res, gwork, n_its, ws->cgs_reduction, robin, bcf);
    }
    launch_k(ws, k_add_interior<T, D, F>, node_grid<D>(lv.g), 128, 0, lv.g, cur_sol<T>(lv), scratch);
    count(ws, 2);
    return apply_bc<T>(ws, lv);
}

template <typename T, int D, int F>
int exec_norm(mg_workspace* ws, int level, int squares = 0) {
    Level& lv = ws->L(level);
    int rc = upload_operator<T>(ws, lv);
    if (rc) return rc;
    if (!(ws->use_fast && fast_norm<T, D, F>(ws, lv))) {
        launch_k(ws, k_norm_partial<T, D, F>, lv.norm_blocks, 256, 0, 
            lv.g, static_cast<const OperatorTable<T>*>(lv.d_op), cur_sol<T>(lv), fieldset<T>(lv.rhs),
            static_cast<T*>(lv.partial));
        lv.norm_count = lv.norm_blocks;
        count(ws);
    }
    launch_k(ws, k_norm_final<T>, 1, 256, 0, static_cast<const T*>(lv.partial), lv.norm_count, ws->d_norm, squares);
    count(ws);
    return MG_OK;
}

// ---- coarse tail: assemble commands instead of launching (mg_tail.cuh) ---------------------------------
inline bool tail_level_ok(const mg_workspace* ws, int level) {
    return ws->use_tail && ws->tail && ws->has_level(level) && level <= ws->tail_max_level &&
           level - ws->min_level < TAIL_MAX_LEVELS;
}

template <typename T>
inline void tail_set(void** dst, const FieldSet<T>& fs) {
    for (int i = 0; i < MG_MAX_FIELDS; ++i) dst[i] = fs.p[i];
}

inline TailCmd& tail_new(mg_workspace* ws, int kind, int level) {
    TailProgram& p = *ws->tail;
    TailCmd& c = p.cmd[p.h.n_cmds++];
    memset(&c, 0, sizeof(c));
    c.kind = kind;
    c.lv = level - ws->min_level;
    c.colour = -1;
    return c;
}

template <typename T, int D, int F>
int tail_flush(mg_workspace* ws) {
    if (!ws->tail || ws->tail->h.n_cmds == 0) return MG_OK;
    TailProgram& p = *ws->tail;
    p.h.cgs_kind = ws->cgs_kind;
    p.h.cgs_max_its = ws->cgs_max_its;
    p.h.cgs_reduction = ws->cgs_reduction;
    p.h.robin = ws->bc_kind == MG_BC_HELMHOLTZ_ROBIN ? 1 : 0;
    p.h.gwork = ws->cgs_work;
    p.h.cgs_in_smem = ws->cgs_smem_bytes > 0 ? 1 : 0;
    bool has_cgs = false;
    for (int i = 0; i < p.h.n_cmds; ++i) has_cgs = has_cgs || p.cmd[i].kind == T_CGS;
    for (int i = 0; i < p.h.n_cmds; ++i) p.cmd[i].next_wide = i + 1 < p.h.n_cmds ? p.cmd[i + 1].wide : 1;
    for (int level = ws->min_level; level <= ws->max_level && level - ws->min_level < TAIL_MAX_LEVELS; ++level) {
        Level& lv = ws->L(level);
        TailLevel& tl = p.h.lv[level - ws->min_level];
        tl.g = lv.g;
        tl.op = lv.d_op;
        double h = 1.0 / (double)(lv.g.n - 1);
        tl.bc_factor = t_div(make_cplx(1.0, 0.0), make_cplx(1.0, -ws->bc_k * h));
        tl.st.c = lv.star_c; tl.st.x = lv.star_x; tl.st.y = lv.star_y; tl.st.z = lv.star_z;
    }
    // upload: header + the commands in use.  A program of a cycle being captured keeps its slot until
    // the cycles are freed; a direct launch takes the next ring slot (the stream is drained when the
    // ring wraps, so no kernel still reads the slot that is overwritten).
    const size_t bytes = sizeof(TailHeader) + (size_t)p.h.n_cmds * sizeof(TailCmd);
    cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
    MG_CUDA(cudaStreamIsCapturing(ws->stream, &capturing));
    unsigned char* slot = nullptr;
    if (capturing != cudaStreamCaptureStatusNone) {
        const size_t need = (bytes + 255) & ~(size_t)255;
        if (ws->tail_pool_used + need > ws->tail_ring_begin) return fail(MG_ERR_STATE, "tail program pool exhausted");
        slot = ws->tail_pool + ws->tail_pool_used;
        ws->tail_pool_used += need;
        // the upload is not part of the cycle graph: it runs now, on its own stream (a synchronising call
        // is only allowed inside a capture sequence of this thread in relaxed mode)
        memcpy(ws->tail_stage, &p, bytes);
        cudaStreamCaptureMode mode = cudaStreamCaptureModeRelaxed;
        MG_CUDA(cudaThreadExchangeStreamCaptureMode(&mode));
        cudaError_t e1 = cudaMemcpyAsync(slot, ws->tail_stage, bytes, cudaMemcpyHostToDevice, ws->aux_stream);
        cudaError_t e2 = cudaStreamSynchronize(ws->aux_stream);
        MG_CUDA(cudaThreadExchangeStreamCaptureMode(&mode));
        if (e1 != cudaSuccess || e2 != cudaSuccess)
            return fail(MG_ERR_CUDA, std::string("tail program upload: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
    } else {
        // stream-ordered copy from the pinned ring slot: no host synchronisation except when the ring wraps
        if (ws->tail_ring_cursor == TAIL_RING_SLOTS) {
            MG_CUDA(cudaStreamSynchronize(ws->stream));
            ws->tail_ring_cursor = 0;
        }
        const size_t k = (size_t)ws->tail_ring_cursor++;
        slot = ws->tail_pool + ws->tail_ring_begin + k * sizeof(TailProgram);
        unsigned char* stage = static_cast<unsigned char*>(ws->tail_stage) + (k + 1) * sizeof(TailProgram);
        memcpy(stage, &p, bytes);
        MG_CUDA(cudaMemcpyAsync(slot, stage, bytes, cudaMemcpyHostToDevice, ws->stream));
    }
    // lean interpreter when every command is one a scalar real star-stencil cycle uses (mg_tail.cuh)
    bool lean = F == 1 && sizeof(T) == sizeof(double) && ws->cgs_kind == MG_CGS_CG && ws->tail_lean;
    for (int i = 0; i < p.h.n_cmds && lean; ++i) {
        const TailCmd& c = p.cmd[i];
        switch (c.kind) {
            case T_SMOOTH_POINT: case T_RESIDUAL: case T_RESIDUAL_RESTRICT: lean = c.star != 0; break;
            case T_CGS: lean = ws->use_fast && ws->L(ws->min_level).star; break;       // the lean CG holds the star in registers
            case T_RESTRICT: case T_ZERO: case T_PROLONG: break;
            default: lean = false;
        }
    }
    if (F == 1 && D == 2 && sizeof(T) == sizeof(cplx) && ws->cgs_kind == MG_CGS_BICGSTAB && ws->tail_lean && !ws->vc_kind) {
        // complex scalar cycle (Helmholtz): the lean interpreter holds the register-cached scalar paths, which need a
        // 5-entry operator on every level a command touches and single-slot point smoothers
        lean = true;
        for (int i = 0; i < p.h.n_cmds && lean; ++i) {
            const TailCmd& c = p.cmd[i];
            const int level = ws->min_level + c.lv;
            const int top = (c.kind == T_RESIDUAL_RESTRICT) ? level + 1 : level;
            const int ne = ws->has_level(top) ? ws->L(top).hop.n_entries[0][0] : 0;
            switch (c.kind) {
                case T_SMOOTH_POINT: lean = c.scalar1 != 0 && ne == 5; break;
                case T_RESIDUAL: case T_RESIDUAL_RESTRICT: lean = ne == 5; break;
                case T_RESTRICT: case T_ZERO: case T_PROLONG: case T_CGS: case T_BC: break;
                default: lean = false;
            }
        }
    }
    {
        // one cluster of tail_ctas CTAs (mg_tail.cuh); every CTA gets the Krylov scratch of CTA 0
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(ws->tail_ctas, 1, 1);
        cfg.blockDim = dim3(TAIL_NT, 1, 1);
        cfg.dynamicSmemBytes = has_cgs ? ws->cgs_smem_bytes : 0;
        cfg.stream = ws->stream;
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (ws->tail_ctas > 1) {
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = ws->tail_ctas;
            attr[na].val.clusterDim.y = 1;
            attr[na].val.clusterDim.z = 1;
            ++na;
        }
        if (ws->use_pdl) {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
        const TailProgram* dev = reinterpret_cast<const TailProgram*>(slot);
        if (lean) cudaLaunchKernelEx(&cfg, k_tail<T, D, F, true>, dev);
        else cudaLaunchKernelEx(&cfg, k_tail<T, D, F, false>, dev);
    }
    count(ws);
    p.h.n_cmds = 0;
    return MG_OK;
}

// -> 1: op appended to the tail program, 0: not eligible (caller flushes and launches), < 0: error
template <typename T, int D, int F>
int tail_append(mg_workspace* ws, const mg_op& op) {
    if (!ws->use_tail || !ws->tail) return 0;
    int top = op.level;                                        // highest level the op touches
    if (op.kind == MG_OP_RESTRICT || op.kind == MG_OP_RESIDUAL_RESTRICT) top = op.level + 1;
    int bottom = (op.kind == MG_OP_PROLONG_ADD || op.kind == MG_OP_PROLONG) ? op.level - 1 : op.level;
    if (!tail_level_ok(ws, top) || !ws->has_level(bottom)) return 0;
    if (ws->tail->h.n_cmds + 6 > TAIL_MAX_CMDS) {
        int rc = tail_flush<T, D, F>(ws);
        if (rc) return rc;
    }
    const bool robin = ws->bc_kind == MG_BC_HELMHOLTZ_ROBIN && sizeof(T) == sizeof(cplx);
    Level& lv = ws->L(op.level);
    // commands appended below for this op: wide (all CTAs of the cluster) unless the op stays on the
    // smallest levels; the Krylov solve and the boundary update always run on CTA 0
    struct WideMarker {
        TailProgram& p; int begin; bool wide;
        ~WideMarker() {
            for (int i = begin; i < p.h.n_cmds; ++i)
                p.cmd[i].wide = (wide && p.cmd[i].kind != T_CGS && p.cmd[i].kind != T_BC) ? 1 : 0;
        }
    } marker{*ws->tail, ws->tail->h.n_cmds, top > ws->tail_solo_level};
    switch (op.kind) {
        case MG_OP_SMOOTH: {
            if (op.smoother < 0 || op.smoother >= (int)ws->smoothers.size()) return fail(MG_ERR_INVALID, "bad smoother handle");
            Smoother& sm = ws->smoothers[op.smoother];
            if (sm.level != op.level) return fail(MG_ERR_INVALID, "smoother registered for another level");
            if (lv.op_dirty) return fail(MG_ERR_STATE, "operator not uploaded (prepare not run)");
            const bool coloured = sm.point && sm.colored;
            const int c_begin = (coloured && op.src == 2) ? 1 : 0;
            const int c_end = coloured ? (op.src == 1 ? 1 : 2) : 1;
            for (int c = c_begin; c < c_end; ++c) {
                FieldSet<T> src = cur_sol<T>(lv);
                FieldSet<T> dst = src;
                bool flip[MG_MAX_FIELDS] = {};
                for (int s = 0; s < sm.m; ++s) flip[sm.slot_field[s]] = true;
                for (int f = 0; f < F; ++f)
                    if (flip[f]) dst.p[f] = static_cast<T*>(lv.sol[lv.cur[f] ^ 1][f]);
                TailCmd& cmd = tail_new(ws, sm.point ? T_SMOOTH_POINT : T_SMOOTH_TILE, op.level);
                cmd.colour = (sm.point && sm.colored) ? c : -1;
                cmd.star = (ws->use_fast && lv.star && sm.point && sm.m == 1) ? 1 : 0;
                if (cmd.star && cmd.colour >= 0 && F == 1) {      // in place: no buffer flip, only the colour's nodes
                    cmd.inplace = 1;
                    dst = src;
                    flip[0] = false;
                }
                // complex scalar 5-point star (Helmholtz): the colour pass of the register-cached path runs in place
                // too -- a node of one colour reads only nodes of the other
                bool star_offsets = F == 1 && D == 2 && lv.hop.n_entries[0][0] == 5;
                {
                    static const int want2[5][2] = {{0, -1}, {-1, 0}, {0, 0}, {1, 0}, {0, 1}};
                    for (int k = 0; k < 5 && star_offsets; ++k)
                        star_offsets = lv.hop.offs[0][0][k][0] == want2[k][0] && lv.hop.offs[0][0][k][1] == want2[k][1];
                }
                if (sizeof(T) == sizeof(cplx) && star_offsets && sm.point && sm.m == 1 && cmd.colour >= 0 && !ws->vc_kind) {
                    cmd.inplace = 1;
                    dst = src;
                    flip[0] = false;
                }
                cmd.omega = sm.w00;
                cmd.sm = sm.d_table;
                cmd.w = sm.d_w;
                tail_set<T>(cmd.a, src);
                tail_set<T>(cmd.b, fieldset<T>(lv.rhs));
                tail_set<T>(cmd.c, dst);
                for (int f = 0; f < F; ++f)
                    if (flip[f]) lv.cur[f] ^= 1;
                // Robin columns: written by the sweep itself where the kernel's scalar path applies (mg_tail.cuh),
                // else by a boundary command after it
                const int ne00 = lv.hop.n_entries[0][0];
                const bool fused_bc = robin && F == 1 && D == 2 && sm.point && sm.m == 1 && !ws->vc_kind && ne00 == 5;
                cmd.bc = fused_bc ? 1 : 0;
                cmd.scalar1 = (sm.point && sm.m == 1) ? 1 : 0;
                if (robin && !fused_bc) tail_set<T>(tail_new(ws, T_BC, op.level).c, cur_sol<T>(lv));
            }
            return 1;
        }
        case MG_OP_RESIDUAL: {
            if (lv.op_dirty) return fail(MG_ERR_STATE, "operator not uploaded (prepare not run)");
            TailCmd& cmd = tail_new(ws, T_RESIDUAL, op.level);
            cmd.star = (ws->use_fast && lv.star) ? 1 : 0;
            tail_set<T>(cmd.a, cur_sol<T>(lv));
            tail_set<T>(cmd.b, fieldset<T>(lv.rhs));
            tail_set<T>(cmd.c, fieldset<T>(lv.res));
            return 1;
        }
        case MG_OP_RESTRICT:
        case MG_OP_RESIDUAL_RESTRICT: {
            Level& lf = ws->L(op.level + 1);
            if (lf.op_dirty && op.kind == MG_OP_RESIDUAL_RESTRICT) return fail(MG_ERR_STATE, "operator not uploaded");
            TailCmd& cmd = tail_new(ws, op.kind == MG_OP_RESTRICT ? T_RESTRICT : T_RESIDUAL_RESTRICT, op.level);
            cmd.star = (ws->use_fast && lf.star) ? 1 : 0;
            if (op.kind == MG_OP_RESTRICT) {
                tail_set<T>(cmd.a, field_by_kind<T>(ws, lf, op.src));
            } else {
                tail_set<T>(cmd.a, cur_sol<T>(lf));
                tail_set<T>(cmd.b, fieldset<T>(lf.rhs));
            }
            tail_set<T>(cmd.c, fieldset<T>(lv.rhs));
            return 1;
        }
        case MG_OP_ZERO: {
            for (int f = 0; f < F; ++f) lv.cur[f] = 0;
            tail_set<T>(tail_new(ws, T_ZERO, op.level).c, cur_sol<T>(lv));
            return 1;
        }
        case MG_OP_PROLONG_ADD:
        case MG_OP_PROLONG: {
            Level& lc = ws->L(op.level - 1);
            const bool add = op.kind == MG_OP_PROLONG_ADD;
            TailCmd& cmd = tail_new(ws, T_PROLONG, op.level);
            cmd.omega = op.omega;
            cmd.add = add ? 1 : 0;
            tail_set<T>(cmd.a, field_by_kind<T>(ws, lc, op.src));
            tail_set<T>(cmd.c, add ? cur_sol<T>(lv) : field_by_kind<T>(ws, lv, MG_COR));
            if (add && robin) tail_set<T>(tail_new(ws, T_BC, op.level).c, cur_sol<T>(lv));
            return 1;
        }
        case MG_OP_CGS: {
            if (ws->slab) return fail(MG_ERR_INVALID, "the coarse-grid solver runs on whole grids, not on a z-slab");
            if (op.level != ws->min_level) return fail(MG_ERR_INVALID, "coarse-grid solver op must be on the workspace's min level");
            if (lv.op_dirty) return fail(MG_ERR_STATE, "operator not uploaded (prepare not run)");
            for (int f = 0; f < F; ++f) lv.cur[f] = 0;
            TailCmd& cmd = tail_new(ws, T_CGS, op.level);
            tail_set<T>(cmd.b, fieldset<T>(lv.rhs));
            tail_set<T>(cmd.c, cur_sol<T>(lv));
            return 1;
        }
        default: return 0;
    }
}

template <typename T, int D, int F>
int exec_op(mg_workspace* ws, const mg_op& op) {
    if (!ws->has_level(op.level)) return fail(MG_ERR_INVALID, "op level out of range");
    int rc = tail_append<T, D, F>(ws, op);
    if (rc != 0) return rc < 0 ? rc : MG_OK;
    rc = tail_flush<T, D, F>(ws);
    if (rc) return rc;
    if (op.kind != MG_OP_SMOOTH && op.kind != MG_OP_ZERO)      // the op may read a lazily zeroed solution
        for (int l = op.level - 1; l <= op.level + 1; ++l)
            if (ws->has_level(l)) {
                rc = materialise_zero<T>(ws, ws->L(l));
                if (rc) return rc;
            }
    switch (op.kind) {
        case MG_OP_SMOOTH: return exec_smooth<T, D, F>(ws, op);
        case MG_OP_RESIDUAL: return exec_residual<T, D, F>(ws, op.level);
        case MG_OP_RESTRICT: return exec_restrict<T, D, F>(ws, op, false);
        case MG_OP_RESIDUAL_RESTRICT: return exec_restrict<T, D, F>(ws, op, true);
        case MG_OP_ZERO: return exec_zero<T>(ws, op.level, ws->lazy_zero && ws->use_fast && !ws->slab);
        case MG_OP_PROLONG_ADD: return exec_prolong<T, D, F>(ws, op, true);
        case MG_OP_PROLONG: return exec_prolong<T, D, F>(ws, op, false);
        case MG_OP_CGS: return exec_cgs<T, D, F>(ws, op.level);
        case MG_OP_KRYLOV: return exec_krylov<T, D, F>(ws, op);
        default: return fail(MG_ERR_INVALID, "unknown op kind");
    }
}

}  // namespace mg
