// mg_workspace.cuh -- device storage of one solver instance: level hierarchy, operator tables,
// smoother tables, captured cycles.  Layout in HBM (DESIGN.md "Data layout"):
//   per level and field: sol[0], sol[1] (ping-pong for out-of-place sweeps), rhs, res, and on the
//   finest level cor and init; each array is n^(d-1) rows of `pitch` elements, pitch = n rounded
//   up to 16 elements so that every row starts on a 128-byte boundary (256 B for complex).
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <string>
#include <vector>

#include "../../include/mgengine.h"
#include "mg_common.cuh"
#include "mg_tail.cuh"

namespace mg {

extern thread_local std::string g_last_error;

inline int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define MG_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return mg::fail(MG_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));      \
    } while (0)

struct HostOperator {          // dtype-agnostic host copy (values as re/im pairs)
    int n_entries[MG_MAX_FIELDS][MG_MAX_FIELDS] = {};
    int offs[MG_MAX_FIELDS][MG_MAX_FIELDS][MG_MAX_ENTRIES][MG_MAX_DIM] = {};
    double val[MG_MAX_FIELDS][MG_MAX_FIELDS][MG_MAX_ENTRIES][2] = {};
    bool set[MG_MAX_FIELDS][MG_MAX_FIELDS] = {};
};

struct Level {
    int level = 0;
    Geom g{};
    void* slab = nullptr;
    size_t slab_bytes = 0;
    void* sol[2][MG_MAX_FIELDS] = {};
    void* rhs[MG_MAX_FIELDS] = {};
    void* res[MG_MAX_FIELDS] = {};
    void* cor[MG_MAX_FIELDS] = {};      // finest only
    void* init[MG_MAX_FIELDS] = {};     // finest only: saved initial guess
    int cur[MG_MAX_FIELDS] = {};        // which sol buffer is current
    bool zero_pending = false;          // sol[0] is logically zero; the fill has not been issued (mg_exec.cuh)
    bool wide_ghost = false;            // z-slab: one more ghost plane lies before local plane 0 and after nz-1
    HostOperator hop;
    void* d_op = nullptr;               // OperatorTable<T>
    bool op_dirty = true;
    // fast-path classification (filled when the operator is uploaded)
    bool star = false;                  // scalar real (2d+1)-point stencil with symmetric structure
    double star_c = 0, star_x = 0, star_y = 0, star_z = 0;
    void* partial = nullptr;            // norm partials
    int norm_blocks = 0;
    int norm_count = 0;                 // partials written by the last norm kernel
};

struct Smoother {
    int level = 0;
    bool jacobi = true, colored = false;
    int m = 0;
    bool point = true;
    int period[MG_MAX_DIM] = {1, 1, 1};
    int slot_field[MG_MAX_SLOTS] = {};
    void* d_table = nullptr;            // SmootherTable<T>
    void* d_w = nullptr;                // W tables
    // point smoother on a star stencil, single slot: scale = W[0][0] (real)
    double w00 = 0;
    bool same_colour_coupling = false;  // stencil couples nodes of equal colour -> no in-place colour sweep
    int full_index = -1;                // index of the all-valid mask in the W tables
    bool canonical_slots = false;       // slot s sits at (s / period[1], s % period[1]) of field 0 (2D scalar tiles)
};

// domain decomposition (mg_dd.cuh, mg_engine_dd.inc): what a rank knows about a peer's level
struct DDPeerLevel {
    char* base = nullptr;                   // the peer's level allocation, mapped into this process
    long long off_sol[2] = {0, 0}, off_rhs = 0, off_res = 0;       // byte offsets of its field arrays (field 0)
    int nz = 0, z0 = 0;
};

struct GraphVariant {
    cudaGraphExec_t exec = nullptr;
    int end_state = 0;
    long long kernels = 0;
};

// kernels captured between mg_capture_begin / mg_capture_end (the FAS statements, mg_enqueue_ops): valid for the
// ping-pong state of the solution buffers it was captured in
struct CapturedGraph {
    cudaGraphExec_t exec = nullptr;
    std::vector<int> begin, end;               // lv.cur[f] of every level and field before / after the captured calls
    long long kernels = 0;
};

struct Cycle {
    std::vector<mg_op> ops;
    std::map<int, GraphVariant> variants;      // keyed by the finest level's buffer state
};

}  // namespace mg

struct mg_workspace {
    int dim = 2, min_level = 0, max_level = 0, n_fields = 1, dtype = MG_F64, device = 0;
    size_t elem = 8;
    bool slab = false;                          // 3D z-slab of a domain-decomposed grid (mg_create_slab)
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;          // set once a caller's stream has been installed (mg_set_stream)
    std::vector<mg::Level> levels;              // index = level - min_level
    std::vector<mg::Smoother> smoothers;
    std::vector<mg::Cycle> cycles;
    std::vector<mg::CapturedGraph> captured;    // mg_capture_begin / mg_capture_end
    bool capturing = false;
    std::vector<int> capture_begin_state;
    long long capture_begin_count = 0;
    int bc_kind = MG_BC_DIRICHLET;
    double bc_k = 0.0;
    int vc_kind = MG_VC_NONE;                   // variable-coefficient operator (mg_set_variable_coefficient)
    double vc_kappa = 0.0;
    int cgs_kind = MG_CGS_CG, cgs_max_its = 1000;
    double cgs_reduction = 1e-12;
    void* cgs_work = nullptr;                   // global scratch when the coarse level exceeds smem
    size_t cgs_work_bytes = 0;
    void* krylov_work = nullptr;                // global scratch of the Krylov smoother ops (MG_OP_KRYLOV)
    size_t krylov_work_bytes = 0;
    double* d_norm = nullptr;                   // device scalar
    double* h_norm = nullptr;                   // pinned host scalar
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, mark[2] = {nullptr, nullptr};
    void* flush_buf = nullptr;
    size_t flush_bytes = 0;
    double time_limit_s = 0.0;                  // wall-clock cap of one solve (mg_set_time_limit; 0 = none)
    long long launches = 0;                     // kernels launched (incl. graph replays)
    long long pending_count = 0;                // kernels issued since the counter was last sampled
    bool use_fast = true;                       // specialised kernels where applicable
    bool lazy_zero = true;              // defer zero fills so that the next smoother can skip reading u
    bool use_pdl = false;                       // programmatic dependent launch between consecutive kernels (MG_PDL=1):
                                                // measured neutral inside captured graphs, so off by default
    bool use_tail = true;                       // small levels run inside the single-CTA tail kernel
    int tail_max_level = 6;                     // levels <= this are "small" (MG_TAIL_LEVEL overrides)
    int tail_ctas = 1;                          // CTAs of the tail kernel's cluster (MG_TAIL_CTAS; 1: single CTA)
    int tail_solo_level = 4;                    // levels <= this are run by CTA 0 alone (MG_TAIL_SOLO)
    bool tail_lean = true;                      // use the lean interpreter when a program qualifies (MG_TAIL_LEAN)
    mg::TailProgram* tail = nullptr;            // command list being assembled (host copy)
    // device copies of the tail programs: [0, tail_ring_begin) holds the programs of captured cycles
    // (bump allocation, reset when the cycles are freed), the rest is a ring for direct launches
    unsigned char* tail_pool = nullptr;
    size_t tail_pool_bytes = 0, tail_pool_used = 0, tail_ring_begin = 0;
    int tail_ring_cursor = 0;
    void* tail_stage = nullptr;                 // pinned staging buffer for the upload
    cudaStream_t aux_stream = nullptr;          // uploads that must not be captured into the cycle graph
    // preconditioned BiCGStab outer solver (mg_outer.cuh)
    mg::HostOperator hop_outer;                 // unshifted operator A of the finest level
    bool outer_set = false, outer_dirty = true;
    void* d_op_outer = nullptr;
    void* outer_slab = nullptr;
    void* outer_vec[16] = {};
    void* d_outer_scalars = nullptr;
    cudaGraphExec_t outer_graph = nullptr;
    int outer_graph_cycle = -1;
    long long outer_graph_kernels = 0;
    size_t cgs_smem_bytes = 0;                  // dynamic shared memory of the coarse solver (0: global scratch)
    // domain decomposition: this workspace is the z-slab of `dd_rank` (mg_engine_dd.inc)
    int dd_rank = -1, dd_world = 0;
    int dd_local_peers = 0;                     // other ranks living in this process (virtual ranks on the same GPU)
    void* dd_comm = nullptr;                    // mg::DDComm of this rank (device)
    void* dd_peer_comm[16] = {};                // every rank's DDComm, mapped (own included)
    std::vector<mg::DDPeerLevel> dd_lo, dd_hi;  // neighbours' levels, index = level - min_level
    std::vector<mg::DDPeerLevel> dd_base;       // every rank's coarsest slab level (index = rank): gather source
    mg_workspace* dd_coarse = nullptr;          // replicated whole-grid hierarchy below the slab levels
    std::map<std::string, void*> dd_ipc_opened; // CUDA IPC handle -> mapping (a handle is opened once)
    std::vector<mg::Cycle> dd_cycles;
    bool dd_use_graph = true;
    // nonlinear FAS stages (mg_fas.cuh): Approximation field per level, allocated on first use
    std::vector<void*> fas_approx;              // index = level - min_level (points past the guard rows)
    std::vector<void*> fas_raw;                 // the allocations behind them

    mg::Level& L(int level) { return levels[level - min_level]; }
    bool has_level(int level) const { return level >= min_level && level <= max_level; }
    int finest_state() {
        int s = 0;
        for (int f = 0; f < n_fields; ++f) s |= L(max_level).cur[f] << f;
        return s;
    }
    void set_finest_state(int s) {
        for (int f = 0; f < n_fields; ++f) L(max_level).cur[f] = (s >> f) & 1;
    }
};

namespace mg {

inline void count(mg_workspace* ws, int k = 1) { ws->pending_count += k; }

// every kernel goes through here: launch on the workspace stream, with the programmatic-dependent-
// launch attribute (mg_common.cuh pdl_sync) unless disabled (MG_PDL=0)
template <typename... KArgs, typename... Args>
inline void launch_k(mg_workspace* ws, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ws->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = ws->use_pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace mg
