/*
 * mgengine.h -- C ABI of the B200-native multigrid cycle engine.
 *
 * This is the only native boundary of the system.  It replaces, for one evolved
 * solver, everything the reference does *below* ProgramGenerator.generate_and_evaluate
 * (/root/reference/evostencils/code_generation/exastencils.py:485-537): ExaSlang
 * emission (:318-336, :684-925), the two ExaStencils compiler runs (:381-409, :495-503),
 * `make` (:411-415) and the runs of the generated binary whose stdout is parsed
 * (:417-443, :539-584).  The reference has no FFI for this path (it shells out to
 * java / make / ./exastencils); the entry points below are what a binding for the path
 * needs: set up the level hierarchy and operators once per problem, hand over one
 * lowered op list per individual, solve, read back the residual history.
 *
 * Conventions
 *   - every function returns 0 on success and a negative mg_status on failure; nothing
 *     throws across the boundary; mg_last_error() returns a thread-local message;
 *   - plain pointers and sizes only; host arrays are dense, row-major
 *     [i_{d-1}]...[i1][i0] with n = 2^level + 1 nodes per dimension
 *     (exastencils.py:97-103), complex values as interleaved (re, im) doubles;
 *   - arrays passed in are borrowed for the duration of the call;
 *   - one workspace is bound to one CUDA device and one stream; calls on the same
 *     workspace must be serialised by the caller (the reference evaluates individuals
 *     strictly sequentially per generator, optimization/program.py:491, 577).
 */
#ifndef MGENGINE_H
#define MGENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mg_workspace mg_workspace;

typedef enum {
    MG_OK = 0,
    MG_ERR_INVALID = -1,      /* bad argument / unsupported construct */
    MG_ERR_CUDA = -2,         /* a CUDA runtime call failed */
    MG_ERR_STATE = -3,        /* call sequence violated (e.g. operator missing) */
    MG_ERR_NOMEM = -4,
    MG_ERR_TIMEOUT = -5       /* a solve exceeded the limit set by mg_set_time_limit */
} mg_status;

typedef enum { MG_F64 = 0, MG_C128 = 1 } mg_dtype;

/* field kinds of one level's storage (CycleStorage, exastencils.py:14-26) */
typedef enum { MG_SOL = 0, MG_RHS = 1, MG_RES = 2, MG_COR = 3 } mg_field;

/* boundary treatment of solution-type fields */
typedef enum {
    MG_BC_DIRICHLET = 0,        /* boundary nodes are never written */
    MG_BC_HELMHOLTZ_ROBIN = 1   /* x-boundaries slaved: u_0 = u_1 / (1 - i k h); y-boundaries 0
                                   (Helmholtz/2D_FD_Helmholtz_fromL3.exa4:25-59) */
} mg_bc;

typedef enum { MG_CGS_CG = 0, MG_CGS_BICGSTAB = 1 } mg_cgs_kind;

/* variable-coefficient operators (stencils/gallery.py:87-185) */
typedef enum {
    MG_VC_NONE = 0,
    MG_VC_EXP_KAPPA = 1         /* -div(c grad u), c(x) = exp(kappa * prod_d (x_d - x_d^2)) on the unit square / cube
                                   (get_coefficient_2D :87-90, get_coefficient_3D :137-140) */
} mg_vc_kind;

/* op kinds: the statement vocabulary generate_multigrid emits (exastencils.py:684-925) */
typedef enum {
    MG_OP_SMOOTH = 0,             /* solve locally ... relax w         (:769-822)  */
    MG_OP_RESIDUAL = 1,           /* res = rhs - A*sol                  (:827-853)  */
    MG_OP_RESTRICT = 2,           /* rhs@level = R * src@(level+1)      (:855-873)  */
    MG_OP_RESIDUAL_RESTRICT = 3,  /* fused 1+2, residual field not stored            */
    MG_OP_ZERO = 4,               /* sol@level = 0                      (:706-709)  */
    MG_OP_PROLONG_ADD = 5,        /* sol@level += w * P * src@(level-1) (:727-743)  */
    MG_OP_PROLONG = 6,            /* cor@level = P * src@(level-1)      (:855-873)  */
    MG_OP_CGS = 7,                /* coarse-grid solve on min_level     (:874-911)  */
    MG_OP_KRYLOV = 8              /* `<Name>_<n>@level()` in the place of a smoother (:758-768;
                                     ir/krylov_subspace.py:4-45): src = mg_cgs_kind, omega = n.
                                     res = rhs - A sol; at most n iterations of the coarse-grid solver's
                                     method on A e = res from e = 0 (early exit at the reduction set by
                                     mg_set_coarse_solver); sol += e.  One CTA: meant for the small levels */
} mg_op_kind;

typedef struct {
    int32_t kind;        /* mg_op_kind */
    int32_t level;       /* level of the field that is written */
    double  omega;       /* relaxation factor (SMOOTH uses the one baked into its W table); KRYLOV: iterations */
    int32_t src;         /* mg_field read by RESTRICT / PROLONG*; SMOOTH with a coloured smoother:
                            0 = both colour passes, 1 = first colour only, 2 = second colour only;
                            KRYLOV: mg_cgs_kind */
    int32_t smoother;    /* SMOOTH: handle returned by mg_add_smoother */
} mg_op;

/* ---- lifetime ------------------------------------------------------------------------ */
int mg_create(mg_workspace** out, int dimension, int min_level, int max_level, int n_fields,
              int dtype, int device);
/* 3D z-slab of a domain-decomposed grid (BASELINE.json configs[2]; no reference counterpart: the
 * reference runs one block / one fragment, example_problems/lib/domain_onePatch.knowledge:4-9).
 * On level min_level + i the workspace owns the global interior planes z_begin[i] .. z_end[i] and
 * holds one more plane on either side (a neighbour's data, filled by the caller's halo exchange, or
 * the global Dirichlet plane).  Host arrays of mg_upload / mg_download are [nz][n][n]. */
int mg_create_slab(mg_workspace** out, int min_level, int max_level, int n_fields, int dtype, int device,
                   const int32_t* z_begin, const int32_t* z_end);
int mg_destroy(mg_workspace* ws);
const char* mg_last_error(void);
/* number of kernels launched (directly or through graph replays) since creation */
int mg_launch_count(mg_workspace* ws, int64_t* out);

/* ---- problem definition (once per problem / level range) ------------------------------ */
/* Constant stencil of block (row, col) on `level`; offsets[n][dimension] with component 0
 * = x; values[n] (or [n][2] for MG_C128), in the accumulation order to be used. */
int mg_set_operator(mg_workspace* ws, int level, int row, int col, int n_entries,
                    const int32_t* offsets, const double* values);
/* Variable-coefficient Poisson operator on every level of a scalar MG_F64 workspace: replaces the
 * `Operator A from Stencil` text of Poisson2DVariableCoefficients / Poisson3DVariableCoefficients.generate_exa3
 * (stencils/gallery.py:113-131, :163-185) and the getCoefficient calls ExaStencils expands from it.  The
 * kernels evaluate the coefficient at the half-grid positions around each node (nothing is stored: 24 B/DoF
 * per sweep as for constant coefficients): off-centre value (-1.0 * c_face) / (h*h), centre
 * ((cE + cW)/(h*h)) + ((cN + cS)/(h*h)) [+ ((cT + cB)/(h*h))].  Values passed to mg_set_operator are then
 * ignored (its offsets must be the (2d+1)-point star in memory order; set here when absent).  Point
 * smoothers only (Jacobi / red-black): their W table holds omega alone (1 slot, 1 mask) and the kernels
 * divide by the node's centre value.  kind = MG_VC_NONE restores constant stencils. */
int mg_set_variable_coefficient(mg_workspace* ws, int kind, double kappa);
int mg_set_boundary(mg_workspace* ws, int bc_kind, double k);
int mg_set_coarse_solver(mg_workspace* ws, int kind, int max_iterations, double reduction);

/* One `solve locally` statement shape.  period[d] = tile shape (1 = point smoother);
 * slots = unknowns of the local system (field, position in tile); the table holds
 * W = omega * inv(L) for every validity mask of the slots that occurs
 * (bit s of mask set <=> slot s is an interior node): w_table[n_masks][n_slots][n_slots]
 * doubles (x2 for complex), row-major.  jacobi != 0: values are read from the old
 * iterate; colored != 0: red-black, colour (sum i_d) % 2 == 0 first. */
int mg_add_smoother(mg_workspace* ws, int level, int jacobi, int colored, const int32_t* period,
                    int n_slots, const int32_t* slot_field, const int32_t* slot_pos,
                    int n_masks, const uint32_t* masks, const double* w_table, int* handle);
int mg_clear_smoothers(mg_workspace* ws);

/* ---- data ------------------------------------------------------------------------------ */
int mg_upload(mg_workspace* ws, int field_kind, int level, int field, const void* host);
int mg_download(mg_workspace* ws, int field_kind, int level, int field, void* host);
/* copy of device-resident dense data (same layout as the host arrays) */
int mg_upload_device(mg_workspace* ws, int field_kind, int level, int field, const void* device_ptr);
/* device address and geometry of a field array (current solution buffer for MG_SOL): rows of `pitch`
 * elements, n rows per plane, nz planes, local plane 0 = global plane z0.  For zero-copy halo
 * exchange by the caller (NCCL send/recv of whole planes, which are contiguous). */
int mg_field_pointer(mg_workspace* ws, int field_kind, int level, int field, void** device_ptr, int32_t* n,
                     int32_t* pitch, int32_t* nz, int32_t* z0);
/* remember / restore the finest-level solution (initial guess for repeated samples) */
int mg_save_initial_guess(mg_workspace* ws);
int mg_restore_initial_guess(mg_workspace* ws);

/* ---- execution ------------------------------------------------------------------------- */
/* run ops immediately on the workspace stream (no graph); used by the parity tests */
int mg_exec_ops(mg_workspace* ws, const mg_op* ops, int n_ops);
/* issue the ops on the workspace stream without waiting for them; mg_set_stream installs a stream
 * of the caller (NULL: back to the workspace's own), so that the caller's NCCL halo exchanges and
 * the engine's kernels are ordered by the stream, not by host synchronisation */
int mg_enqueue_ops(mg_workspace* ws, const mg_op* ops, int n_ops);
int mg_set_stream(mg_workspace* ws, void* cuda_stream);
/* capture the op list (plus the fused residual-norm of the finest level) as a CUDA graph */
int mg_build_cycle(mg_workspace* ws, const mg_op* ops, int n_ops, int* cycle);
int mg_free_cycles(mg_workspace* ws);
/* Caller-driven capture (the FAS statements below are separate calls, exastencils_FAS.py:99-319): the kernels that
 * mg_fas_smooth / mg_fas_residual / mg_fas_restrict / mg_fas_subtract_approximation / mg_enqueue_ops issue between
 * begin and end become ONE CUDA graph instead of running; synchronising calls (mg_exec_ops, norms, uploads) are not
 * allowed in between, and everything the calls allocate on first use must exist (run the sequence once before).
 * mg_graph_launch enqueues n replays; it returns MG_ERR_STATE when the solution buffers are not in the ping-pong
 * state of the capture (a sequence with an odd number of Jacobi sweeps on a level alternates between two states:
 * capture both).  mg_free_cycles drops the graphs. */
int mg_capture_begin(mg_workspace* ws);
int mg_capture_end(mg_workspace* ws, int* graph);
int mg_graph_launch(mg_workspace* ws, int graph, int n);
/* replay `n_cycles` cycles without convergence checks (throughput measurements) */
int mg_run_cycles(mg_workspace* ws, int cycle, int n_cycles, float* ms);
/* sqrt(sum over interior nodes and fields of (rhs - A sol)^2) on `level` */
int mg_residual_norm(mg_workspace* ws, int level, double* out);
/* outer iteration of the generated solver: res_hist[0] = initial residual norm, then one
 * cycle + norm per iteration until norm < tol * res_hist[0] (strict), a non-finite norm, or
 * max_its.  *n_its = iterations done; *converged = 1 iff the tolerance was met.
 * res_hist must hold max_its + 1 doubles.  *ms = device time of the solve phase. */
int mg_solve(mg_workspace* ws, int cycle, double tol, int max_its, double* res_hist,
             int* n_its, int* converged, float* ms);
/* Wall-clock cap (seconds, 0 = none) of one mg_solve / mg_solve_bicgstab call: the reference kills the
 * generated binary after `timeout_evaluate` seconds and scores the individual with the sentinel triple
 * (code_generation/exastencils.py:42, 424-431).  Checked after every outer iteration; an expired solve
 * returns MG_ERR_TIMEOUT with res_hist / n_its filled up to that iteration. */
int mg_set_time_limit(mg_workspace* ws, double seconds);
/* Preconditioned BiCGStab outer solver of the Helmholtz configuration
 * (example_problems/Helmholtz/2D_FD_Helmholtz_fromL3.exa3:144-200): the cycle is the preconditioner
 * (u = 0; f = p; cycle), `A` the unshifted operator of the finest level set by mg_set_outer_operator.
 * Same history / stopping conventions as mg_solve; one CUDA graph per BiCGStab iteration. */
int mg_set_outer_operator(mg_workspace* ws, int n_entries, const int32_t* offsets, const double* values);
int mg_solve_bicgstab(mg_workspace* ws, int cycle, double tol, int max_its, double* res_hist, int* n_its,
                      int* converged, float* ms);
int mg_synchronize(mg_workspace* ws);

/* ---- measurement (CUDA events on the workspace stream) ---------------------------------- */
/* issue the op list `reps` times back to back; *ms = device time of all repetitions */
int mg_time_ops(mg_workspace* ws, const mg_op* ops, int n_ops, int reps, float* ms);
/* run the op list once with a CUDA event after every op: ms_per_op[i] = device time of op i
 * (launch gaps included); a per-stage view of one cycle without a profiler */
int mg_profile_ops(mg_workspace* ws, const mg_op* ops, int n_ops, float* ms_per_op);
/* record mark 0 / mark 1 on the workspace stream; elapsed = time between the two marks */
int mg_mark(mg_workspace* ws, int which);
int mg_elapsed(mg_workspace* ws, float* ms);
/* write `bytes` bytes of scratch (allocated on first use) to evict the L2 between timed runs */
int mg_flush_l2(mg_workspace* ws, size_t bytes);
/* bit 0: specialised kernels where applicable, bit 1: single-CTA tail kernel for the small levels
 * (default 3 = both; 0 = generic grid kernels only) */
int mg_set_fast_kernels(mg_workspace* ws, int enabled);

/* ---- 3D domain decomposition with peer-to-peer halo exchange (BASELINE.json configs[2]) ------------ */
/* No reference counterpart (one block / one fragment: example_problems/lib/domain_onePatch.knowledge:4-9,
 * parallelization_pureOmp.knowledge:5).  A rank = a z-slab workspace (mg_create_slab; it carries two ghost
 * planes on either side) + a whole-grid workspace of the levels below, run redundantly by every rank.
 * Ranks store halo planes straight into each other's memory (NVLink; CUDA IPC between processes) and
 * order them with flag words -- no collective-library call inside a cycle, one CUDA graph per cycle and rank.
 * Setup: mg_dd_init on every rank; mg_dd_describe for the comm block (level < 0) and every slab level; the
 * caller ships descriptors + IPC handles between the ranks; mg_dd_attach for every peer's comm block, the two
 * neighbours' slab levels and every rank's coarsest slab level.  Then every rank enqueues the same sequence of
 * mg_dd_exchange / mg_dd_norm / mg_dd_launch calls.  Results are bitwise those of the single-domain solve
 * (norms: the rank-ordered sum of the ranks' partial sums of squares). */
int mg_dd_init(mg_workspace* slab, int rank, int world, mg_workspace* coarse);
int mg_dd_describe(mg_workspace* slab, int level, uint64_t* desc8, void* ipc_handle64);
int mg_dd_attach(mg_workspace* slab, int peer, int level, const uint64_t* desc8, const void* ipc_handle64);
/* halo exchange of 1 or 2 boundary planes of a field (MG_SOL / MG_RHS / MG_RES) with both neighbours, enqueued */
int mg_dd_exchange(mg_workspace* slab, int field_kind, int level, int planes);
/* the unfused op list of one cycle (every rank the same; smoother handles of ops on levels <= the coarsest slab
 * level are the coarse workspace's); the engine inserts exchanges, the gather of the coarsest slab level's
 * right-hand side and the fused kernels where they apply */
int mg_dd_build_cycle(mg_workspace* slab, const mg_op* ops, int n_ops, int* cycle);
int mg_dd_prepare(mg_workspace* slab, int cycle);              /* instantiate the graphs outside a timed region */
int mg_dd_launch(mg_workspace* slab, int cycle, int n);        /* enqueue n cycles, each followed by the norm */
int mg_dd_norm(mg_workspace* slab);                            /* enqueue the combined residual norm alone */
int mg_dd_read_norm(mg_workspace* slab, double* out);          /* synchronise; MG_ERR_STATE if a peer never arrived */
int mg_dd_error(mg_workspace* slab, int* flag);

/* ---- nonlinear FAS stages ------------------------------------------ */
/* -Laplace(u) + gamma u e^u = f: the statements of example_problems/FAS_2D_Basic/FAS_2D_Basic.exa4
 * (replaces the Layer-4 text ProgramGeneratorFAS.traverse_graph emits, code_generation/
 * exastencils_FAS.py:99-319, and the ExaStencils compile behind it).  2D scalar MG_F64 workspaces whose
 * operator on the level is the 5-point Laplace star; direct launches on the workspace stream. */
/* `sweeps` smoother sweeps.  newton_steps >= 1: Jacobi-Newton with that many local Newton updates per
 * node, newton_steps == 0: Jacobi-Picard (exastencils_FAS.py:206-223); colored != 0: red-black in place,
 * colour (i0 + i1) % 2 == 0 first (:236-246), else Jacobi into the other solution slot.
 * (omega 0.8, newton_steps 1, colored 0 = Smoother@all of the template, .exa4:65-75; 30 of those =
 * CGS@coarsest, :58-63) */
int mg_fas_smooth(mg_workspace* ws, int level, double omega, double gamma, int sweeps, int newton_steps,
                  int colored);
/* Residual@level = RHS - (L u + gamma e^u u), boundary 0 (.exa4:77-82) */
int mg_fas_residual(mg_workspace* ws, int level, double gamma);
/* `level` = the coarse level: Approximation = R Solution@(level+1); Solution@level = Approximation;
 * RHS@level = R Residual@(level+1) + (L Approximation + gamma e^Approximation Approximation)
 * (.exa4:219-227; exastencils_FAS.py:120-149) */
int mg_fas_restrict(mg_workspace* ws, int level, double gamma);
/* Solution@level -= Approximation@level (.exa4:247-249; exastencils_FAS.py:174-184); the correction is
 * then prolongated with an MG_OP_PROLONG_ADD op */
int mg_fas_subtract_approximation(mg_workspace* ws, int level);
/* sqrt(sum over interior nodes of (RHS - (L u + gamma e^u u))^2) (.exa4:83-89); synchronises */
int mg_fas_residual_norm(mg_workspace* ws, int level, double gamma, double* out);

/* ---- allocation-free residual norm ------------------------------------------------------------ */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * out_device[0] = sqrt(sum over interior nodes of (rhs - A u)^2) of one scalar fp64 field on an n^dimension
 * node grid -- what mg_residual_norm computes for a workspace level -- without a workspace and without any
 * allocation: u_device / rhs_device are the caller's device arrays (rows of `pitch` elements, n rows per
 * plane, n planes in 3D; e.g. from mg_field_pointer), out_device a device double of the caller, the stencil
 * as for mg_set_operator (offsets[n_entries][dimension], values[n_entries], accumulated in that order).
 * The per-row partial sums live in a static device array of MG_NOALLOC_MAX_ROWS = 266240 doubles (2.03 MiB
 * per GPU that loads the library, until exit): grids with more interior rows (3D: n > 517) are rejected with
 * MG_ERR_INVALID.  Enqueued on `cuda_stream` (NULL: the default stream) of the CURRENT device, no
 * synchronisation.  Every call shares the array: calls that overlap on the GPU (two streams, two threads)
 * race; nothing in it needs clearing between calls.  The first call on a device looks the array's address
 * up (cudaGetSymbolAddress): make it outside a stream capture.  Row sums are formed in a fixed order, so
 * the result is deterministic; it agrees with mg_residual_norm to rounding (<= 1e-12 relative). */
#define MG_NOALLOC_MAX_ROWS 266240
int mg_residual_norm_noalloc(int dimension, int n, int pitch, int n_entries, const int32_t* offsets,
                             const double* values, const void* u_device, const void* rhs_device,
                             double* out_device, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* MGENGINE_H */
