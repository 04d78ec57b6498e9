/*
 * ORACLE / CPU BASELINE (test infrastructure, never on the product path) -- PARITY UNPINNED.
 *
 * C + OpenMP restatement of the op-list interpreter of oracle/mg_oracle.py, one loop nest per
 * unfused statement -- the granularity at which ExaStencils emits C++ for the statements
 * ProgramGenerator.generate_multigrid writes (one statement per line,
 * /root/reference/evostencils/code_generation/exastencils.py:684-925; OpenMP over the loop
 * dimensions, example_problems/lib/parallelization_pureOmp.knowledge:2-9).  The reference's real
 * generated code cannot be produced here (ExaStencils v1.1 needs java + sbt, README.md:21-32), so
 * this is a "port", not the reference.  It exports the same C ABI as include/mgengine.h (scalar
 * real fields, Dirichlet boundaries, CG coarse solver) so that the Python binding can drive it;
 * it is validated bit-for-bit against the numpy oracle in tests/test_oracle_c.py and timed by
 * bench.py's cpu_baseline / --impl reference legs only.
 *
 * Build: gcc -O3 -march=native -fopenmp -ffp-contract=off -shared -fPIC (oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/mgengine.h"

#define MAXE 27
#define MAXS 8
#define LANES 256

typedef struct {
    int n;
    long long total;
    double *sol, *alt, *rhs, *res, *cor, *init;
    double* approx;                 /* FAS: restricted fine solution (allocated on first use) */
    int ne;
    int off[MAXE][3];
    double val[MAXE];
    int set;
} level_t;

typedef struct {
    int level, jacobi, colored, m, point;
    int period[3];
    int pos[MAXS][3];
    int mask_index[1 << MAXS];
    double* w;
} smoother_t;

typedef struct {
    mg_op* ops;
    int n;
} cycle_t;

struct mg_workspace {
    int dim, lo, hi;
    level_t* lv;
    smoother_t* sm;
    int n_sm, cap_sm;
    cycle_t* cy;
    int n_cy, cap_cy;
    int cgs_max;
    double cgs_red;
};

static char g_err[512] = "";
static int fail(int code, const char* msg) {
    snprintf(g_err, sizeof g_err, "%s", msg);
    return code;
}
const char* mg_last_error(void) { return g_err; }

static level_t* L(mg_workspace* ws, int level) { return &ws->lv[level - ws->lo]; }
static inline long long idx_of(const level_t* l, int dim, int x, int y, int z) {
    return dim == 3 ? ((long long)z * l->n + y) * l->n + x : (long long)y * l->n + x;
}

int mg_create(mg_workspace** out, int dimension, int min_level, int max_level, int n_fields, int dtype, int device) {
    (void)device;
    if (n_fields != 1 || dtype != MG_F64) return fail(MG_ERR_INVALID, "CPU oracle port: one real field only");
    if (dimension < 2 || dimension > 3) return fail(MG_ERR_INVALID, "dimension");
    mg_workspace* ws = calloc(1, sizeof *ws);
    ws->dim = dimension; ws->lo = min_level; ws->hi = max_level;
    ws->cgs_max = 1000; ws->cgs_red = 1e-12;
    ws->lv = calloc(max_level - min_level + 1, sizeof(level_t));
    for (int level = min_level; level <= max_level; ++level) {
        level_t* l = L(ws, level);
        l->n = (1 << level) + 1;
        l->total = (long long)l->n * l->n * (dimension == 3 ? l->n : 1);
        size_t b = (size_t)l->total * sizeof(double);
        l->sol = calloc(1, b); l->alt = calloc(1, b); l->rhs = calloc(1, b); l->res = calloc(1, b);
        if (level == max_level) { l->cor = calloc(1, b); l->init = calloc(1, b); }
        if (!l->sol || !l->alt || !l->rhs || !l->res) return fail(MG_ERR_NOMEM, "calloc");
    }
    *out = ws;
    return MG_OK;
}

int mg_clear_smoothers(mg_workspace* ws) {
    for (int i = 0; i < ws->n_sm; ++i) free(ws->sm[i].w);
    ws->n_sm = 0;
    for (int i = 0; i < ws->n_cy; ++i) free(ws->cy[i].ops);
    ws->n_cy = 0;
    return MG_OK;
}
int mg_free_cycles(mg_workspace* ws) {
    for (int i = 0; i < ws->n_cy; ++i) free(ws->cy[i].ops);
    ws->n_cy = 0;
    return MG_OK;
}

int mg_destroy(mg_workspace* ws) {
    if (!ws) return MG_OK;
    mg_clear_smoothers(ws);
    for (int level = ws->lo; level <= ws->hi; ++level) {
        level_t* l = L(ws, level);
        free(l->sol); free(l->alt); free(l->rhs); free(l->res); free(l->cor); free(l->init); free(l->approx);
    }
    free(ws->lv); free(ws->sm); free(ws->cy); free(ws);
    return MG_OK;
}

int mg_launch_count(mg_workspace* ws, int64_t* out) { (void)ws; *out = 0; return MG_OK; }
int mg_synchronize(mg_workspace* ws) { (void)ws; return MG_OK; }
int mg_set_fast_kernels(mg_workspace* ws, int e) { (void)ws; (void)e; return MG_OK; }
int mg_flush_l2(mg_workspace* ws, size_t b) { (void)ws; (void)b; return MG_OK; }
int mg_mark(mg_workspace* ws, int w) { (void)ws; (void)w; return MG_OK; }
int mg_elapsed(mg_workspace* ws, float* ms) { (void)ws; *ms = 0; return MG_OK; }
int mg_upload_device(mg_workspace* ws, int k, int l, int f, const void* p) {
    (void)ws; (void)k; (void)l; (void)f; (void)p;
    return fail(MG_ERR_INVALID, "no device in the CPU oracle port");
}

int mg_set_operator(mg_workspace* ws, int level, int row, int col, int n, const int32_t* offs, const double* vals) {
    if (row || col || n > MAXE) return fail(MG_ERR_INVALID, "operator block");
    level_t* l = L(ws, level);
    l->ne = n;
    for (int k = 0; k < n; ++k) {
        for (int c = 0; c < 3; ++c) l->off[k][c] = c < ws->dim ? offs[k * ws->dim + c] : 0;
        l->val[k] = vals[k];
    }
    l->set = 1;
    return MG_OK;
}
int mg_set_boundary(mg_workspace* ws, int kind, double k) {
    (void)ws; (void)k;
    return kind == MG_BC_DIRICHLET ? MG_OK : fail(MG_ERR_INVALID, "Dirichlet only");
}
int mg_set_coarse_solver(mg_workspace* ws, int kind, int max_its, double red) {
    if (kind != MG_CGS_CG) return fail(MG_ERR_INVALID, "CG only");
    ws->cgs_max = max_its; ws->cgs_red = red;
    return MG_OK;
}

int mg_add_smoother(mg_workspace* ws, int level, int jacobi, int colored, const int32_t* period, int m,
                    const int32_t* slot_field, const int32_t* slot_pos, int n_masks, const uint32_t* masks,
                    const double* w, int* handle) {
    (void)slot_field;
    if (ws->n_sm == ws->cap_sm) {
        ws->cap_sm = ws->cap_sm ? 2 * ws->cap_sm : 16;
        ws->sm = realloc(ws->sm, ws->cap_sm * sizeof(smoother_t));
    }
    smoother_t* s = &ws->sm[ws->n_sm];
    memset(s, 0, sizeof *s);
    s->level = level; s->jacobi = jacobi; s->colored = colored; s->m = m; s->point = 1;
    for (int d = 0; d < 3; ++d) {
        s->period[d] = d < ws->dim ? period[d] : 1;
        if (s->period[d] != 1) s->point = 0;
    }
    for (int i = 0; i < m; ++i)
        for (int d = 0; d < 3; ++d) s->pos[i][d] = d < ws->dim ? slot_pos[i * ws->dim + d] : 0;
    for (int i = 0; i < (1 << MAXS); ++i) s->mask_index[i] = -1;
    for (int k = 0; k < n_masks; ++k) s->mask_index[masks[k]] = k;
    size_t b = (size_t)n_masks * m * m * sizeof(double);
    s->w = malloc(b);
    memcpy(s->w, w, b);
    *handle = ws->n_sm++;
    return MG_OK;
}

static double* field_ptr(mg_workspace* ws, int kind, int level) {
    level_t* l = L(ws, level);
    switch (kind) {
        case MG_SOL: return l->sol;
        case MG_RHS: return l->rhs;
        case MG_RES: return l->res;
        case MG_COR: return level == ws->hi ? l->cor : l->sol;
    }
    return NULL;
}
int mg_upload(mg_workspace* ws, int kind, int level, int field, const void* host) {
    (void)field;
    memcpy(field_ptr(ws, kind, level), host, (size_t)L(ws, level)->total * sizeof(double));
    return MG_OK;
}
int mg_download(mg_workspace* ws, int kind, int level, int field, void* host) {
    (void)field;
    memcpy(host, field_ptr(ws, kind, level), (size_t)L(ws, level)->total * sizeof(double));
    return MG_OK;
}
int mg_save_initial_guess(mg_workspace* ws) {
    level_t* l = L(ws, ws->hi);
    memcpy(l->init, l->sol, (size_t)l->total * sizeof(double));
    return MG_OK;
}
int mg_restore_initial_guess(mg_workspace* ws) {
    level_t* l = L(ws, ws->hi);
    memcpy(l->sol, l->init, (size_t)l->total * sizeof(double));
    return MG_OK;
}

/* ---- stages: one loop nest per statement ------------------------------------------------------ */
static inline double apply_at(const level_t* l, const long long* lin, const double* u, long long i) {
    double acc = l->val[0] * u[i + lin[0]];
    for (int k = 1; k < l->ne; ++k) acc = acc + l->val[k] * u[i + lin[k]];
    return acc;
}
static void linear_offsets(const level_t* l, int dim, long long* lin) {
    for (int k = 0; k < l->ne; ++k)
        lin[k] = l->off[k][0] + (long long)l->off[k][1] * l->n + (dim == 3 ? (long long)l->off[k][2] * l->n * l->n : 0);
}

static void st_residual(mg_workspace* ws, int level) {
    level_t* l = L(ws, level);
    long long lin[MAXE];
    linear_offsets(l, ws->dim, lin);
    const int n = l->n, zs = ws->dim == 3 ? 1 : 0, ze = ws->dim == 3 ? n - 2 : 0;
#pragma omp parallel for collapse(2) schedule(static)
    for (int z = zs; z <= ze; ++z)
        for (int y = 1; y <= n - 2; ++y)
            for (int x = 1; x <= n - 2; ++x) {
                long long i = idx_of(l, ws->dim, x, y, z);
                l->res[i] = l->rhs[i] - apply_at(l, lin, l->sol, i);
            }
}

static int st_smooth(mg_workspace* ws, const smoother_t* s) {
    level_t* l = L(ws, s->level);
    long long lin[MAXE];
    linear_offsets(l, ws->dim, lin);
    const int n = l->n, dim = ws->dim, zs = dim == 3 ? 1 : 0, ze = dim == 3 ? n - 2 : 0;
    if (s->point) {
        const double w = s->w[(size_t)s->mask_index[1] * 1];
        if (!s->colored) {
            memcpy(l->alt, l->sol, (size_t)l->total * sizeof(double));
#pragma omp parallel for collapse(2) schedule(static)
            for (int z = zs; z <= ze; ++z)
                for (int y = 1; y <= n - 2; ++y)
                    for (int x = 1; x <= n - 2; ++x) {
                        long long i = idx_of(l, dim, x, y, z);
                        double r = l->rhs[i] - apply_at(l, lin, l->sol, i);
                        l->alt[i] = l->sol[i] + w * r;
                    }
            double* t = l->sol; l->sol = l->alt; l->alt = t;
        } else {
            for (int k = 0; k < l->ne; ++k) {
                int par = (l->off[k][0] + l->off[k][1] + l->off[k][2]) & 1;
                int centre = !l->off[k][0] && !l->off[k][1] && !l->off[k][2];
                if (!par && !centre) return fail(MG_ERR_INVALID, "same-colour coupling: in-place colouring undefined");
            }
            for (int colour = 0; colour < 2; ++colour) {
#pragma omp parallel for collapse(2) schedule(static)
                for (int z = zs; z <= ze; ++z)
                    for (int y = 1; y <= n - 2; ++y)
                        for (int x = 1 + ((1 + y + z + colour) & 1); x <= n - 2; x += 2) {
                            long long i = idx_of(l, dim, x, y, z);
                            double r = l->rhs[i] - apply_at(l, lin, l->sol, i);
                            l->sol[i] = l->sol[i] + w * r;
                        }
            }
        }
        return MG_OK;
    }
    /* tile smoother (Jacobi) */
    memcpy(l->alt, l->sol, (size_t)l->total * sizeof(double));
    int tiles[3] = {1, 1, 1};
    for (int d = 0; d < dim; ++d) tiles[d] = (n - 1) / s->period[d] + 1;
    const int m = s->m;
#pragma omp parallel for collapse(2) schedule(static)
    for (int tz = 0; tz < tiles[2]; ++tz)
        for (int ty = 0; ty < tiles[1]; ++ty)
            for (int tx = 0; tx < tiles[0]; ++tx) {
                double r[MAXS];
                long long id[MAXS];
                unsigned mask = 0;
                for (int a = 0; a < m; ++a) {
                    int x = tx * s->period[0] + s->pos[a][0], y = ty * s->period[1] + s->pos[a][1];
                    int z = dim == 3 ? tz * s->period[2] + s->pos[a][2] : 0;
                    int in = x >= 1 && x <= n - 2 && y >= 1 && y <= n - 2 && (dim < 3 || (z >= 1 && z <= n - 2));
                    id[a] = -1; r[a] = 0.0;
                    if (in) {
                        id[a] = idx_of(l, dim, x, y, z);
                        mask |= 1u << a;
                        r[a] = l->rhs[id[a]] - apply_at(l, lin, l->sol, id[a]);
                    }
                }
                if (!mask) continue;
                const double* w = s->w + (size_t)s->mask_index[mask] * m * m;
                for (int a = 0; a < m; ++a) {
                    if (!((mask >> a) & 1u)) continue;
                    double delta = 0.0;
                    int first = 1;
                    for (int b = 0; b < m; ++b) {
                        if (!((mask >> b) & 1u)) continue;
                        double term = w[a * m + b] * r[b];
                        delta = first ? term : delta + term;
                        first = 0;
                    }
                    l->alt[id[a]] = l->sol[id[a]] + delta;
                }
            }
    double* t = l->sol; l->sol = l->alt; l->alt = t;
    return MG_OK;
}

static void st_restrict(mg_workspace* ws, int coarse, int src_kind) {
    level_t* lc = L(ws, coarse);
    level_t* lf = L(ws, coarse + 1);
    const double* fine = field_ptr(ws, src_kind, coarse + 1);
    const int n = lc->n, dim = ws->dim, zs = dim == 3 ? 1 : 0, ze = dim == 3 ? n - 2 : 0;
    static const double w1[3] = {0.25, 0.5, 0.25};
#pragma omp parallel for collapse(2) schedule(static)
    for (int Z = zs; Z <= ze; ++Z)
        for (int Y = 1; Y <= n - 2; ++Y)
            for (int X = 1; X <= n - 2; ++X) {
                double acc = 0.0;
                int first = 1;
                for (int oz = dim == 3 ? -1 : 0; oz <= (dim == 3 ? 1 : 0); ++oz)
                    for (int oy = -1; oy <= 1; ++oy)
                        for (int ox = -1; ox <= 1; ++ox) {
                            double w = w1[ox + 1] * w1[oy + 1];
                            if (dim == 3) w = w1[oz + 1] * w;
                            double term = w * fine[idx_of(lf, dim, 2 * X + ox, 2 * Y + oy, 2 * Z + oz)];
                            acc = first ? term : acc + term;
                            first = 0;
                        }
                lc->rhs[idx_of(lc, dim, X, Y, Z)] = acc;
            }
}

static void st_prolong(mg_workspace* ws, int fine, int src_kind, double omega, int add) {
    level_t* lf = L(ws, fine);
    level_t* lc = L(ws, fine - 1);
    const double* coarse = field_ptr(ws, src_kind, fine - 1);
    double* dst = add ? lf->sol : field_ptr(ws, MG_COR, fine);
    const int n = lf->n, dim = ws->dim, zs = dim == 3 ? 1 : 0, ze = dim == 3 ? n - 2 : 0;
#pragma omp parallel for collapse(2) schedule(static)
    for (int z = zs; z <= ze; ++z)
        for (int y = 1; y <= n - 2; ++y)
            for (int x = 1; x <= n - 2; ++x) {
                int px = x & 1, py = y & 1, pz = dim == 3 ? (z & 1) : 0;
                double acc = 0.0;
                int first = 1;
                for (int cz = 0; cz <= pz; ++cz)
                    for (int cy = 0; cy <= py; ++cy)
                        for (int cx = 0; cx <= px; ++cx) {
                            double term = coarse[idx_of(lc, dim, (x >> 1) + cx, (y >> 1) + cy, (z >> 1) + cz)];
                            acc = first ? term : acc + term;
                            first = 0;
                        }
                int k = px + py + pz;
                double pe = k == 0 ? acc : (k == 1 ? 0.5 : (k == 2 ? 0.25 : 0.125)) * acc;
                long long i = idx_of(lf, dim, x, y, z);
                dst[i] = add ? dst[i] + omega * pe : pe;
            }
}

static double butterfly(double* p) {
    double w[8];
    for (int q = 0; q < 8; ++q) {
        double* v = p + 32 * q;
        for (int off = 16; off > 0; off >>= 1)
            for (int i = 0; i < off; ++i) v[i] = v[i] + v[i + off];
        w[q] = v[0];
    }
    for (int k = 4; k > 0; k >>= 1)
        for (int i = 0; i < k; ++i) w[i] = w[i] + w[i + k];
    return w[0];
}
static double lane_dot(const double* a, const double* b, long long n) {
    double part[LANES];
    for (int i = 0; i < LANES; ++i) part[i] = 0.0;
    for (long long e = 0; e < n; ++e) part[e % LANES] = part[e % LANES] + a[e] * b[e];
    return butterfly(part);
}

static void st_cg(mg_workspace* ws, int level) {
    level_t* l = L(ws, level);
    long long lin[MAXE];
    linear_offsets(l, ws->dim, lin);
    const int n = l->n, dim = ws->dim, m = n - 2;
    const long long N = (long long)m * m * (dim == 3 ? m : 1);
    double *x = calloc(N, 8), *r = malloc(N * 8), *p = malloc(N * 8), *ap = malloc(N * 8);
    double* pf = calloc(l->total, 8);
    long long* map = malloc(N * sizeof(long long));
    long long e = 0;
    for (int z = dim == 3 ? 1 : 0; z <= (dim == 3 ? n - 2 : 0); ++z)
        for (int y = 1; y <= n - 2; ++y)
            for (int xx = 1; xx <= n - 2; ++xx) map[e++] = idx_of(l, dim, xx, y, z);
    for (e = 0; e < N; ++e) { r[e] = l->rhs[map[e]]; p[e] = r[e]; }
    double rs = lane_dot(r, r, N);
    double r0 = sqrt(rs);
    if (r0 != 0.0) {
        for (int it = 1; it <= ws->cgs_max; ++it) {
            for (e = 0; e < N; ++e) pf[map[e]] = p[e];
            for (e = 0; e < N; ++e) ap[e] = apply_at(l, lin, pf, map[e]);
            double alpha = rs / lane_dot(p, ap, N);
            for (e = 0; e < N; ++e) { x[e] = x[e] + alpha * p[e]; r[e] = r[e] - alpha * ap[e]; }
            double rs_new = lane_dot(r, r, N);
            if (sqrt(rs_new) <= ws->cgs_red * r0) break;
            double beta = rs_new / rs;
            for (e = 0; e < N; ++e) p[e] = r[e] + beta * p[e];
            rs = rs_new;
        }
    }
    memset(l->sol, 0, (size_t)l->total * 8);
    for (e = 0; e < N; ++e) l->sol[map[e]] = x[e];
    free(x); free(r); free(p); free(ap); free(pf); free(map);
}

static int run_ops(mg_workspace* ws, const mg_op* ops, int n_ops) {
    for (int i = 0; i < n_ops; ++i) {
        const mg_op* op = &ops[i];
        if (op->level < ws->lo || op->level > ws->hi) return fail(MG_ERR_INVALID, "level");
        switch (op->kind) {
            case MG_OP_SMOOTH: {
                int rc = st_smooth(ws, &ws->sm[op->smoother]);
                if (rc) return rc;
                break;
            }
            case MG_OP_RESIDUAL: st_residual(ws, op->level); break;
            case MG_OP_RESTRICT: st_restrict(ws, op->level, op->src); break;
            case MG_OP_RESIDUAL_RESTRICT: st_residual(ws, op->level + 1); st_restrict(ws, op->level, MG_RES); break;
            case MG_OP_ZERO: memset(L(ws, op->level)->sol, 0, (size_t)L(ws, op->level)->total * 8); break;
            case MG_OP_PROLONG_ADD: st_prolong(ws, op->level, op->src, op->omega, 1); break;
            case MG_OP_PROLONG: st_prolong(ws, op->level, op->src, 1.0, 0); break;
            case MG_OP_CGS: st_cg(ws, op->level); break;
            default: return fail(MG_ERR_INVALID, "op kind");
        }
    }
    return MG_OK;
}

int mg_exec_ops(mg_workspace* ws, const mg_op* ops, int n_ops) { return run_ops(ws, ops, n_ops); }

int mg_residual_norm(mg_workspace* ws, int level, double* out) {
    level_t* l = L(ws, level);
    long long lin[MAXE];
    linear_offsets(l, ws->dim, lin);
    const int n = l->n, dim = ws->dim;
    const int rows = dim == 3 ? n * n : n;
    double* rowsum = calloc(rows, 8);
#pragma omp parallel for schedule(static)
    for (int row = 0; row < rows; ++row) {
        int z = dim == 3 ? row / n : 0, y = dim == 3 ? row % n : row;
        if (y < 1 || y > n - 2 || (dim == 3 && (z < 1 || z > n - 2))) continue;
        double acc = 0.0;
        for (int x = 1; x <= n - 2; ++x) {
            long long i = idx_of(l, dim, x, y, z);
            double r = l->rhs[i] - apply_at(l, lin, l->sol, i);
            acc += r * r;
        }
        rowsum[row] = acc;
    }
    double total = 0.0;
    for (int row = 0; row < rows; ++row) total += rowsum[row];
    free(rowsum);
    *out = sqrt(total);
    return MG_OK;
}

int mg_build_cycle(mg_workspace* ws, const mg_op* ops, int n_ops, int* cycle) {
    if (ws->n_cy == ws->cap_cy) {
        ws->cap_cy = ws->cap_cy ? 2 * ws->cap_cy : 8;
        ws->cy = realloc(ws->cy, ws->cap_cy * sizeof(cycle_t));
    }
    ws->cy[ws->n_cy].ops = malloc(n_ops * sizeof(mg_op));
    memcpy(ws->cy[ws->n_cy].ops, ops, n_ops * sizeof(mg_op));
    ws->cy[ws->n_cy].n = n_ops;
    *cycle = ws->n_cy++;
    return MG_OK;
}

/* OpenMP team size, set explicitly by the caller (torchrun exports OMP_NUM_THREADS=1 to its workers, which
 * would silently time a one-thread run); the getter reports what a parallel region really gets. */
#ifdef _OPENMP
#include <omp.h>
#endif
int mg_oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
    return MG_OK;
}
int mg_oracle_threads(void) {
    int used = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp single
        used = omp_get_num_threads();
    }
#endif
    return used;
}

static double now_ms(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return 1e3 * t.tv_sec + 1e-6 * t.tv_nsec;
}

int mg_run_cycles(mg_workspace* ws, int cycle, int n_cycles, float* ms) {
    double t0 = now_ms();
    for (int i = 0; i < n_cycles; ++i) {
        int rc = run_ops(ws, ws->cy[cycle].ops, ws->cy[cycle].n);
        if (rc) return rc;
        double r;
        mg_residual_norm(ws, ws->hi, &r);
    }
    if (ms) *ms = (float)(now_ms() - t0);
    return MG_OK;
}

int mg_time_ops(mg_workspace* ws, const mg_op* ops, int n_ops, int reps, float* ms) {
    double t0 = now_ms();
    for (int i = 0; i < reps; ++i) {
        int rc = run_ops(ws, ops, n_ops);
        if (rc) return rc;
    }
    *ms = (float)(now_ms() - t0);
    return MG_OK;
}

int mg_solve(mg_workspace* ws, int cycle, double tol, int max_its, double* hist, int* n_its, int* converged,
             float* ms) {
    double t0 = now_ms();
    mg_residual_norm(ws, ws->hi, &hist[0]);
    *n_its = 0; *converged = 0;
    for (int it = 1; it <= max_its; ++it) {
        int rc = run_ops(ws, ws->cy[cycle].ops, ws->cy[cycle].n);
        if (rc) return rc;
        mg_residual_norm(ws, ws->hi, &hist[it]);
        *n_its = it;
        if (!isfinite(hist[it])) break;
        if (hist[it] < tol * hist[0]) { *converged = 1; break; }
    }
    if (ms) *ms = (float)(now_ms() - t0);
    return MG_OK;
}


/* ---- nonlinear FAS stages (example_problems/FAS_2D_Basic/FAS_2D_Basic.exa4; oracle/fas_oracle.py is the
 * numpy statement of the same arithmetic): -Laplace(u) + gam u e^u = f, 2D, one loop nest per statement ---- */
typedef struct { double cc, cx, cy, gam; } fas_star_t;

static int fas_prepare(mg_workspace* ws, int level, double gamma, level_t** out, fas_star_t* s) {
    if (ws->dim != 2) return fail(MG_ERR_INVALID, "FAS stages: 2D only");
    if (level < ws->lo || level > ws->hi) return fail(MG_ERR_INVALID, "level out of range");
    level_t* l = L(ws, level);
    if (!l->set || l->ne != 5) return fail(MG_ERR_INVALID, "FAS stages need a 5-point star operator");
    s->gam = gamma; s->cc = s->cx = s->cy = 0.0;
    for (int k = 0; k < 5; ++k) {
        if (l->off[k][0] == 0 && l->off[k][1] == 0) s->cc = l->val[k];
        else if (l->off[k][1] == 0) s->cx = l->val[k];
        else s->cy = l->val[k];
    }
    if (!l->approx) l->approx = calloc(1, (size_t)l->total * sizeof(double));
    *out = l;
    return MG_OK;
}

/* (L u + gam e^u u): centre, E, W, N, S in the order the .exa4 declares the stencil (:18-24) */
static inline double fas_apply(const fas_star_t* s, const double* u, long long i, int n) {
    const double c = u[i];
    double a = s->cc * c;
    a = a + s->cx * u[i + 1];
    a = a + s->cx * u[i - 1];
    a = a + s->cy * u[i + n];
    a = a + s->cy * u[i - n];
    return a + (s->gam * exp(c)) * c;
}

static inline double fas_relax(const fas_star_t* s, const double* u, const double* f, long long i, int n, double w,
                               int newton_steps) {
    const double e = s->cx * u[i + 1], wt = s->cx * u[i - 1], nt = s->cy * u[i + n], st = s->cy * u[i - n];
    double v = u[i];
    const int steps = newton_steps > 0 ? newton_steps : 1;
    for (int k = 0; k < steps; ++k) {
        double a = s->cc * v;
        a = a + e; a = a + wt; a = a + nt; a = a + st;
        const double ev = exp(v);
        const double num = f[i] - (a + (s->gam * ev) * v);
        const double den = newton_steps > 0 ? s->cc + s->gam * (v + 1.0) * ev : s->cc;
        v = v + w * (num / den);
    }
    return v;
}

int mg_fas_smooth(mg_workspace* ws, int level, double omega, double gamma, int sweeps, int newton_steps, int colored) {
    level_t* l; fas_star_t s;
    int rc = fas_prepare(ws, level, gamma, &l, &s);
    if (rc) return rc;
    const int n = l->n;
    for (int k = 0; k < sweeps; ++k) {
        if (colored) {
            for (int colour = 0; colour < 2; ++colour) {
#pragma omp parallel for schedule(static)
                for (int y = 1; y <= n - 2; ++y)
                    for (int x = 1 + ((1 + y + colour) & 1); x <= n - 2; x += 2) {
                        const long long i = (long long)y * n + x;
                        l->sol[i] = fas_relax(&s, l->sol, l->rhs, i, n, omega, newton_steps);
                    }
            }
        } else {
            memcpy(l->alt, l->sol, (size_t)l->total * sizeof(double));       /* boundary values */
#pragma omp parallel for schedule(static)
            for (int y = 1; y <= n - 2; ++y)
                for (int x = 1; x <= n - 2; ++x) {
                    const long long i = (long long)y * n + x;
                    l->alt[i] = fas_relax(&s, l->sol, l->rhs, i, n, omega, newton_steps);
                }
            double* t = l->sol; l->sol = l->alt; l->alt = t;
        }
    }
    return MG_OK;
}

int mg_fas_residual(mg_workspace* ws, int level, double gamma) {
    level_t* l; fas_star_t s;
    int rc = fas_prepare(ws, level, gamma, &l, &s);
    if (rc) return rc;
    const int n = l->n;
    memset(l->res, 0, (size_t)l->total * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int y = 1; y <= n - 2; ++y)
        for (int x = 1; x <= n - 2; ++x) {
            const long long i = (long long)y * n + x;
            l->res[i] = l->rhs[i] - fas_apply(&s, l->sol, i, n);
        }
    return MG_OK;
}

static inline double fas_restrict_at(const double* fine, int nf, int X, int Y) {
    static const double w1[3] = {0.25, 0.5, 0.25};
    double acc = 0.0;
    int first = 1;
    for (int oy = -1; oy <= 1; ++oy)
        for (int ox = -1; ox <= 1; ++ox) {
            const double term = (w1[ox + 1] * w1[oy + 1]) * fine[(long long)(2 * Y + oy) * nf + 2 * X + ox];
            acc = first ? term : acc + term;
            first = 0;
        }
    return acc;
}

int mg_fas_restrict(mg_workspace* ws, int level, double gamma) {
    level_t *lc, *lf; fas_star_t sc, sf;
    int rc = fas_prepare(ws, level, gamma, &lc, &sc);
    if (!rc) rc = fas_prepare(ws, level + 1, gamma, &lf, &sf);
    if (rc) return rc;
    const int n = lc->n;
    memset(lc->approx, 0, (size_t)lc->total * sizeof(double));
    memset(lc->sol, 0, (size_t)lc->total * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int Y = 1; Y <= n - 2; ++Y)
        for (int X = 1; X <= n - 2; ++X) {
            const long long i = (long long)Y * n + X;
            lc->approx[i] = fas_restrict_at(lf->sol, lf->n, X, Y);
            lc->sol[i] = lc->approx[i];
        }
#pragma omp parallel for schedule(static)
    for (int Y = 1; Y <= n - 2; ++Y)
        for (int X = 1; X <= n - 2; ++X) {
            const long long i = (long long)Y * n + X;
            lc->rhs[i] = fas_restrict_at(lf->res, lf->n, X, Y) + fas_apply(&sc, lc->approx, i, n);
        }
    return MG_OK;
}

int mg_fas_subtract_approximation(mg_workspace* ws, int level) {
    level_t* l; fas_star_t s;
    int rc = fas_prepare(ws, level, 0.0, &l, &s);
    if (rc) return rc;
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < l->total; ++i) l->sol[i] = l->sol[i] - l->approx[i];
    return MG_OK;
}

int mg_fas_residual_norm(mg_workspace* ws, int level, double gamma, double* out) {
    level_t* l; fas_star_t s;
    int rc = fas_prepare(ws, level, gamma, &l, &s);
    if (rc) return rc;
    const int n = l->n;
    double* rows = calloc((size_t)n, sizeof(double));
#pragma omp parallel for schedule(static)
    for (int y = 1; y <= n - 2; ++y) {
        double acc = 0.0;
        for (int x = 1; x <= n - 2; ++x) {
            const long long i = (long long)y * n + x;
            const double r = l->rhs[i] - fas_apply(&s, l->sol, i, n);
            acc = acc + r * r;
        }
        rows[y] = acc;
    }
    double sum = 0.0;
    for (int y = 1; y <= n - 2; ++y) sum = sum + rows[y];
    free(rows);
    *out = sqrt(sum);
    return MG_OK;
}
