"""TEST INFRASTRUCTURE -- numpy restatement of the reference's nonlinear FAS example
(/root/reference/example_problems/FAS_2D_Basic/FAS_2D_Basic.exa4), SURVEY.md 8f rank 3.

PARITY UNPINNED: the arithmetic of the reference lives in ExaStencils-generated C++ (not buildable
here); this file follows the hand-written / generated Layer-4 text statement by
statement:

  * problem  -Laplace(u) + gam * u * exp(u) = f on the unit square, u = 0 on the boundary, gam = 10,
    right-hand side and exact solution of the "enhanced test problem"           (.exa4:18-33, 46-56)
  * smoother: damped Jacobi-Newton, Solution<next> = u + w (f - (L u + gam e^u u)) / (diag(L) +
    gam (1 + u) e^u), w = 0.8                                                   (.exa4:68-75)
  * residual f - (L u + gam e^u u), boundary 0                                   (.exa4:77-82, 214-218)
  * FAS coarse problem: approximation = R u_fine, coarse solution starts from it, coarse
    rhs = R r_fine + (L_c approx + gam e^approx approx); after the coarse solve the solution
    becomes the correction (u_c -= approx) and is prolongated onto the fine solution
                                                                               (.exa4:219-229, 246-253)
    (the generated text restricts `Solution@coarse` with the fine level's restriction stencil,
    exastencils_FAS.py:120-128 -- read here as the restriction of the fine solution, which is what
    the full approximation scheme needs)
  * coarse-grid solver = 30 smoother sweeps                                     (.exa4:58-63)
  * outer loop: until res < 1e-10 res0 or 100 iterations, norm = sqrt(sum r^2)  (.exa4:83-89, 147-148)

Stencil entries are accumulated in the order the .exa4 declares them (centre, E, W, N, S) with
separate multiply and add; restriction / prolongation weights as in oracle/mg_oracle.py."""
import numpy as np

GAMMA = 10.0
SMOOTHER_WEIGHT = 0.8
CGS_SWEEPS = 30


def nodes(level):
    return 2 ** level + 1


def exact_solution(level):
    n = nodes(level)
    x = np.linspace(0.0, 1.0, n)[None, :]
    y = np.linspace(0.0, 1.0, n)[:, None]
    return (x ** 2 - x ** 3) * np.sin(3.0 * np.pi * y)


def right_hand_side(level, gamma=GAMMA):
    """rhsFct of the enhanced test problem (.exa4:46-53); arrays are [y, x]."""
    n = nodes(level)
    x = np.linspace(0.0, 1.0, n)[None, :]
    y = np.linspace(0.0, 1.0, n)[:, None]
    sol = (x ** 2 - x ** 3) * np.sin(3.0 * np.pi * y)
    return ((9.0 * np.pi ** 2 + gamma * np.exp(sol)) * (x ** 2 - x ** 3) + 6.0 * x - 2.0) * np.sin(3.0 * np.pi * y)


def apply_operator(u, level, gamma):
    """(L u + gam e^u u) on the interior nodes; 0 elsewhere."""
    h2 = float(4 ** level)                      # 1 / h^2
    out = np.zeros_like(u)
    c = u[1:-1, 1:-1]
    a = (2.0 * h2 + 2.0 * h2) * c
    a = a + (-h2) * u[1:-1, 2:]
    a = a + (-h2) * u[1:-1, :-2]
    a = a + (-h2) * u[2:, 1:-1]
    a = a + (-h2) * u[:-2, 1:-1]
    out[1:-1, 1:-1] = a + (gamma * np.exp(c)) * c
    return out


def residual(u, f, level, gamma):
    r = np.zeros_like(u)
    r[1:-1, 1:-1] = f[1:-1, 1:-1] - apply_operator(u, level, gamma)[1:-1, 1:-1]
    return r


def _local_steps(c, nb, f_int, h2, gamma, weight, newton_steps):
    """`newton_steps` in-place updates of the centre values c with the neighbour sum terms fixed
    (exastencils_FAS.py:206-231): c += w (f - (L u + gam e^c c)) / (diag [+ gam (c + 1) e^c]);
    newton_steps = 0: Jacobi-Picard (one step, denominator diag(L) only)."""
    diag = 2.0 * h2 + 2.0 * h2
    for _ in range(max(1, newton_steps)):
        a = diag * c
        for term in nb:
            a = a + term
        num = f_int - (a + (gamma * np.exp(c)) * c)
        den = diag + gamma * (c + 1.0) * np.exp(c) if newton_steps > 0 else diag
        c = c + weight * (num / den)
    return c


def smooth(u, f, level, gamma, weight=SMOOTHER_WEIGHT, newton_steps=1, colored=False):
    """One sweep of the smoother `traverse_graph` emits (exastencils_FAS.py:186-246): Jacobi (the old
    values of the neighbours, `Solution<next>`) or red-black in place, colour (i0 + i1) % 2 == 0 first."""
    h2 = float(4 ** level)

    def neighbours(v):
        return [(-h2) * v[1:-1, 2:], (-h2) * v[1:-1, :-2], (-h2) * v[2:, 1:-1], (-h2) * v[:-2, 1:-1]]

    if not colored:
        new = u.copy()
        new[1:-1, 1:-1] = _local_steps(u[1:-1, 1:-1], neighbours(u), f[1:-1, 1:-1], h2, gamma, weight, newton_steps)
        return new
    new = u.copy()
    n = u.shape[0]
    yy, xx = np.meshgrid(np.arange(1, n - 1), np.arange(1, n - 1), indexing="ij")
    for colour in (0, 1):
        mask = ((xx + yy) % 2) == colour
        upd = _local_steps(new[1:-1, 1:-1], neighbours(new), f[1:-1, 1:-1], h2, gamma, weight, newton_steps)
        new[1:-1, 1:-1] = np.where(mask, upd, new[1:-1, 1:-1])
    return new


def restrict(fine):
    """Full weighting in memory order of the offsets (oracle/mg_oracle.py conventions); interior only."""
    n = (fine.shape[0] - 1) // 2 + 1
    out = np.zeros((n, n))
    w1 = (0.25, 0.5, 0.25)
    acc = None
    for oy in (-1, 0, 1):
        for ox in (-1, 0, 1):
            term = (w1[ox + 1] * w1[oy + 1]) * fine[2 + oy:fine.shape[0] - 2 + oy + 1:2, 2 + ox:fine.shape[1] - 2 + ox + 1:2]
            acc = term if acc is None else acc + term
    out[1:-1, 1:-1] = acc
    return out


def prolong_add(u, coarse, omega=1.0):
    """u += omega * (P coarse) on the interior fine nodes: P = 2^-k * (sum of the enclosing coarse nodes in
    memory order)."""
    n = u.shape[0]
    out = u.copy()
    e = coarse
    out[2:n - 1:2, 2:n - 1:2] += omega * e[1:-1, 1:-1]                                 # even, even
    out[2:n - 1:2, 1:n - 1:2] += omega * (0.5 * (e[1:-1, :-1] + e[1:-1, 1:]))          # odd x
    out[1:n - 1:2, 2:n - 1:2] += omega * (0.5 * (e[:-1, 1:-1] + e[1:, 1:-1]))          # odd y
    out[1:n - 1:2, 1:n - 1:2] += omega * (0.25 * (((e[:-1, :-1] + e[:-1, 1:]) + e[1:, :-1]) + e[1:, 1:]))
    return out


def norm(r):
    return float(np.sqrt(np.sum(r[1:-1, 1:-1] * r[1:-1, 1:-1])))


def cycle(u, f, level, min_level, pre, post, gamma=GAMMA, weight=SMOOTHER_WEIGHT, cgs_sweeps=CGS_SWEEPS,
          cgc_omega=1.0):
    """One FAS V(pre, post) cycle (gen_mgCycle of the .exa4, any depth)."""
    if level == min_level:
        for _ in range(cgs_sweeps):
            u = smooth(u, f, level, gamma, weight)
        return u
    for _ in range(pre):
        u = smooth(u, f, level, gamma, weight)
    r = residual(u, f, level, gamma)
    approx = restrict(u)
    f_c = restrict(r)
    f_c[1:-1, 1:-1] = f_c[1:-1, 1:-1] + apply_operator(approx, level - 1, gamma)[1:-1, 1:-1]
    u_c = cycle(approx.copy(), f_c, level - 1, min_level, pre, post, gamma, weight, cgs_sweeps, cgc_omega)
    u_c = u_c - approx
    u = prolong_add(u, u_c, cgc_omega)
    for _ in range(post):
        u = smooth(u, f, level, gamma, weight)
    return u


def solve(level, min_level, pre=2, post=2, gamma=GAMMA, weight=SMOOTHER_WEIGHT, cgs_sweeps=CGS_SWEEPS,
          target=1e-10, max_iterations=100, cgc_omega=1.0):
    """-> (solution, residual history); the Solve function of the .exa4 (:118-187)."""
    f = right_hand_side(level, gamma)
    u = np.zeros_like(f)
    history = [norm(residual(u, f, level, gamma))]
    while not (history[-1] < target * history[0] or len(history) - 1 >= max_iterations):
        u = cycle(u, f, level, min_level, pre, post, gamma, weight, cgs_sweeps, cgc_omega)
        history.append(norm(residual(u, f, level, gamma)))
    return u, history


def solve_ops(ops, max_level, min_level, gamma=GAMMA, target=1e-10, max_iterations=100):
    """Execute a statement list (`evostencils_b200.fas.FasOp` records: kind, level, omega, sweeps,
    newton_steps, colored) as the cycle of the `.exa4`'s Solve loop.  -> (solution, residual history)."""
    f = {max_level: right_hand_side(max_level, gamma)}
    u = {max_level: np.zeros_like(f[max_level])}
    r, approx = {}, {}

    def run_cycle():
        for op in ops:
            lv = op.level
            if op.kind == "smooth":
                for _ in range(op.sweeps):
                    u[lv] = smooth(u[lv], f[lv], lv, gamma, op.omega, op.newton_steps, op.colored)
            elif op.kind == "residual":
                r[lv] = residual(u[lv], f[lv], lv, gamma)
            elif op.kind == "restrict":
                approx[lv] = restrict(u[lv + 1])
                u[lv] = approx[lv].copy()
                f[lv] = restrict(r[lv + 1])
                f[lv][1:-1, 1:-1] = f[lv][1:-1, 1:-1] + apply_operator(approx[lv], lv, gamma)[1:-1, 1:-1]
            elif op.kind == "subtract":
                u[lv] = u[lv] - approx[lv]
            elif op.kind == "prolong_add":
                u[lv] = prolong_add(u[lv], u[lv - 1], op.omega)
            else:
                raise ValueError(op.kind)

    history = [norm(residual(u[max_level], f[max_level], max_level, gamma))]
    while not (history[-1] < target * history[0] or len(history) - 1 >= max_iterations):
        run_cycle()
        history.append(norm(residual(u[max_level], f[max_level], max_level, gamma)))
        if not np.isfinite(history[-1]):
            break
    return u[max_level], history
