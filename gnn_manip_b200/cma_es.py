"""Host-side (mu/mu_w, lambda)-CMA-ES with an ask/tell interface: the optimiser loop behind
TrajectoryCMAsolver.optimize_trajectory (cma.fmin2, gnn_manip/utils/traj_utils.py:257) and
InterpolatedCMAsolver.optimize_trajectory (cma.fmin_con, traj_utils.py:336).

The reference delegates to the `cma` package (unpinned pip dependency, environment.yml:9; not vendored, not installed
here) and evaluates the population serially.  Here a generation is ONE batched objective call: every candidate of the
population is rolled out as a disjoint graph of one rollout plan (and the population is split over the ranks, see
sharding.py), so the optimiser only ever sees [popsize] objective values.

The update equations are the published ones (Hansen, "The CMA Evolution Strategy: A Tutorial", 2016: default
strategy parameters, rank-one + rank-mu update, cumulative step-size adaptation); pycma's default box handling
(BoundTransform) and its augmented-Lagrangian constraint handler are replaced by the simpler statements documented on
`CMAES.ask` and `AugmentedLagrangian` -- parity with pycma is UNPINNED (library absent) and sample streams differ.
"""
import numpy as np


class CMAES:
    def __init__(self, x0, sigma0, popsize=None, seed=1234, bounds=None, maxiter=100):
        self.mean = np.asarray(x0, dtype=np.float64).copy()
        self.n = n = self.mean.size
        self.sigma = float(sigma0)
        self.lam = int(popsize) if popsize else 4 + int(3 * np.log(n))
        self.maxiter = int(maxiter)
        self.rng = np.random.RandomState(seed)
        self.bounds = None if bounds is None else (float(bounds[0]), float(bounds[1]))
        mu = self.lam // 2
        w = np.log(self.lam / 2.0 + 0.5) - np.log(np.arange(1, mu + 1))
        self.weights = w / w.sum()
        self.mu = mu
        self.mueff = 1.0 / np.sum(self.weights ** 2)
        self.cc = (4 + self.mueff / n) / (n + 4 + 2 * self.mueff / n)
        self.cs = (self.mueff + 2) / (n + self.mueff + 5)
        self.c1 = 2 / ((n + 1.3) ** 2 + self.mueff)
        self.cmu = min(1 - self.c1, 2 * (self.mueff - 2 + 1 / self.mueff) / ((n + 2) ** 2 + self.mueff))
        self.damps = 1 + 2 * max(0.0, np.sqrt((self.mueff - 1) / (n + 1)) - 1) + self.cs
        self.chi_n = np.sqrt(n) * (1 - 1 / (4.0 * n) + 1 / (21.0 * n * n))
        self.pc, self.ps = np.zeros(n), np.zeros(n)
        self.C = np.eye(n)
        self.B, self.D = np.eye(n), np.ones(n)
        self.iteration = 0
        self.evaluations = 0
        self._eig_at = 0
        self.best_x, self.best_f = self.mean.copy(), np.inf
        self._asked = None

    # ------------------------------------------------------------------------------------------------
    def ask(self):
        """[popsize, n] candidates x_k = m + sigma B D z_k.  Box bounds (one [lo, hi] pair for every coordinate, as the
        reference passes them, traj_utils.py:326) are enforced by clipping the sampled point; the unclipped sample is
        what the distribution update sees with a quadratic distance penalty added to its value in tell()."""
        self._update_eig()
        z = self.rng.randn(self.lam, self.n)
        y = (z * self.D) @ self.B.T
        x = self.mean + self.sigma * y
        self._raw = x
        if self.bounds is not None:
            x = np.clip(x, self.bounds[0], self.bounds[1])
        self._asked = x
        return x.copy()

    def tell(self, values):
        f = np.asarray(values, dtype=np.float64).copy()
        if f.shape != (self.lam,):
            raise ValueError(f"tell() needs {self.lam} objective values, got shape {f.shape}")
        x = self._raw
        if self.bounds is not None:
            f = f + 1e3 * np.sum((self._raw - self._asked) ** 2, axis=1)
        self.evaluations += self.lam
        order = np.argsort(f, kind="stable")
        if f[order[0]] < self.best_f:
            self.best_f, self.best_x = float(f[order[0]]), self._asked[order[0]].copy()
        n, sel = self.n, x[order[: self.mu]]
        old = self.mean
        self.mean = self.weights @ sel
        ymean = (self.mean - old) / self.sigma
        inv_sqrt_c_y = self.B @ ((self.B.T @ ymean) / self.D)
        self.ps = (1 - self.cs) * self.ps + np.sqrt(self.cs * (2 - self.cs) * self.mueff) * inv_sqrt_c_y
        self.iteration += 1
        hsig = (np.linalg.norm(self.ps) / np.sqrt(1 - (1 - self.cs) ** (2 * self.iteration)) / self.chi_n
                < 1.4 + 2.0 / (n + 1))
        self.pc = (1 - self.cc) * self.pc + hsig * np.sqrt(self.cc * (2 - self.cc) * self.mueff) * ymean
        ys = (sel - old) / self.sigma
        rank_mu = (ys * self.weights[:, None]).T @ ys
        self.C = ((1 - self.c1 - self.cmu + (1 - hsig) * self.c1 * self.cc * (2 - self.cc)) * self.C
                  + self.c1 * np.outer(self.pc, self.pc) + self.cmu * rank_mu)
        self.sigma *= float(np.exp((self.cs / self.damps) * (np.linalg.norm(self.ps) / self.chi_n - 1)))
        return order

    def _update_eig(self):
        if self.iteration - self._eig_at < max(1, int(1.0 / ((self.c1 + self.cmu) * self.n * 10.0))) and self.iteration:
            return
        self._eig_at = self.iteration
        self.C = np.triu(self.C) + np.triu(self.C, 1).T
        d2, self.B = np.linalg.eigh(self.C)
        self.D = np.sqrt(np.maximum(d2, 1e-30))

    def stop(self):
        return self.iteration >= self.maxiter or self.sigma * float(self.D.max()) < 1e-12

    @property
    def result(self):
        return self.best_x, self.best_f, self.evaluations, self.iteration, self.mean.copy(), self.sigma


class AugmentedLagrangian:
    """Inequality constraints g_i(x) <= 0 folded into the objective (the role of cma.fmin_con's handler,
    traj_utils.py:336-337):  f_al(x) = f(x) + sum_i { lam_i g_i + mu_i g_i^2 / 2  if lam_i + mu_i g_i >= 0
                                                      -lam_i^2 / (2 mu_i)         otherwise },
    multipliers updated once per generation from the constraint values at the best candidate:
    lam_i <- max(0, lam_i + mu_i g_i); mu_i doubled while g_i stays violated, halved (floored) once it is inactive."""

    def __init__(self, n_constraints, mu0=1.0):
        self.lam = np.zeros(n_constraints)
        self.mu = np.full(n_constraints, float(mu0))

    def __call__(self, f, g):
        g = np.asarray(g, dtype=np.float64)
        active = self.lam + self.mu * g >= 0
        terms = np.where(active, self.lam * g + 0.5 * self.mu * g * g, -self.lam ** 2 / (2 * self.mu))
        return np.asarray(f, dtype=np.float64) + terms.sum(axis=-1)

    def update(self, g_best):
        g = np.asarray(g_best, dtype=np.float64)
        self.lam = np.maximum(0.0, self.lam + self.mu * g)
        self.mu = np.where(g > 0, self.mu * 2.0, np.maximum(self.mu * 0.5, 1e-3))


def fmin_batch(objective_batch, x0, sigma0, options=None, constraints_batch=None):
    """Minimise with one objective_batch(X [popsize, n]) -> [popsize] call per generation (every rank of a
    torch.distributed job must call this with the same arguments: the sample stream is seeded identically everywhere).
    options: popsize, seed, maxiter, bounds -- the keys the reference sets on cma.CMAOptions (traj_utils.py:72-76, 326).
    constraints_batch(X) -> [popsize, n_constraints] switches the augmented Lagrangian on (fmin_con).
    Returns (xbest, es) like cma.fmin2."""
    options = dict(options or {})
    es = CMAES(x0, sigma0, popsize=options.get("popsize"), seed=options.get("seed", 1234),
               bounds=options.get("bounds"), maxiter=options.get("maxiter", 100))
    al = None
    es.history = []
    best_feasible = (np.inf, None)
    while not es.stop():
        X = es.ask()
        f = np.asarray(objective_batch(X), dtype=np.float64)
        if constraints_batch is not None:
            g = np.asarray(constraints_batch(X), dtype=np.float64)
            if al is None:
                al = AugmentedLagrangian(g.shape[1])
            order = es.tell(al(f, g))
            al.update(g[order[0]])
            feasible = np.all(g <= 0, axis=1)
            if feasible.any():
                k = int(np.argmin(np.where(feasible, f, np.inf)))
                if f[k] < best_feasible[0]:
                    best_feasible = (float(f[k]), X[k].copy())
        else:
            order = es.tell(f)
        es.history.append(float(f[order[0]]))
    if constraints_batch is not None:
        # like cma.fmin_con: the best feasible candidate seen, else the final distribution mean
        es.best_feasible = best_feasible
        return (best_feasible[1] if best_feasible[1] is not None else es.mean.copy()), es
    return es.best_x, es
