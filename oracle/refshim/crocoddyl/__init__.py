"""Minimal stand-in for the parts of crocoddyl (3.0 Python API) that the reference's
NumPy solvers and LQ test problems touch.  TEST INFRASTRUCTURE ONLY -- see
oracle/refshim/README.md.  Nothing here is on the hot path: these classes only
produce the per-knot INPUTS (Fx, Fu, Lxx, ... g, Gx, Gu) that the reference's own
python/sqp.py, python/csqp.py and python/qp_solvers/stagewise_qp.py consume.
"""
import numpy as np


class StateVector:
    """crocoddyl::StateVector: x in R^nx, diff(a,b) = b - a, integrate(x,dx) = x + dx."""

    def __init__(self, nx):
        self.nx = nx
        self.ndx = nx
        self.nq = nx // 2
        self.nv = nx - self.nq

    def zero(self):
        return np.zeros(self.nx)

    def diff(self, x0, x1):
        return np.asarray(x1, dtype=float) - np.asarray(x0, dtype=float)

    def integrate(self, x, dx):
        return np.asarray(x, dtype=float) + np.asarray(dx, dtype=float)


class _Data:
    pass


class ActionDataAbstract(_Data):
    def __init__(self, model):
        ndx, nu, ng = model.state.ndx, model.nu, model.ng
        self.cost = 0.0
        self.xnext = model.state.zero()
        self.Fx = np.zeros((ndx, ndx))
        self.Fu = np.zeros((ndx, nu))
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))
        self.g = np.zeros(ng)
        self.Gx = np.zeros((ng, ndx))
        self.Gu = np.zeros((ng, nu))


class ActionModelAbstract:
    def __init__(self, state, nu, nr=0, ng=0, nh=0):
        self.state = state
        self.nu = nu
        self.nr = nr
        self.ng = ng
        self.nh = nh
        self.g_lb = np.full(ng, -np.inf)
        self.g_ub = np.full(ng, np.inf)

    def createData(self):
        return ActionDataAbstract(self)


class DifferentialActionDataAbstract(_Data):
    def __init__(self, model):
        ndx, nv, nu, ng = model.state.ndx, model.state.nv, model.nu, model.ng
        self.cost = 0.0
        self.xout = np.zeros(nv)
        self.Fx = np.zeros((nv, ndx))
        self.Fu = np.zeros((nv, nu))
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))
        self.g = np.zeros(ng)
        self.Gx = np.zeros((ng, ndx))
        self.Gu = np.zeros((ng, nu))


class DifferentialActionModelAbstract:
    def __init__(self, state, nu, nr=0, ng=0, nh=0):
        self.state = state
        self.nu = nu
        self.nr = nr
        self.ng = ng
        self.nh = nh
        self.g_lb = np.full(ng, -np.inf)
        self.g_ub = np.full(ng, np.inf)

    def createData(self):
        return DifferentialActionDataAbstract(self)


class IntegratedActionDataEuler(ActionDataAbstract):
    def __init__(self, model):
        super().__init__(model)
        self.differential = model.differential.createData()
        self.dx = np.zeros(model.state.ndx)


class IntegratedActionModelEuler(ActionModelAbstract):
    """Symplectic-Euler discretisation of a differential action model, restated from
    crocoddyl 3.0: cost and its derivatives scaled by dt,
    dx = [v dt + a dt^2 ; a dt], Fx = I + [[dt^2 a_q, dt I + dt^2 a_v],[dt a_q, dt a_v]],
    Fu = [dt^2 a_u ; dt a_u]; g, Gx, Gu unscaled.  Terminal use (calc(x) only): unscaled
    differential cost, xnext = x."""

    def __init__(self, differential, dt=1e-3, with_cost_residual=True):
        super().__init__(differential.state, differential.nu, differential.nr,
                         differential.ng, differential.nh)
        self.differential = differential
        self.dt = float(dt)
        self.enable_integration = self.dt != 0.0

    @property
    def g_lb(self):
        return self.differential.g_lb

    @g_lb.setter
    def g_lb(self, v):
        pass

    @property
    def g_ub(self):
        return self.differential.g_ub

    @g_ub.setter
    def g_ub(self, v):
        pass

    def createData(self):
        return IntegratedActionDataEuler(self)

    def calc(self, data, x, u=None):
        dd = data.differential
        nv = self.state.nv
        x = np.asarray(x, dtype=float)
        if u is None:
            self.differential.calc(dd, x)
            data.dx = np.zeros(self.state.ndx)
            data.xnext = x.copy()
            data.cost = dd.cost
        else:
            self.differential.calc(dd, x, u)
            if self.enable_integration:
                a = np.asarray(dd.xout, dtype=float)
                v = x[-nv:]
                dt = self.dt
                data.dx = np.concatenate([v * dt + a * dt * dt, a * dt])
                data.xnext = self.state.integrate(x, data.dx)
                data.cost = dt * dd.cost
            else:
                data.dx = np.zeros(self.state.ndx)
                data.xnext = x.copy()
                data.cost = dd.cost
        if self.ng:
            data.g = np.array(dd.g, dtype=float)

    def calcDiff(self, data, x, u=None):
        dd = data.differential
        nv, ndx = self.state.nv, self.state.ndx
        if u is None:
            self.differential.calcDiff(dd, x)
            data.Fx = np.eye(ndx)
            data.Fu = np.zeros((ndx, self.nu))
            data.Lx = np.array(dd.Lx, dtype=float)
            data.Lxx = np.array(dd.Lxx, dtype=float)
            if self.ng:
                data.Gx = np.array(dd.Gx, dtype=float)
            return
        self.differential.calcDiff(dd, x, u)
        if self.enable_integration:
            dt = self.dt
            da_dx = np.asarray(dd.Fx, dtype=float)
            da_du = np.asarray(dd.Fu, dtype=float).reshape(nv, self.nu)
            Fx = np.zeros((ndx, ndx))
            Fx[:nv, :] = da_dx * dt * dt
            Fx[nv:, :] = da_dx * dt
            Fx[:nv, ndx - nv:] += dt * np.eye(nv)
            data.Fx = Fx + np.eye(ndx)
            data.Fu = np.concatenate([da_du * dt * dt, da_du * dt], axis=0)
            data.Lx = dt * np.asarray(dd.Lx, dtype=float)
            data.Lu = dt * np.asarray(dd.Lu, dtype=float)
            data.Lxx = dt * np.asarray(dd.Lxx, dtype=float)
            data.Lxu = dt * np.asarray(dd.Lxu, dtype=float)
            data.Luu = dt * np.asarray(dd.Luu, dtype=float)
        else:
            data.Fx = np.eye(ndx)
            data.Fu = np.zeros((ndx, self.nu))
            data.Lx = np.array(dd.Lx, dtype=float)
            data.Lu = np.array(dd.Lu, dtype=float)
            data.Lxx = np.array(dd.Lxx, dtype=float)
            data.Lxu = np.array(dd.Lxu, dtype=float)
            data.Luu = np.array(dd.Luu, dtype=float)
        if self.ng:
            data.Gx = np.array(dd.Gx, dtype=float)
            data.Gu = np.array(dd.Gu, dtype=float)


class FrozenActionModel(ActionModelAbstract):
    """OUR ADDITION (not in crocoddyl): a knot whose calc/calcDiff return fixed,
    pre-drawn blocks.  Lets synthetic problems of any (nx, nu, ng) be pushed through the
    reference's NumPy solvers.  `blocks` keys: Fx Fu Lx Lu Lxx Lxu Luu g Gx Gu cost xnext."""

    def __init__(self, nx, nu, ng, blocks, g_lb=None, g_ub=None):
        super().__init__(StateVector(nx), nu, 0, ng, 0)
        self.blocks = blocks
        if ng:
            self.g_lb = np.array(g_lb, dtype=float)
            self.g_ub = np.array(g_ub, dtype=float)

    def calc(self, data, x, u=None):
        b = self.blocks
        data.cost = float(b.get("cost", 0.0))
        data.xnext = np.array(b["xnext"], dtype=float) if "xnext" in b else np.array(x, dtype=float)
        if self.ng:
            data.g = np.array(b["g"], dtype=float)

    def calcDiff(self, data, x, u=None):
        b = self.blocks
        for k in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu", "Gx", "Gu"):
            if k in b:
                setattr(data, k, np.array(b[k], dtype=float))


class ShootingProblem:
    def __init__(self, x0, runningModels, terminalModel):
        self.x0 = np.array(x0, dtype=float)
        self.runningModels = list(runningModels)
        self.terminalModel = terminalModel
        self.T = len(self.runningModels)
        self.runningDatas = [m.createData() for m in self.runningModels]
        self.terminalData = terminalModel.createData()
        self.nthreads = 1

    def calc(self, xs, us):
        c = 0.0
        for m, d, x, u in zip(self.runningModels, self.runningDatas, xs, us):
            m.calc(d, x, u)
            c += d.cost
        self.terminalModel.calc(self.terminalData, xs[-1])
        return c + self.terminalData.cost

    def calcDiff(self, xs, us):
        c = 0.0
        for m, d, x, u in zip(self.runningModels, self.runningDatas, xs, us):
            m.calcDiff(d, x, u)
            c += d.cost
        self.terminalModel.calcDiff(self.terminalData, xs[-1])
        return c + self.terminalData.cost


class SolverAbstract:
    """crocoddyl::SolverAbstract surface used by the reference prototypes:
    problem, xs, us, setCandidate."""

    def __init__(self, problem):
        self.problem = problem
        self.xs = [m.state.zero() for m in problem.runningModels] + [problem.terminalModel.state.zero()]
        self.us = [np.zeros(m.nu) for m in problem.runningModels]
        self.is_feasible = False
        self.iter = 0
        self.cost = 0.0

    def setCandidate(self, xs=None, us=None, is_feasible=False):
        if xs is not None and len(xs):
            self.xs = [np.array(x, dtype=float) for x in xs]
        if us is not None and len(us):
            self.us = [np.array(u, dtype=float) for u in us]
        self.is_feasible = is_feasible
