"""TEST INFRASTRUCTURE ONLY -- CPU model of the coefficient-space L-BFGS that
``styletransfer_b200/csrc/optim.cu`` runs on the device (``lbfgs.DeviceLBFGS``).

The reference optimises the Gatys image with ``torch.optim.LBFGS`` at its defaults
(reference network.py:435 via get_content_optimizer :403-409: lr 1, max_iter 20,
history 100, no line search).  torch's ``step`` walks the history twice with one
dot + one axpy kernel per pair (the two-loop recursion).  The device version
keeps the small matrices SY[i][j] = s_i.y_j and YY[i][j] = y_i.y_j and runs the
two loops on COEFFICIENTS: every vector the recursion touches is a combination of
the current gradient and the stored pairs, so one pass over the history (dots of
the new gradient with every s_j, y_j, the previous gradient and the previous
direction) gives every inner product the recursion and the next history update
need, and one more pass forms the direction.  This file restates that algorithm
with numpy float64 scalars and float32 vectors, structured like the kernels
(physical ring slots), and ``tests/test_cpu.py`` pins it against
``torch.optim.LBFGS`` itself.  Nothing in the product imports it.
"""
import numpy as np


class LBFGSModel:
    def __init__(self, x, lr=1.0, max_iter=20, history_size=100, tolerance_grad=1e-7, tolerance_change=1e-9):
        self.x = x                      # float32 numpy vector, updated in place
        n, m = x.size, history_size
        self.m, self.lr, self.max_iter = m, lr, max_iter
        self.tol_grad, self.tol_change = tolerance_grad, tolerance_change
        self.S = np.zeros((m, n), np.float32)
        self.Y = np.zeros((m, n), np.float32)
        self.g_prev = np.zeros(n, np.float32)
        self.d = np.zeros(n, np.float32)
        self.SY = np.zeros((m, m))      # SY[i][j] = s_i . y_j
        self.YY = np.zeros((m, m))
        self.ro = np.zeros(m)
        self.cy, self.cs, self.cg = np.zeros(m), np.zeros(m), 0.0   # d = cg*g_prev + sum cy_j y_j + sum cs_j s_j
        self.sg_prev, self.yg_prev = np.zeros(m), np.zeros(m)
        self.k = self.head = 0
        self.n_iter = 0
        self.t = 0.0
        self.H = 1.0
        self.dg_prev = 0.0              # d . g_prev
        self.gg_prev = 0.0
        self.loss_prev = 0.0
        self.dmax = 0.0

    def _order(self):
        return [(self.head + i) % self.m for i in range(self.k)]      # oldest -> newest

    def iteration(self, g, loss, first_of_step):
        """One (dots, solve, update) round on the gradient ``g`` of the current ``x``.
        Returns False when torch's ``step`` would stop iterating (the rest of the step is skipped)."""
        m = self.m
        g64 = g.astype(np.float64)
        # ---- pass A: dots of g with the history, g_prev, d, itself
        sg = self.S.astype(np.float64) @ g64
        yg = self.Y.astype(np.float64) @ g64
        gg, ggp, dg = g64 @ g64, g64 @ self.g_prev.astype(np.float64), g64 @ self.d.astype(np.float64)
        gabs_sum, gmax = np.abs(g64).sum(), np.abs(g64).max() if g64.size else 0.0
        # ---- solve
        if gmax <= self.tol_grad:
            return False
        if not first_of_step:
            if self.dmax <= self.tol_change or abs(loss - self.loss_prev) < self.tol_change:
                return False
        self.n_iter += 1
        pushed, new = False, -1
        if self.n_iter == 1:
            self.cg, self.cy[:], self.cs[:] = -1.0, 0.0, 0.0
            self.H = 1.0
        else:
            ys = self.t * (dg - self.dg_prev)
            yy = gg - 2.0 * ggp + self.gg_prev
            if ys > 1e-10:
                old = self._order()
                # s_new . y_j = t * d . y_j with d in coefficient form (uses rows of the pair about to be dropped)
                dy = self.cg * self.yg_prev + self.cy @ self.YY + self.cs @ self.SY
                row_sy = self.t * dy
                col_sy = sg - self.sg_prev                  # s_j . y_new
                col_yy = yg - self.yg_prev                  # y_j . y_new
                if self.k == m:
                    new = self.head
                    self.head = (self.head + 1) % m
                else:
                    new = (self.head + self.k) % m
                    self.k += 1
                for j in old:
                    if j == new:
                        continue
                    self.SY[new, j], self.SY[j, new] = row_sy[j], col_sy[j]
                    self.YY[new, j] = self.YY[j, new] = col_yy[j]
                self.SY[new, new], self.YY[new, new] = ys, yy
                self.ro[new] = 1.0 / ys
                sg[new], yg[new] = self.t * dg, gg - ggp
                self.H = ys / yy
                pushed = True
            order = self._order()
            qg, qy = -1.0, np.zeros(m)
            al = np.zeros(m)
            for i in reversed(order):
                al[i] = self.ro[i] * (qg * sg[i] + sum(qy[j] * self.SY[i, j] for j in order))
                qy[i] -= al[i]
            rs = np.zeros(m)
            for i in order:
                be = self.ro[i] * (self.H * (qg * yg[i] + sum(qy[j] * self.YY[i, j] for j in order))
                                   + sum(rs[j] * self.SY[j, i] for j in order))
                rs[i] += al[i] - be
            self.cg, self.cy, self.cs = self.H * qg, self.H * qy, rs
        order = self._order()
        gtd = self.cg * gg + sum(self.cy[j] * yg[j] + self.cs[j] * sg[j] for j in order)
        t_old = self.t
        self.t = min(1.0, 1.0 / gabs_sum) * self.lr if self.n_iter == 1 else self.lr
        skip_update = gtd > -self.tol_change
        # ---- pass B: write the new pair, form d, move x
        if pushed:
            self.S[new] = (self.d.astype(np.float64) * t_old).astype(np.float32)
            self.Y[new] = g - self.g_prev
        d64 = self.cg * g64
        for j in order:
            d64 += self.cy[j] * self.Y[j].astype(np.float64) + self.cs[j] * self.S[j].astype(np.float64)
        self.d = d64.astype(np.float32)
        self.g_prev = g.copy()
        self.sg_prev, self.yg_prev, self.gg_prev, self.dg_prev = sg, yg, gg, gtd
        self.loss_prev = loss
        if skip_update:
            return False
        self.x += (self.t * d64).astype(np.float32)
        self.dmax = float(np.abs(self.t * d64).max())
        return True

    def step(self, closure):
        """closure() -> (loss, grad float32) at the current self.x; mirrors torch.optim.LBFGS.step."""
        first_loss = None
        for it in range(self.max_iter):
            loss, g = closure()
            if first_loss is None:
                first_loss = loss
            if not self.iteration(g, loss, first_of_step=(it == 0)):
                break
        return first_loss
