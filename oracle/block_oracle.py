"""TEST INFRASTRUCTURE ONLY.  CPU restatement of the multi-mode block operator of the MATLAB application layer
 on top of a CPU `mapping` (oracle/oracleapi.OracleAniso or the compiled reference oracle/refapi.RefAniso):

  forward   aniso.m:121-136    rhs_i = sum_{j=-(K-1)}^{K-1} mapping(u_|j|, |i-j|),  i = -(K-1)..0, block |i|
  mforward  aniso.m:138-156    same with sigma_s .* u_|j| and the weight chi = (g^|j| - g^K)/(1 - g^K)
  gmres     gmres.cpp:53-169   the reference's own restarted GMRES (MGS + Givens), restated for an arbitrary operator;
                               aniso.m:165-169 calls MATLAB's gmres(A, rhs, 400, 1e-11, 400), which is not available here
                               ("the reference's own gmres.cpp, restart 400, tol 1e-11, maxit 400").
MATLAB itself cannot run in this container, so parity of this layer is pinned on the compiled reference `mapping`
underneath these loops, not on MATLAB output.  Never imported by the product package."""
import numpy as np


def chi(g, a, K):
    return (g ** a - g ** K) / (1.0 - g ** K)


def forward(o, u, K, n, weighted=False, g=0.0, sigma_s=None):
    rhs = np.zeros(K * n)
    for i in range(-(K - 1), 1):
        iid = abs(i)
        for j in range(-(K - 1), K):
            a = abs(j)
            blk = u[a * n:(a + 1) * n]
            if weighted:
                theta = o.mapping(sigma_s * blk, abs(i - j))
                rhs[iid * n:(iid + 1) * n] += chi(g, a, K) * theta
            else:
                rhs[iid * n:(iid + 1) * n] += o.mapping(blk, abs(i - j))
    return rhs


def mforward(o, u, K, n, g, sigma_s):
    return forward(o, u, K, n, True, g, sigma_s)


def _rotation(dx, dy):  # GeneratePlaneRotation, gmres.cpp:26-39
    if dy == 0.0:
        return 1.0, 0.0
    if abs(dy) > abs(dx):
        t = dx / dy
        sn = 1.0 / np.sqrt(1.0 + t * t)
        return t * sn, sn
    t = dy / dx
    cs = 1.0 / np.sqrt(1.0 + t * t)
    return cs, t * cs


def gmres(A, b, x0=None, m=400, max_iter=400, tol=1e-11, prec=None):
    """gmres.cpp:53-169.  Returns (x, rc, iterations, residual history); rc 0 = converged.
    prec: left preconditioner r -> M^-1 r (the hook aniso.m:165 passes to MATLAB's gmres; gmres.cpp:7 "todo:
    preconditioner"): the same recurrence on M^-1 A x = M^-1 b, residuals measured as |M^-1 (b - A x)| / |M^-1 b|."""
    if prec is not None:
        A0, b0 = A, b
        A = lambda v: prec(A0(v))  # noqa: E731
        b = prec(b0)
    n = b.size
    x = np.zeros(n) if x0 is None else x0.copy()
    m = min(m, max(max_iter, 1))
    H = np.zeros((m + 1, m))
    normb = np.linalg.norm(b)
    r = b - A(x)
    beta = np.linalg.norm(r)
    if normb == 0.0:
        normb = 1.0
    resid = beta / normb
    hist = []
    if resid <= tol:  # gmres.cpp:90 (<= here, < below)
        return x, 0, 0, np.array([resid])
    V = np.zeros((m + 1, n))
    j = 1
    iters = 0

    def update(k, s):  # Update, gmres.cpp:12-24
        y = s[:k + 1].copy()
        for i in range(k, -1, -1):
            y[i] /= H[i, i]
            for jj in range(i - 1, -1, -1):
                y[jj] -= H[jj, i] * y[i]
        for jj in range(k + 1):
            x[:] += y[jj] * V[jj]

    while j <= max_iter:
        V[0] = r / beta
        s = np.zeros(m + 1)
        s[0] = beta
        cs, sn = np.zeros(m + 1), np.zeros(m + 1)
        i = 0
        while i < m and j <= max_iter:
            hist.append(resid)
            iters = j
            w = A(V[i])
            for k in range(i + 1):
                H[k, i] = np.dot(w, V[k])
                w = w - H[k, i] * V[k]
            H[i + 1, i] = np.linalg.norm(w)
            V[i + 1] = w / H[i + 1, i]
            for k in range(i):
                t = cs[k] * H[k, i] + sn[k] * H[k + 1, i]
                H[k + 1, i] = -sn[k] * H[k, i] + cs[k] * H[k + 1, i]
                H[k, i] = t
            cs[i], sn[i] = _rotation(H[i, i], H[i + 1, i])
            t = cs[i] * H[i, i] + sn[i] * H[i + 1, i]
            H[i + 1, i] = -sn[i] * H[i, i] + cs[i] * H[i + 1, i]
            H[i, i] = t
            t = cs[i] * s[i] + sn[i] * s[i + 1]
            s[i + 1] = -sn[i] * s[i] + cs[i] * s[i + 1]
            s[i] = t
            resid = abs(s[i + 1]) / normb
            if resid < tol:
                update(i, s)
                hist.append(resid)
                return x, 0, iters, np.array(hist)
            i += 1
            j += 1
        update(i - 1, s)
        r = b - A(x)
        beta = np.linalg.norm(r)
        resid = beta / normb
        if resid < tol:
            hist.append(resid)
            return x, 0, iters, np.array(hist)
    hist.append(resid)
    return x, 1, iters, np.array(hist)
