"""numpy emulation of csrc/eigh_cluster_kernels.cu: one-sided block Jacobi, odd-even transposition over 32 block
positions with position swap, Gram eigensolve per pair; reports sweeps and accuracy."""
import numpy as np

def run(m, b, seed=0, reorder=True):
    rng = np.random.RandomState(seed)
    a = rng.standard_normal((m, m)); a = 0.5 * (a + a.T) * 3.0 + np.diag(np.linspace(-40.0, 25.0, m))
    npos = 32; mp = npos * b
    A = np.zeros((mp, mp)); A[:m, :m] = a
    rad = np.abs(A).sum(0) - np.abs(np.diag(A))
    lo = (np.diag(A) - rad)[:m].min(); hi = (np.diag(A) + rad)[:m].max()
    spread = max(hi - lo, 1.0); sigma = spread - lo
    G = A + sigma * np.eye(mp); V = np.eye(mp)
    blocks = [np.arange(p * b, (p + 1) * b) for p in range(npos)]      # column indices held at each position
    for sweep in range(30):
        smax = 0.0
        for rnd in range(npos):
            for p in range(rnd & 1, npos - 1, 2):
                cols = np.concatenate([blocks[p], blocks[p + 1]])
                X = G[:, cols]
                C = X.T @ X
                d = np.sqrt(np.outer(np.diag(C), np.diag(C)))
                rel = np.abs(C) / d; np.fill_diagonal(rel, 0.0)
                smax = max(smax, rel.max())
                w, Q = np.linalg.eigh(C)
                Q = Q[:, rng.permutation(2 * b)]        # arbitrary column order, like the warp Jacobi's positions
                if reorder:
                    # greedy: slot k takes the unassigned column with the largest |Q[k][c]| -> Q as close to I as possible
                    free = list(range(2 * b)); order = []
                    for k in range(2 * b):
                        c = max(free, key=lambda c: abs(Q[k, c])); free.remove(c); order.append(c)
                    Q = Q[:, order]
                G[:, cols] = X @ Q
                V[:, cols] = V[:, cols] @ Q
                # new first half -> right position, second half -> left
                blocks[p], blocks[p + 1] = cols[b:], cols[:b]
        print('  sweep', sweep, 'max rel %.3e' % smax)
        if smax <= 1e-9:
            break
    lam = np.einsum('ij,ij->j', V, G) - sigma
    pad = np.abs(V[m:, :]).sum(0) > 0.5
    ev = np.sort(lam[~pad])
    assert ev.shape[0] == m, (ev.shape, m)
    ref = np.linalg.eigvalsh(a)
    print('m', m, 'b', b, 'sweeps', sweep + 1, 'max eig err %.2e' % np.abs(ev - ref).max())

run(130, 8)
run(240, 8)
run(200, 8)
run(125, 4)
