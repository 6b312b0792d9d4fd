"""experiment: phased local moving + restarted multilevel passes (CPH_COMMUNITY_ITERS) against the sequential oracle on
noisy planted graphs; dQ is RELATIVE, (Q_gpu - Q_oracle) / |Q_oracle|"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from scipy.sparse import csr_matrix
def planted(n, g, p_in, p_out, seed, wmax=5.0):
    rng = np.random.default_rng(seed)
    truth = np.arange(n) % g
    P = np.where(truth[:, None] == truth[None, :], p_in, p_out)
    U = np.triu(rng.random((n, n)) < P, 1)
    W = np.triu(rng.uniform(0.5, wmax, size=(n, n)), 1) * U
    return csr_matrix(W + W.T), truth
from oracle import community_oracle as co
from cphasing_b200 import community
import torch
combos = [(8, 0), ('e', 0.9), ('e', 0.95), ('e', 0.98)]   # fixed phases | adaptive with eta_min
worst = {c: 0.0 for c in combos}
for (n, g, pi, po, seed) in [(400, 8, 0.25, 0.03, 2), (2000, 20, 0.1, 0.01, 3), (4000, 48, 0.3, 0.02, 4), (4000, 48, 0.2, 0.02, 5),
                             (3000, 30, 0.15, 0.015, 6)]:
    A, truth = planted(n, g, pi, po, seed)
    for method in ("louvain", "leiden"):
        for gamma in (0.5, 1.0, 2.0):
            t0 = time.perf_counter(); mo, _ = co.community(A, gamma, method); to = time.perf_counter() - t0
            qo = co.modularity(A, mo, gamma)
            row = [f"n={n} {method} g={gamma} oracle Q={qo:.5f} ({to*1e3:.0f} ms)"]
            for (P, it) in combos:
                if isinstance(P, str):
                    os.environ["CPH_COMMUNITY_PHASES"] = "0"
                    os.environ["CPH_COMMUNITY_ETA"] = str(it)
                else:
                    os.environ["CPH_COMMUNITY_PHASES"] = str(P)
                community.cluster(A, gamma, method)
                torch.cuda.synchronize(); t0 = time.perf_counter()
                vc = community.cluster(A, gamma, method)
                torch.cuda.synchronize(); tg = time.perf_counter() - t0
                d = (vc.quality - qo) / abs(qo)
                worst[(P, it)] = min(worst[(P, it)], d)
                row.append(f"P={P},{it}: dQ={d:+.4f} {tg*1e3:.1f}ms")
            print(" | ".join(row), flush=True)
print("worst relative dQ per combo:", {f"P={p},{i}": round(v, 4) for (p, i), v in worst.items()})
