"""numpy replay of the multifrontal schedule produced by the host symbolic analysis
(`plate_inverse_problem_b200/csrc/symbolic.cpp`), with the same factor format as the CUDA
kernels (kernels.cuh header): pivot tiles of width 32, diagonal tiles stored inverted,
Lt = F21 D^-1, U kept as is.  CPU test infrastructure: it exercises every index map the
GPU kernels consume (entry slots, relmaps, row lists, level order) without a GPU."""
import numpy as np

TW = 32


class Emulator:
    def __init__(self, S, N):
        self.S, self.N = S, N
        nf = S["nfronts"]
        self.children = [[] for _ in range(nf)]
        for f in range(nf):
            p = S["f_parent"][f]
            if p >= 0:
                assert p > f
                self.children[p].append(f)
        # level order, as the kernels are launched
        self.order = sorted(range(nf), key=lambda f: (S["f_level"][f], f))

    def rows(self, f):
        S = self.S
        return S["rowidx"][S["f_rowptr"][f]:S["f_rowptr"][f + 1]]

    def factor(self, data):
        S = self.S
        self.F = {}
        for f in self.order:
            s, k, ld = int(S["f_s"][f]), int(S["f_k"][f]), int(S["f_ld"][f])
            F = np.zeros(s * ld, dtype=np.complex128)
            e0, e1 = S["f_entptr"][f], S["f_entptr"][f + 1]
            F[S["ent_slot"][e0:e1]] = data[S["ent_src"][e0:e1]]
            F = F.reshape(s, ld)[:, :s].copy()
            for c in self.children[f]:
                kc = int(S["f_k"][c])
                mc = int(S["f_s"][c]) - kc
                rel = S["relmap"][S["f_reloff"][c]:S["f_reloff"][c] + mc]
                F[np.ix_(rel, rel)] += self.F[c][kc:, kc:]
            for j0 in range(0, k, TW):
                j1 = min(j0 + TW, k)
                if j0 > 0:
                    F[j0:, j0:j1] -= F[j0:, :j0] @ F[:j0, j0:j1]
                    F[j0:j1, j1:] -= F[j0:j1, :j0] @ F[:j0, j1:]
                Dinv = np.linalg.inv(F[j0:j1, j0:j1])
                F[j0:j1, j0:j1] = Dinv
                F[j1:, j0:j1] = F[j1:, j0:j1] @ Dinv
            if s > k:
                F[k:, k:] -= F[k:, :k] @ F[:k, k:]
            self.F[f] = F

    def solve(self, b, trans=False):
        S = self.S
        perm = S["perm"]
        V = {}
        for f in self.order:
            s, k = int(S["f_s"][f]), int(S["f_k"][f])
            rows = perm[self.rows(f)]
            v = np.zeros(s, dtype=np.complex128)
            v[:k] = b[rows[:k]]
            for c in self.children[f]:
                kc = int(S["f_k"][c])
                mc = int(S["f_s"][c]) - kc
                rel = S["relmap"][S["f_reloff"][c]:S["f_reloff"][c] + mc]
                v[rel] += V[c][kc:]
            F = self.F[f]
            for j0 in range(0, k, TW):
                j1 = min(j0 + TW, k)
                if not trans:
                    v[j1:] -= F[j1:, j0:j1] @ v[j0:j1]
                else:
                    v[j0:j1] = F[j0:j1, j0:j1].T @ v[j0:j1]
                    v[j1:] -= F[j0:j1, j1:].T @ v[j0:j1]
            V[f] = v
        x = np.zeros(self.N, dtype=np.complex128)
        for f in reversed(self.order):
            s, k = int(S["f_s"][f]), int(S["f_k"][f])
            rows = perm[self.rows(f)]
            v = V[f].copy()
            v[k:] = x[rows[k:]]
            F = self.F[f]
            for j0 in reversed(range(0, k, TW)):
                j1 = min(j0 + TW, k)
                if not trans:
                    v[j0:j1] = F[j0:j1, j0:j1] @ (v[j0:j1] - F[j0:j1, j1:] @ v[j1:])
                else:
                    v[j0:j1] -= F[j1:, j0:j1].T @ v[j1:]
            x[rows[:k]] = v[:k]
        return x
