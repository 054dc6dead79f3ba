// Per-thread bodies of the batched multifrontal kernels ("phases").
//
// Data layout: systems are interleaved in groups of BL ("chunk");
// element `slot` of system b lives at  buf[(chunk * slots_per_system + slot) * BL + lane],
// chunk = b / BL, lane = b % BL, every element one complex double (16 B).  A CUDA thread
// is (worker, lane): the BL lanes of a worker run the same control flow on BL different
// systems, so every global access of a worker is one contiguous BL*16-byte segment and
// all index arithmetic is shared by the lanes.
//
// Every function here is a *phase*: it only reads what earlier phases (separated by a
// block barrier) wrote.  engine.cu wraps the phases in __global__ kernels with
// __syncthreads() between them; tests/emul/emul.cpp runs the same phases serially on the
// host so that the index logic is testable without a GPU.
#pragma once
#include <cstdint>
#include <cmath>

#if defined(__CUDACC__)
#define JPB_HD __host__ __device__ __forceinline__
#else
#define JPB_HD inline
#endif

namespace jpb {

struct alignas(16) cd { double re, im; };

JPB_HD cd cmake(double r, double i) { cd z; z.re = r; z.im = i; return z; }
JPB_HD cd cmul(cd a, cd b) {
    return cmake(fma(a.re, b.re, -(a.im * b.im)), fma(a.re, b.im, a.im * b.re));
}
JPB_HD void cadd(cd& c, cd a) { c.re += a.re; c.im += a.im; }
JPB_HD void csub(cd& c, cd a) { c.re -= a.re; c.im -= a.im; }
// c += a*b
JPB_HD void cfma(cd& c, cd a, cd b) {
    c.re = fma(a.re, b.re, c.re); c.re = fma(-a.im, b.im, c.re);
    c.im = fma(a.re, b.im, c.im); c.im = fma(a.im, b.re, c.im);
}
// c -= a*b
JPB_HD void cfnma(cd& c, cd a, cd b) {
    c.re = fma(-a.re, b.re, c.re); c.re = fma(a.im, b.im, c.re);
    c.im = fma(-a.re, b.im, c.im); c.im = fma(-a.im, b.re, c.im);
}
JPB_HD cd crcp(cd a) {
    double d = fma(a.re, a.re, a.im * a.im);
    double r = 1.0 / d;
    return cmake(a.re * r, -a.im * r);
}
JPB_HD cd cneg(cd a) { return cmake(-a.re, -a.im); }

struct FrontDesc {
    int64_t off;     // first slot of the s*s row-major front
    int64_t invoff;  // this front's inverse map into its parent (length s_parent), -1 = root
    int s, k, c0, voff;
    int rowptr;      // into rowidx
    int childptr, nchild;
    int parent;
};

struct SymDev {
    const FrontDesc* fr;
    const int* rowidx;
    const int* children;
    const int* invmap;
    const int* perm;   // perm[new] = old
    int64_t front_slots;
    int64_t vec_slots;
    int N;
};

// one unit of work of a launch; meaning of the fields depends on the kernel
struct Item { int front, jb0, jb1, flag, ti0, tj0, r0, r1; };
enum { ITEM_INLINE_UPDATE = 1 };

// tile t of a front dimension: pivot tiles first (last one may be short), then the
// contribution rows, again in fours.
JPB_HD void tile_span(int t, int k, int s, int KT, int& r0, int& n) {
    if (t < KT) { r0 = 4 * t; n = k - r0; }
    else { r0 = k + 4 * (t - KT); n = s - r0; }
    n = n < 4 ? n : 4;
}

template <int BL>
JPB_HD void load_tile(const cd* F, int s, int lane, int r0, int nr, int c0, int nc,
                      cd (&T)[4][4]) {
    const cd* p = F + ((int64_t)r0 * s + c0) * BL + lane;
    if (nr == 4 && nc == 4) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) T[i][j] = p[((int64_t)i * s + j) * BL];
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                T[i][j] = (i < nr && j < nc) ? p[((int64_t)i * s + j) * BL] : cmake(0.0, 0.0);
    }
}

template <int BL>
JPB_HD void store_tile(cd* F, int s, int lane, int r0, int nr, int c0, int nc,
                       const cd (&T)[4][4]) {
    cd* p = F + ((int64_t)r0 * s + c0) * BL + lane;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (i < nr && j < nc) p[((int64_t)i * s + j) * BL] = T[i][j];
}

// C -= F[r0.., p] * F[p, c0..]  for p in [p0, p1)
template <int BL>
JPB_HD void gemm_acc(cd (&C)[4][4], const cd* F, int s, int lane, int r0, int nr, int c0,
                     int nc, int p0, int p1) {
    const cd* a = F + ((int64_t)r0 * s + p0) * BL + lane;
    const cd* b = F + ((int64_t)p0 * s + c0) * BL + lane;
    const int64_t rs = (int64_t)s * BL;
    if (nr == 4 && nc == 4) {
#pragma unroll 2
        for (int p = p0; p < p1; ++p) {
            cd av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = a[i * rs];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = b[j * BL];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) cfnma(C[i][j], av[i], bv[j]);
            a += BL;
            b += rs;
        }
    } else {
        for (int p = p0; p < p1; ++p) {
            cd av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = i < nr ? a[i * rs] : cmake(0.0, 0.0);
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = j < nc ? b[j * BL] : cmake(0.0, 0.0);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) cfnma(C[i][j], av[i], bv[j]);
            a += BL;
            b += rs;
        }
    }
}

// ------------------------------------------------------------------ numeric factorisation
// shared scratch of the panel kernel: 2 parities x {invL, invU} x 16 entries x BL lanes
template <int BL> struct PanelSmem { static constexpr int kElems = 2 * 32 * BL; };

// Diagonal pivot tile J: left-looking update with the pivot tiles [jb0, J) of the current
// block, 4x4 LU without pivoting (static pivoting: the plate matrices have a definite
// imaginary part), reciprocals of the pivots stored on the diagonal,
// explicit inverses of both triangles written to `inv` for the off-diagonal tiles.
template <int BL>
JPB_HD void panel_diag(cd* F, int s, int k, int lane, int jb0, int J, cd* inv) {
    const int r0 = 4 * J;
    const int n = (k - r0) < 4 ? (k - r0) : 4;
    cd D[4][4];
    load_tile<BL>(F, s, lane, r0, n, r0, n, D);
    gemm_acc<BL>(D, F, s, lane, r0, n, r0, n, 4 * jb0, r0);
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i >= n) D[i][i] = cmake(1.0, 0.0);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        cd r = crcp(D[p][p]);
        D[p][p] = r;
#pragma unroll
        for (int i = p + 1; i < 4; ++i) {
            D[i][p] = cmul(D[i][p], r);
#pragma unroll
            for (int j = p + 1; j < 4; ++j) cfnma(D[i][j], D[i][p], D[p][j]);
        }
    }
    store_tile<BL>(F, s, lane, r0, n, r0, n, D);
    // X = L^-1 (unit lower), Y = U^-1 (upper; D[i][i] already holds 1/u_ii)
    cd X[4][4], Y[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            X[i][j] = cmake(i == j ? 1.0 : 0.0, 0.0);
            Y[i][j] = cmake(0.0, 0.0);
        }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = j + 1; i < 4; ++i) {
            cd acc = cmake(0.0, 0.0);
#pragma unroll
            for (int m = j; m < i; ++m) cfnma(acc, D[i][m], X[m][j]);
            X[i][j] = acc;
        }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        Y[j][j] = D[j][j];
#pragma unroll
        for (int i = j - 1; i >= 0; --i) {
            cd acc = cmake(0.0, 0.0);
#pragma unroll
            for (int m = i + 1; m <= j; ++m) cfnma(acc, D[i][m], Y[m][j]);
            Y[i][j] = cmul(acc, D[i][i]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            inv[(i * 4 + j) * BL + lane] = X[i][j];
            inv[(16 + i * 4 + j) * BL + lane] = Y[i][j];
        }
}

// kind 0: L tile (I, J) = S * U_JJ^-1 ; kind 1: U tile (J, I) = L_JJ^-1 * S
template <int BL>
JPB_HD void panel_offdiag(cd* F, int s, int k, int KT, int lane, int jb0, int J, int I,
                          int kind, const cd* inv) {
    int jr0, jn, ir0, in_;
    tile_span(J, k, s, KT, jr0, jn);
    tile_span(I, k, s, KT, ir0, in_);
    cd S[4][4], R[4][4];
    if (kind == 0) {
        load_tile<BL>(F, s, lane, ir0, in_, jr0, jn, S);
        gemm_acc<BL>(S, F, s, lane, ir0, in_, jr0, jn, 4 * jb0, jr0);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                cd acc = cmake(0.0, 0.0);
#pragma unroll
                for (int m = 0; m <= j; ++m) cfma(acc, S[i][m], inv[(16 + m * 4 + j) * BL + lane]);
                R[i][j] = acc;
            }
        store_tile<BL>(F, s, lane, ir0, in_, jr0, jn, R);
    } else {
        load_tile<BL>(F, s, lane, jr0, jn, ir0, in_, S);
        gemm_acc<BL>(S, F, s, lane, jr0, jn, ir0, in_, 4 * jb0, jr0);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                cd acc = cmake(0.0, 0.0);
#pragma unroll
                for (int m = 0; m <= i; ++m) cfma(acc, inv[(i * 4 + m) * BL + lane], S[m][j]);
                R[i][j] = acc;
            }
        store_tile<BL>(F, s, lane, jr0, jn, ir0, in_, R);
    }
}

// One step of the left-looking panel factorisation of pivot tile column/row J inside the
// block [jb0, jb1).  Worker 0 finishes the two tiles next to the diagonal first and then
// factors the next diagonal tile, so that a single barrier per step suffices.
template <int BL>
JPB_HD void panel_step(cd* F, int s, int k, int lane, int w, int NW, int jb0, int jb1, int J,
                       cd* sh) {
    const int KT = (k + 3) >> 2;
    const int T = KT + ((s - k + 3) >> 2);
    const cd* inv = sh + (J & 1) * 32 * BL;
    const int nitems = 2 * (T - J - 1);
    if (w == 0) {
        if (nitems > 0) {
            panel_offdiag<BL>(F, s, k, KT, lane, jb0, J, J + 1, 0, inv);
            panel_offdiag<BL>(F, s, k, KT, lane, jb0, J, J + 1, 1, inv);
        }
        if (NW == 1)
            for (int q = 2; q < nitems; ++q)
                panel_offdiag<BL>(F, s, k, KT, lane, jb0, J, J + 1 + (q >> 1), q & 1, inv);
        if (J + 1 < jb1) panel_diag<BL>(F, s, k, lane, jb0, J + 1, sh + ((J + 1) & 1) * 32 * BL);
    } else {
        for (int q = 2 + (w - 1); q < nitems; q += NW - 1)
            panel_offdiag<BL>(F, s, k, KT, lane, jb0, J, J + 1 + (q >> 1), q & 1, inv);
    }
}

// trailing tile (I, I2) -= L(I, block) * U(block, I2)
template <int BL>
JPB_HD void trailing_tile(cd* F, int s, int k, int KT, int lane, int jb0, int jb1, int I,
                          int I2) {
    int ir0, in_, cr0, cn;
    tile_span(I, k, s, KT, ir0, in_);
    tile_span(I2, k, s, KT, cr0, cn);
    int p1 = 4 * jb1 < k ? 4 * jb1 : k;
    cd C[4][4];
    load_tile<BL>(F, s, lane, ir0, in_, cr0, cn, C);
    gemm_acc<BL>(C, F, s, lane, ir0, in_, cr0, cn, 4 * jb0, p1);
    store_tile<BL>(F, s, lane, ir0, in_, cr0, cn, C);
}

// ------------------------------------------------------------------ front assembly
// Rows [r0, r1) of a front: every slot is written exactly once with the sum of the
// children's contribution-block entries that land on it (gather through the children's
// inverse maps, children in ascending id order => deterministic).  Entries of A are added
// afterwards by the entry kernel.
template <int BL>
JPB_HD void assemble_rows(const SymDev& sd, const FrontDesc& fd, cd* Fsys /* chunk base */,
                          int lane, int w, int NW, int r0, int r1) {
    cd* F = Fsys + fd.off * BL;
    const int s = fd.s;
    for (int i = r0; i < r1; ++i) {
        for (int j = w; j < s; j += NW) {
            cd v = cmake(0.0, 0.0);
            for (int c = 0; c < fd.nchild; ++c) {
                const FrontDesc& ch = sd.fr[sd.children[fd.childptr + c]];
                const int* inv = sd.invmap + ch.invoff;
                int ci = inv[i];
                if (ci < 0) continue;
                int cj = inv[j];
                if (cj < 0) continue;
                cadd(v, Fsys[(ch.off + (int64_t)ci * ch.s + cj) * BL + lane]);
            }
            F[((int64_t)i * s + j) * BL + lane] = v;
        }
    }
}

// ------------------------------------------------------------------ triangular solves
// element (i, j) of the triangular factor in use: L or U as stored (TRANS = false) or the
// plain transposes U^T / L^T (TRANS = true, the reference's UMFPACK_Aat solve).
template <int BL, bool TRANS>
JPB_HD cd fel(const cd* F, int s, int lane, int i, int j) {
    return TRANS ? F[((int64_t)j * s + i) * BL + lane] : F[((int64_t)i * s + j) * BL + lane];
}

// forward sweep, phase 0: local vector = rhs on the pivot rows + children's updates
template <int BL>
JPB_HD void fwd_gather(const SymDev& sd, const FrontDesc& fd, cd* Vsys, const cd* rhs_sys,
                       int lane, int w, int NW) {
    cd* V = Vsys + (int64_t)fd.voff * BL;
    for (int i = w; i < fd.s; i += NW) {
        cd v = cmake(0.0, 0.0);
        if (i < fd.k) v = rhs_sys[(int64_t)sd.perm[fd.c0 + i] * BL + lane];
        for (int c = 0; c < fd.nchild; ++c) {
            const FrontDesc& ch = sd.fr[sd.children[fd.childptr + c]];
            int ci = sd.invmap[ch.invoff + i];
            if (ci >= 0) cadd(v, Vsys[((int64_t)ch.voff + ci) * BL + lane]);
        }
        V[(int64_t)i * BL + lane] = v;
    }
}

// in-tile forward substitution on pivot tile J (lower triangle; unit diagonal unless TRANS)
template <int BL, bool TRANS>
JPB_HD void fwd_solve_tile(const cd* F, int s, int k, cd* V, int lane, int J) {
    const int r0 = 4 * J;
    const int n = (k - r0) < 4 ? (k - r0) : 4;
    cd v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = i < n ? V[(int64_t)(r0 + i) * BL + lane] : cmake(0.0, 0.0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i < n) {
#pragma unroll
            for (int m = 0; m < i; ++m) cfnma(v[i], fel<BL, TRANS>(F, s, lane, r0 + i, r0 + m), v[m]);
            if (TRANS) v[i] = cmul(v[i], F[((int64_t)(r0 + i) * s + r0 + i) * BL + lane]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < n) V[(int64_t)(r0 + i) * BL + lane] = v[i];
}

// v_I -= T(I rows, [p0,p1)) * v[p0..p1)
template <int BL, bool TRANS>
JPB_HD void vec_update_tile(const cd* F, int s, cd* V, int lane, int ir0, int in_, int p0,
                            int p1) {
    cd acc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = cmake(0.0, 0.0);
    for (int p = p0; p < p1; ++p) {
        cd y = V[(int64_t)p * BL + lane];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < in_) cfnma(acc[i], fel<BL, TRANS>(F, s, lane, ir0 + i, p), y);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < in_) cadd(V[(int64_t)(ir0 + i) * BL + lane], acc[i]);
}

template <int BL, bool TRANS>
JPB_HD void fwd_tri_step(const cd* F, int s, int k, cd* V, int lane, int w, int NW, int J) {
    // pivot tile J is solved; update pivot tiles I in (J, KT), worker 0 solves tile J+1
    const int KT = (k + 3) >> 2;
    const int p0 = 4 * J, p1 = 4 * J + 4;   // J < KT-1 => full tile
    if (w == 0) {
        int I = J + 1;
        int ir0 = 4 * I, in_ = (k - ir0) < 4 ? (k - ir0) : 4;
        vec_update_tile<BL, TRANS>(F, s, V, lane, ir0, in_, p0, p1);
        if (NW == 1)
            for (int I2 = J + 2; I2 < KT; ++I2) {
                int r0 = 4 * I2, n = (k - r0) < 4 ? (k - r0) : 4;
                vec_update_tile<BL, TRANS>(F, s, V, lane, r0, n, p0, p1);
            }
        fwd_solve_tile<BL, TRANS>(F, s, k, V, lane, I);
    } else {
        for (int I = J + 2 + (w - 1); I < KT; I += NW - 1) {
            int r0 = 4 * I, n = (k - r0) < 4 ? (k - r0) : 4;
            vec_update_tile<BL, TRANS>(F, s, V, lane, r0, n, p0, p1);
        }
    }
}

// contribution rows: v_I -= T(I, pivots) y   (after all pivot tiles are solved)
template <int BL, bool TRANS>
JPB_HD void fwd_rect(const cd* F, int s, int k, cd* V, int lane, int w, int NW) {
    const int MT = (s - k + 3) >> 2;
    for (int u = w; u < MT; u += NW) {
        int r0 = k + 4 * u, n = (s - r0) < 4 ? (s - r0) : 4;
        vec_update_tile<BL, TRANS>(F, s, V, lane, r0, n, 0, k);
    }
}

// backward sweep, phase 0: fetch the already known unknowns of the contribution rows
template <int BL>
JPB_HD void bwd_gather(const SymDev& sd, const FrontDesc& fd, cd* Vsys, const cd* Xsys,
                       int lane, int w, int NW) {
    cd* V = Vsys + (int64_t)fd.voff * BL;
    const int* rows = sd.rowidx + fd.rowptr;
    for (int i = fd.k + w; i < fd.s; i += NW)
        V[(int64_t)i * BL + lane] = Xsys[(int64_t)rows[i] * BL + lane];
}

// number of column segments the rectangular part of the backward sweep is split into
JPB_HD int bwd_nsplit(int KT, int NW) { return KT >= NW ? 1 : NW / KT; }

// phase 1a: partial sums of  T(I rows, contribution columns) * w
template <int BL, bool TRANS>
JPB_HD void bwd_rect_partial(const cd* F, int s, int k, cd* V, cd* sh, int lane, int w,
                             int NW) {
    const int KT = (k + 3) >> 2;
    const int ns = bwd_nsplit(KT, NW);
    const int m = s - k;
    if (ns == 1) {
        for (int I = w; I < KT; I += NW) {
            int r0 = 4 * I, n = (k - r0) < 4 ? (k - r0) : 4;
            vec_update_tile<BL, TRANS>(F, s, V, lane, r0, n, k, s);
        }
        return;
    }
    if (w >= KT * ns) return;
    const int I = w / ns, part = w % ns;
    const int seg = (m + ns - 1) / ns;
    int p0 = k + part * seg, p1 = p0 + seg;
    if (p1 > s) p1 = s;
    const int r0 = 4 * I, n = (k - r0) < 4 ? (k - r0) : 4;
    cd acc[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = cmake(0.0, 0.0);
    for (int p = p0; p < p1; ++p) {
        cd y = V[(int64_t)p * BL + lane];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < n) cfnma(acc[i], fel<BL, TRANS>(F, s, lane, r0 + i, p), y);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) sh[((int64_t)w * 4 + i) * BL + lane] = acc[i];
}

// phase 1b: ordered reduction of the partial sums
template <int BL>
JPB_HD void bwd_rect_reduce(int k, cd* V, const cd* sh, int lane, int w, int NW) {
    const int KT = (k + 3) >> 2;
    const int ns = bwd_nsplit(KT, NW);
    if (ns == 1 || w >= KT) return;
    const int I = w;
    const int r0 = 4 * I, n = (k - r0) < 4 ? (k - r0) : 4;
    for (int i = 0; i < n; ++i) {
        cd v = V[(int64_t)(r0 + i) * BL + lane];
        for (int part = 0; part < ns; ++part)
            cadd(v, sh[((int64_t)(I * ns + part) * 4 + i) * BL + lane]);
        V[(int64_t)(r0 + i) * BL + lane] = v;
    }
}

// in-tile back substitution on pivot tile J (upper triangle; unit diagonal if TRANS)
template <int BL, bool TRANS>
JPB_HD void bwd_solve_tile(const cd* F, int s, int k, cd* V, int lane, int J) {
    const int r0 = 4 * J;
    const int n = (k - r0) < 4 ? (k - r0) : 4;
    cd v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = i < n ? V[(int64_t)(r0 + i) * BL + lane] : cmake(0.0, 0.0);
#pragma unroll
    for (int i = 3; i >= 0; --i) {
        if (i < n) {
#pragma unroll
            for (int m = i + 1; m < 4; ++m)
                if (m < n) cfnma(v[i], fel<BL, TRANS>(F, s, lane, r0 + i, r0 + m), v[m]);
            if (!TRANS) v[i] = cmul(v[i], F[((int64_t)(r0 + i) * s + r0 + i) * BL + lane]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < n) V[(int64_t)(r0 + i) * BL + lane] = v[i];
}

template <int BL, bool TRANS>
JPB_HD void bwd_tri_step(const cd* F, int s, int k, cd* V, int lane, int w, int NW, int J) {
    // pivot tile J (>= 1) is solved; update tiles I < J, worker 0 solves tile J-1
    const int p0 = 4 * J;
    const int p1 = (p0 + 4) < k ? (p0 + 4) : k;
    if (w == 0) {
        vec_update_tile<BL, TRANS>(F, s, V, lane, 4 * (J - 1), 4, p0, p1);
        if (NW == 1)
            for (int I = J - 2; I >= 0; --I) vec_update_tile<BL, TRANS>(F, s, V, lane, 4 * I, 4, p0, p1);
        bwd_solve_tile<BL, TRANS>(F, s, k, V, lane, J - 1);
    } else {
        for (int I = J - 2 - (w - 1); I >= 0; I -= NW - 1)
            vec_update_tile<BL, TRANS>(F, s, V, lane, 4 * I, 4, p0, p1);
    }
}

template <int BL>
JPB_HD void bwd_write(const FrontDesc& fd, const cd* Vsys, cd* Xsys, int lane, int w, int NW) {
    const cd* V = Vsys + (int64_t)fd.voff * BL;
    for (int i = w; i < fd.k; i += NW)
        Xsys[(int64_t)(fd.c0 + i) * BL + lane] = V[(int64_t)i * BL + lane];
}

}  // namespace jpb
