// sm_100a kernels of the batched multifrontal complex-FP64 LU and everything around it.
//
// Storage.  For every system of a wave the fronts are stored one
// after the other; front f is a row-major s x ld complex128 matrix at slot offset off(f).
// After the numeric phase the front holds, with pivot tiles of width 32 (the last one of
// a front may be narrower):
//     diagonal pivot tile J          : inverse of the (updated) diagonal block, D_J^-1
//     rows below tile J, its columns : Lt = F[rows, J] * D_J^-1        ("unit block lower")
//     rows of tile J, columns right  : U  = updated F[J, cols]          (as is)
//     rows/cols >= k                 : Schur complement, added into the parent front
// so A = Lt * Ubar with Ubar block upper triangular with the D_J on its diagonal.  The
// trailing updates C -= Lt * U are done on the FP64 tensor cores (DMMA m8n8k4).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace jpb {

typedef double2 cd;   // x = re, y = im

struct FrontDesc {
    int64_t off;      // first slot of the front
    int64_t entptr;   // first A entry of the front (into ent_src / ent_slot)
    int s, k, ld, nent;
    int rowptr;       // into rowidx_orig
    int voff;         // first slot of the front-local vector
    int childptr, nchild;
    int reloff;       // this front's CB rows -> parent-local rows (into relmap)
    int parent;
    int npan;         // as a child: how many of its contribution rows land on pivots of the parent
    int invoff;       // as a child: first slot of its inverse map in cbinv
};

struct SymDev {
    const FrontDesc* fr;
    const int* rowidx_orig;
    const int* children;
    const int* relmap;
    const int* ent_src;
    const int* ent_slot;
    // for child c and contribution row q of its parent (q = parent row - k_parent):
    // cbinv[invoff(c) + q] = contribution row of c that lands there, or -1
    const int* cbinv;
    int64_t front_slots;
    int64_t vec_slots;
    int64_t nnz;
    int N;
};

__device__ __forceinline__ cd cmul(cd a, cd b) {
    return make_double2(fma(a.x, b.x, -(a.y * b.y)), fma(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ void cfma(cd& c, cd a, cd b) {
    c.x = fma(a.x, b.x, c.x); c.x = fma(-a.y, b.y, c.x);
    c.y = fma(a.x, b.y, c.y); c.y = fma(a.y, b.x, c.y);
}
__device__ __forceinline__ cd crcp(cd a) {
    // scaled to survive |a|^2 over/underflow
    double m = fmax(fabs(a.x), fabs(a.y));
    double sx = a.x / m, sy = a.y / m;
    double d = 1.0 / (m * fma(sx, sx, sy * sy));
    return make_double2(sx * d, -sy * d);
}
__device__ __forceinline__ cd warp_sum(cd v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    }
    return v;
}

// ---------------------------------------------------------------------------- DMMA tiles
__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

template <int MB, int NB>
struct Acc {
    double re[MB][NB][2];
    double im[MB][NB][2];
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int m = 0; m < MB; ++m)
#pragma unroll
            for (int n = 0; n < NB; ++n) { re[m][n][0] = re[m][n][1] = im[m][n][0] = im[m][n][1] = 0.0; }
    }
};

// acc += A[0:8MB, 0:K] * B[0:K, 0:8NB];  A(i,p) = A[i*lda + p] valid for i < mrows,
// B(p,j) = B[p*ldb + j] valid for j < ncols (mrows, ncols >= 1).
// Fragment layout of mma.m8n8k4.f64: lane = 4*g + t holds A(g, t), B(t, g) and
// C(g, 2t), C(g, 2t+1).  One 16-byte load per lane yields the re and im fragments.
// Rows / columns past the edge are clamped onto the last valid one: they only feed
// accumulator rows / columns that store_tile never writes, so the main loop carries no
// predicates and advances plain pointers; only a K tail (K % 4) is masked.
template <int MB, int NB>
__device__ __forceinline__ void warp_mma_step(Acc<MB, NB>& acc, const cd (&a)[MB], const cd (&b)[NB]) {
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n) dmma(acc.re[m][n][0], acc.re[m][n][1], a[m].x, b[n].x);
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n) dmma(acc.im[m][n][0], acc.im[m][n][1], a[m].x, b[n].y);
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n) dmma(acc.re[m][n][0], acc.re[m][n][1], -a[m].y, b[n].y);
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n) dmma(acc.im[m][n][0], acc.im[m][n][1], a[m].y, b[n].x);
}

template <int MB, int NB>
__device__ __forceinline__ void warp_gemm(Acc<MB, NB>& acc, const cd* A, int lda, int mrows,
                                          const cd* B, int ldb, int ncols, int K, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const cd* pa[MB];
    const cd* pb[NB];
#pragma unroll
    for (int m = 0; m < MB; ++m) pa[m] = A + (int64_t)min(8 * m + g, mrows - 1) * lda + t;
#pragma unroll
    for (int n = 0; n < NB; ++n) pb[n] = B + (int64_t)t * ldb + min(8 * n + g, ncols - 1);
    const int64_t bstep = 4 * (int64_t)ldb;
    const int nfull = K >> 2;
    cd a[MB], b[NB];
    if (nfull > 0) {
#pragma unroll
        for (int m = 0; m < MB; ++m) a[m] = *pa[m];
#pragma unroll
        for (int n = 0; n < NB; ++n) b[n] = *pb[n];
    }
    // two k-steps per trip so that the operand sets ping-pong without register moves
    int it = 1;
#pragma unroll 1
    for (; it + 1 < nfull; it += 2) {
        cd an[MB], bn[NB];
#pragma unroll
        for (int m = 0; m < MB; ++m) { pa[m] += 4; an[m] = *pa[m]; }
#pragma unroll
        for (int n = 0; n < NB; ++n) { pb[n] += bstep; bn[n] = *pb[n]; }
        warp_mma_step<MB, NB>(acc, a, b);
#pragma unroll
        for (int m = 0; m < MB; ++m) { pa[m] += 4; a[m] = *pa[m]; }
#pragma unroll
        for (int n = 0; n < NB; ++n) { pb[n] += bstep; b[n] = *pb[n]; }
        warp_mma_step<MB, NB>(acc, an, bn);
    }
    if (it < nfull) {
        cd an[MB], bn[NB];
#pragma unroll
        for (int m = 0; m < MB; ++m) { pa[m] += 4; an[m] = *pa[m]; }
#pragma unroll
        for (int n = 0; n < NB; ++n) { pb[n] += bstep; bn[n] = *pb[n]; }
        warp_mma_step<MB, NB>(acc, a, b);
#pragma unroll
        for (int m = 0; m < MB; ++m) a[m] = an[m];
#pragma unroll
        for (int n = 0; n < NB; ++n) b[n] = bn[n];
    }
    const int rem = K & 3;
    if (rem) {      // K tail: both operands masked (0 * garbage could be NaN)
        cd an[MB], bn[NB];
        const cd zero = make_double2(0.0, 0.0);
        const int64_t adv = nfull > 0 ? 4 : 0;
#pragma unroll
        for (int m = 0; m < MB; ++m) an[m] = t < rem ? pa[m][adv] : zero;
#pragma unroll
        for (int n = 0; n < NB; ++n) bn[n] = t < rem ? pb[n][adv * ldb] : zero;
        if (nfull > 0) warp_mma_step<MB, NB>(acc, a, b);
        warp_mma_step<MB, NB>(acc, an, bn);
    } else if (nfull > 0) {
        warp_mma_step<MB, NB>(acc, a, b);
    }
}

// The same product with both operands in shared memory: at ~30 cycles of load latency the
// next k-step's operands need no register double buffer, which frees 24 registers and lets
// more of the small shared-memory-resident fronts be co-resident.
template <int MB, int NB>
__device__ __forceinline__ void warp_gemm_shared(Acc<MB, NB>& acc, const cd* A, int lda, int mrows,
                                                 const cd* B, int ldb, int ncols, int K, int lane) {
    const int g = lane >> 2, t = lane & 3;
    int oa[MB], ob[NB];
#pragma unroll
    for (int m = 0; m < MB; ++m) oa[m] = min(8 * m + g, mrows - 1) * lda + t;
#pragma unroll
    for (int n = 0; n < NB; ++n) ob[n] = t * ldb + min(8 * n + g, ncols - 1);
    const int nfull = K >> 2;
#pragma unroll 1
    for (int it = 0; it < nfull; ++it) {
        cd a[MB], b[NB];
#pragma unroll
        for (int m = 0; m < MB; ++m) a[m] = A[oa[m] + 4 * it];
#pragma unroll
        for (int n = 0; n < NB; ++n) b[n] = B[ob[n] + 4 * it * ldb];
        warp_mma_step<MB, NB>(acc, a, b);
    }
    const int rem = K & 3;
    if (rem) {
        cd a[MB], b[NB];
        const cd zero = make_double2(0.0, 0.0);
#pragma unroll
        for (int m = 0; m < MB; ++m) a[m] = t < rem ? A[oa[m] + 4 * nfull] : zero;
#pragma unroll
        for (int n = 0; n < NB; ++n) b[n] = t < rem ? B[ob[n] + 4 * nfull * ldb] : zero;
        warp_mma_step<MB, NB>(acc, a, b);
    }
}

// Ask L1 for the 16 x (8 NB) tile at C ahead of the k-loop, so that the read-modify-write of
// store_tile<SUB = true> does not wait a full L2 round trip after the last DMMA.
template <int NB>
__device__ __forceinline__ void prefetch_tile(const cd* C, int ldc, int mrows, int ncols, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        const int i = min(8 * m + g, mrows - 1);
        const int j = min(t * (2 * NB), ncols - 1);        // one 128-byte line = 8 complex
        asm volatile("prefetch.global.L1 [%0];" ::"l"(C + (int64_t)i * ldc + j));
    }
}

// SUB = false: C = P ; SUB = true: C -= P   (P = accumulated product)
template <int MB, int NB, bool SUB>
__device__ __forceinline__ void store_tile(const Acc<MB, NB>& acc, cd* C, int ldc, int mrows,
                                           int ncols, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int m = 0; m < MB; ++m) {
        int i = 8 * m + g;
        if (i >= mrows) continue;
#pragma unroll
        for (int n = 0; n < NB; ++n) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                int j = 8 * n + 2 * t + e;
                if (j >= ncols) continue;
                cd* p = C + (int64_t)i * ldc + j;
                if (SUB) {
                    cd v = *p;
                    v.x -= acc.re[m][n][e];
                    v.y -= acc.im[m][n][e];
                    *p = v;
                } else {
                    *p = make_double2(acc.re[m][n][e], acc.im[m][n][e]);
                }
            }
        }
    }
}

// Epilogue of a contribution-block tile: C(i, j) = sum_children child(inv_c[i], inv_c[j]) - P(i, j)
// with P the accumulated product Lt * U.  (rc, cc) is the tile's position inside the block;
// every element of the block is written exactly once and never read here.
struct ChildInfo {      // what the gather needs of a child, staged once per CTA in shared memory
    int64_t cb;         // slot of the child's contribution block inside the system's front storage
    int ld;
    int invoff;
};
constexpr int MAXCH = 8;

template <int MB, int NB>
__device__ __forceinline__ void store_tile_cb(const Acc<MB, NB>& acc, cd* C, int ldc, int mrows, int ncols,
                                              int rc, int cc, const SymDev& sd, const FrontDesc& fd,
                                              const ChildInfo* chs, const cd* Fsys, int lane) {
    const int g = lane >> 2, t = lane & 3;
    cd val[MB][NB][2];
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e) val[m][n][e] = make_double2(-acc.re[m][n][e], -acc.im[m][n][e]);
    for (int ci = 0; ci < fd.nchild; ++ci) {
        ChildInfo ch;
        if (fd.nchild <= MAXCH) {
            ch = chs[ci];
        } else {
            const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
            ch.cb = fc.off + (int64_t)fc.k * fc.ld + fc.k;
            ch.ld = fc.ld;
            ch.invoff = fc.invoff;
        }
        const int* inv = sd.cbinv + ch.invoff;
        const cd* Fc = Fsys + ch.cb;
        int ri[MB], cj[NB][2];
#pragma unroll
        for (int m = 0; m < MB; ++m) ri[m] = 8 * m + g < mrows ? __ldg(inv + rc + 8 * m + g) : -1;
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e)
                cj[n][e] = 8 * n + 2 * t + e < ncols ? __ldg(inv + cc + 8 * n + 2 * t + e) : -1;
#pragma unroll
        for (int m = 0; m < MB; ++m) {
            if (ri[m] < 0) continue;
            const cd* row = Fc + (int64_t)ri[m] * ch.ld;
#pragma unroll
            for (int n = 0; n < NB; ++n)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    if (cj[n][e] >= 0) {
                        const cd a = row[cj[n][e]];
                        val[m][n][e].x += a.x;
                        val[m][n][e].y += a.y;
                    }
        }
    }
#pragma unroll
    for (int m = 0; m < MB; ++m) {
        const int i = 8 * m + g;
        if (i >= mrows) continue;
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int j = 8 * n + 2 * t + e;
                if (j < ncols) C[(int64_t)i * ldc + j] = val[m][n][e];
            }
    }
}

// C -= P for the 16 x 32 warp tile of the CTA-level GEMM: one 8 x 8 accumulator block at a time
// (a compiler barrier between blocks keeps ptxas from hoisting all 16 read-modify-write loads
// above the first store, which at 128 registers spilled the accumulators)
template <int MB, int NB>
__device__ __forceinline__ void store_tile_sub_blockwise(const Acc<MB, NB>& acc, cd* C, int ldc, int mrows,
                                                         int ncols, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int m = 0; m < MB; ++m) {
        const int i = 8 * m + g;
#pragma unroll
        for (int n = 0; n < NB; ++n) {
            const int j = 8 * n + 2 * t;
            if (i < mrows && j < ncols) {
                cd* p = C + (int64_t)i * ldc + j;
                if (j + 1 < ncols) {
                    cd v0 = p[0], v1 = p[1];
                    v0.x -= acc.re[m][n][0]; v0.y -= acc.im[m][n][0];
                    v1.x -= acc.re[m][n][1]; v1.y -= acc.im[m][n][1];
                    p[0] = v0; p[1] = v1;
                } else {
                    cd v0 = p[0];
                    v0.x -= acc.re[m][n][0]; v0.y -= acc.im[m][n][0];
                    p[0] = v0;
                }
            }
            asm volatile("" ::: "memory");
        }
    }
}

// The same epilogue without a second register tile (the 16 x 32 warp tile of the CTA-level
// GEMM below leaves no room for one): the accumulator is negated in place, the children's
// entries are added into it, and it is stored.
template <int MB, int NB>
__device__ __forceinline__ void store_tile_cb_inplace(Acc<MB, NB>& acc, cd* C, int ldc, int mrows, int ncols,
                                                      int rc, int cc, const SymDev& sd, const FrontDesc& fd,
                                                      const ChildInfo* chs, const cd* Fsys, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int m = 0; m < MB; ++m)
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e) { acc.re[m][n][e] = -acc.re[m][n][e]; acc.im[m][n][e] = -acc.im[m][n][e]; }
    for (int ci = 0; ci < fd.nchild; ++ci) {
        ChildInfo ch;
        if (fd.nchild <= MAXCH) {
            ch = chs[ci];
        } else {
            const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
            ch.cb = fc.off + (int64_t)fc.k * fc.ld + fc.k;
            ch.ld = fc.ld;
            ch.invoff = fc.invoff;
        }
        const int* inv = sd.cbinv + ch.invoff;
        const cd* Fc = Fsys + ch.cb;
        int ri[MB];
#pragma unroll
        for (int m = 0; m < MB; ++m) ri[m] = 8 * m + g < mrows ? __ldg(inv + rc + 8 * m + g) : -1;
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int cj = 8 * n + 2 * t + e < ncols ? __ldg(inv + cc + 8 * n + 2 * t + e) : -1;
                if (cj < 0) continue;
#pragma unroll
                for (int m = 0; m < MB; ++m)
                    if (ri[m] >= 0) {
                        const cd a = Fc[(int64_t)ri[m] * ch.ld + cj];
                        acc.re[m][n][e] += a.x;
                        acc.im[m][n][e] += a.y;
                    }
            }
    }
#pragma unroll
    for (int m = 0; m < MB; ++m) {
        const int i = 8 * m + g;
        if (i >= mrows) continue;
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int j = 8 * n + 2 * t + e;
                if (j < ncols) C[(int64_t)i * ldc + j] = make_double2(acc.re[m][n][e], acc.im[m][n][e]);
            }
    }
}

// ------------------------------------------- CTA-level GEMM fed by bulk asynchronous copies
// C[m x n] (-)= A[m x K] * B[K x n] for the large fronts, all operands row-major in global
// memory.  The per-warp loops above read every operand element once per warp tile from
// L1/L2 (4 flop per byte of L1 traffic); here a 256-thread CTA owns a 64 x 64 tile of C, the
// 64 x 16 panel of A and the 16 x 64 panel of B of a K chunk are brought into shared memory
// ONCE by the TMA engine (cp.async.bulk, one contiguous row segment per copy, completion on an
// mbarrier; 16 flop per byte of L2 traffic) while the previous chunk is being multiplied, and
// the DMMA fragments come from shared memory at ~30 cycles instead of an L2 round trip.
// Two stages; one __syncthreads per chunk hands the consumed stage back to the producer
// (warp 0).  Shared leading dimensions (20 and 66 complex) make both fragment loads
// bank-conflict free.  Warp w computes the 16 x 32 sub-tile (w & 3, w >> 2).
constexpr int GB_M = 64, GB_N = 64, GB_K = 16;
constexpr int GB_LDA = GB_K + 4, GB_LDB = GB_N + 2;
constexpr int GB_STAGE = GB_M * GB_LDA + GB_K * GB_LDB;                  // complex elements per stage
constexpr int GB_SMEM = 2 * GB_STAGE * (int)sizeof(cd) + 2 * (int)sizeof(uint64_t);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct GemmPipe {
    cd* buf;            // 2 stages of GB_STAGE complex elements
    uint64_t* full;     // 2 mbarriers (TMA completion per stage)
    uint32_t it;        // chunks issued so far by this CTA (stage = it & 1, parity = (it >> 1) & 1)
};

// epilogue functor: epi(acc, r0, c0, mr, nc, lane) with (r0, c0) the position of the warp's 16 x 32
// sub-tile inside C and mr x nc its valid extent
template <class Epi>
__device__ __noinline__ void cta_gemm(GemmPipe& gp, const cd* A, int lda, int m, const cd* B, int ldb, int n,
                                         int K, int tid, Epi epi) {
    const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int wr = warp & 3, wc = warp >> 2;
    const int ntr = (m + GB_M - 1) / GB_M, ntc = (n + GB_N - 1) / GB_N, nch = (K + GB_K - 1) / GB_K;
    const int NI = ntr * ntc * nch;
    // global memory written through the generic proxy (the strips of the previous phases) is about
    // to be read through the async proxy
    fence_proxy_async();
    __syncthreads();
    // One row segment per bulk copy.  UBLKCP is a warp-level instruction with uniform operands:
    // issued from a lane-strided loop it is serialised lane by lane (80 copies through one warp
    // cost more than the chunk's DMMAs -- measured, the first version was 1.4x slower than the
    // per-warp loops); here every warp issues its share (rows warp, warp + 8, ...) from one elected
    // lane with warp-uniform addresses.
    auto issue = [&](int item, uint32_t it) {
        const int tile = item / nch, ch = item - tile * nch;
        const int R0 = GB_M * (tile % ntr), C0 = GB_N * (tile / ntr), k0 = GB_K * ch;
        const int kv = min(GB_K, K - k0), nv = min(GB_N, n - C0);
        cd* As = gp.buf + (it & 1) * GB_STAGE;
        cd* Bs = As + GB_M * GB_LDA;
        uint64_t* bar = gp.full + (it & 1);
        if (tid == 0) mbar_expect_tx(bar, (uint32_t)((GB_M * kv + kv * nv) * sizeof(cd)));
        if (lane == 0) {
            for (int r = warp; r < GB_M; r += 8)       // rows past the edge repeat the last valid one
                bulk_g2s(As + r * GB_LDA, A + (int64_t)min(R0 + r, m - 1) * lda + k0, (uint32_t)(kv * sizeof(cd)), bar);
            for (int r = warp; r < kv; r += 8)
                bulk_g2s(Bs + r * GB_LDB, B + (int64_t)(k0 + r) * ldb + C0, (uint32_t)(nv * sizeof(cd)), bar);
        }
    };
    if (NI > 0) issue(0, gp.it);
    Acc<2, 4> acc;
    for (int item = 0; item < NI; ++item) {
        const uint32_t it = gp.it + item;
        if (item + 1 < NI) issue(item + 1, it + 1);
        const int tile = item / nch, ch = item - tile * nch;
        if (ch == 0) acc.zero();
        mbar_wait(gp.full + (it & 1), (it >> 1) & 1);
        const cd* As = gp.buf + (it & 1) * GB_STAGE + (16 * wr + g) * GB_LDA + t;
        const cd* Bs = gp.buf + (it & 1) * GB_STAGE + GB_M * GB_LDA + t * GB_LDB + 32 * wc + g;
        const int kv = min(GB_K, K - GB_K * ch);
        const int nfull = kv >> 2, rem = kv & 3;
#pragma unroll
        for (int ks = 0; ks < GB_K / 4; ++ks) {
            if (ks < nfull) {
                cd a[2], b[4];
#pragma unroll
                for (int mi = 0; mi < 2; ++mi) a[mi] = As[8 * mi * GB_LDA + 4 * ks];
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) b[ni] = Bs[4 * ks * GB_LDB + 8 * ni];
                warp_mma_step<2, 4>(acc, a, b);
            }
        }
        if (rem) {      // K tail: both operands masked (stale shared memory beyond the copied part)
            cd a[2], b[4];
            const cd zero = make_double2(0.0, 0.0);
#pragma unroll
            for (int mi = 0; mi < 2; ++mi) a[mi] = t < rem ? As[8 * mi * GB_LDA + 4 * nfull] : zero;
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) b[ni] = t < rem ? Bs[4 * nfull * GB_LDB + 8 * ni] : zero;
            warp_mma_step<2, 4>(acc, a, b);
        }
        if (ch == nch - 1) {
            const int r0 = GB_M * (tile % ntr) + 16 * wr, c0 = GB_N * (tile / ntr) + 32 * wc;
            const int mr = min(16, m - r0), nc = min(32, n - c0);
            if (mr > 0 && nc > 0) epi(acc, r0, c0, mr, nc, lane);
        }
        __syncthreads();      // the stage just read may be refilled (by the issue of item + 2)
    }
    gp.it += NI;
}

// e / d for 0 <= e < 2^24, d >= 1: float reciprocal estimate, corrected to the exact quotient
// (a few instructions instead of the ~20 of an integer division)
__device__ __forceinline__ int fast_div(int e, int d, float inv_d) {
    int q = __float2int_rz((float)e * inv_d);
    if (q * d > e) --q;
    if ((q + 1) * d <= e) ++q;
    return q;
}

// --------------------------------------------------------------- pivot-tile inversion
constexpr int TW = 32;        // pivot tile width
constexpr int TLD = TW + 1;   // shared-memory leading dimension

struct TileSmem {
    cd D[2][TW][TLD];
    int piv[TW];
    int src[TW];
    int sing;
    int swapped;
};

// pivot row of column p among rows p..n-1 of the tile T (leading dimension ldt): argmax of
// max(|re|, |im|), compared on the upper 27 bits of the double (sign-free exponent + 16
// mantissa bits: ample for a pivot choice) so that one integer warp reduction does the
// search; ties go to the lowest row.  Returns the key (0 in its upper bits: no usable pivot).
__device__ __forceinline__ unsigned pivot_search(const cd* T, int ldt, int n, int p, int lane) {
    unsigned key = 0u;
    if (lane >= p && lane < n) {
        const cd v = T[lane * ldt + p];
        const double m = fmax(fabs(v.x), fabs(v.y));
        const unsigned hi = (unsigned)__double2hiint(m);
        if (m == m && hi < 0x7ff00000u) key = ((hi >> 4) << 5) | (unsigned)(31 - lane);
    }
    return __reduce_max_sync(0xffffffffu, key);
}
// 1/z = conj(z) / |z|^2 after scaling z by 2^-e (e = exponent carried by the search key) to
// dodge over/underflow: one division
__device__ __forceinline__ cd pivot_reciprocal(cd z, unsigned key) {
    const int ex = ((int)(key >> 21)) - 1023;
    const double sc = __hiloint2double((1023 - ex) << 20, 0);   // 2^-ex, exact
    const double zx = z.x * sc, zy = z.y * sc;
    // |z|^2 of the scaled pivot lies in [1, 8): the hardware reciprocal seed plus two Newton
    // steps is accurate to the last bit or two and a third of the dependent chain of a division
    const double n2 = fma(zx, zx, zy * zy);
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(n2));
    r = fma(fma(-n2, r, 1.0), r, r);
    r = fma(fma(-n2, r, 1.0), r, r);
    const double q = sc * r;
    return make_double2(zx * q, -zy * q);
}

// Column source map of the row exchanges: A^-1 = B P, i.e. column j of the result is column
// src[j] of B, with src = the transpositions (p, piv[p]) applied to the identity from the last
// pivot to the first.  Every warp derives it redundantly in registers (lane j ends up with
// src[j]; three shuffles per pivot), so nobody waits on a serial loop of one thread.
__device__ __forceinline__ int pivot_column_source(const int* piv, int n, int lane) {
    const int mine = lane < n ? piv[lane] : lane;
    int src = lane;
    for (int p = n - 1; p >= 0; --p) {
        const int r = __shfl_sync(0xffffffffu, mine, p);
        const int vr = __shfl_sync(0xffffffffu, src, r);
        const int vp = __shfl_sync(0xffffffffu, src, p);
        if (lane == p) src = vr;
        else if (lane == r) src = vp;
    }
    return src;
}

// Gauss-Jordan inversion of the n x n block in ts.D[0] with partial pivoting restricted to
// the block (row exchanges only move rows inside the pivot tile, so the static front
// structure is unaffected: D^-1 absorbs the permutation).  Double buffered: step p reads
// D[cur] and writes D[cur^1], so one barrier per step suffices; every warp finds the pivot
// row redundantly.  A thread owns column j = lane of the rows its warp strides over, so the
// normalised pivot-row element is formed once per step and an element update is two shared
// loads, four FMAs and a store.  Returns the buffer index holding D^-1.  A zero /
// non-finite pivot sets ts.sing.  All NT threads of the CTA take part.
template <int NT>
__device__ __forceinline__ int invert_tile(TileSmem& ts, int n, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    constexpr int R = TW / NW;        // rows of a thread: warp, warp + NW, ... (fixed over the pivots)
    int roff[R];
#pragma unroll
    for (int u = 0; u < R; ++u) roff[u] = min(warp + u * NW, n - 1) * TLD;
    int cur = 0;
    for (int p = 0; p < n; ++p) {
        const cd* Dc = &ts.D[cur][0][0];
        cd* Dn = &ts.D[cur ^ 1][0][0];
        const unsigned key = pivot_search(Dc, TLD, n, p, lane);
        const bool ok = (key >> 5) != 0u;
        const int r = ok ? 31 - (int)(key & 31u) : p;
        if (tid == 0) {
            ts.piv[p] = r;
            if (!ok) ts.sing = 1;
            if (r != p) ts.swapped = 1;
        }
        const cd d = ok ? pivot_reciprocal(Dc[r * TLD + p], key) : make_double2(1.0, 0.0);
        if (lane < n) {
            const cd rb = (lane == p) ? d : cmul(Dc[r * TLD + lane], d);
            // four rows per trip: the eight shared loads go out together and the four FMA
            // chains interleave (a warp's progress here is bounded by dependent-issue latency);
            // the row offsets are per-thread constants, so the loop carries no index arithmetic
            const cd* Dl = Dc + lane;
            const cd* Dp = Dc + p;
            cd* Dnl = Dn + lane;
#pragma unroll
            for (int ub = 0; ub < R; ub += 4) {
                if (warp + ub * NW < n) {
                    cd f[4], a[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (ub + u < R) {
                            f[u] = Dp[roff[ub + u]];
                            a[u] = Dl[roff[ub + u]];
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (ub + u < R) {
                            cd out = a[u];
                            if (lane == p) out = make_double2(0.0, 0.0);
                            out.x = fma(-f[u].x, rb.x, out.x); out.x = fma(f[u].y, rb.y, out.x);
                            out.y = fma(-f[u].x, rb.y, out.y); out.y = fma(-f[u].y, rb.x, out.y);
                            if (warp + (ub + u) * NW < n) Dnl[roff[ub + u]] = out;
                        }
                    }
                }
            }
            // rows p and r trade places: row r receives the update of the OLD row p, row p the
            // normalised pivot row; both are written after (and by the warps of) the loop's
            // meaningless updates of those rows, which keeps the loop free of row selects
            if (r != p && warp == r % NW) {
                const cd f = Dp[p * TLD];
                cd out = Dl[p * TLD];
                if (lane == p) out = make_double2(0.0, 0.0);
                out.x = fma(-f.x, rb.x, out.x); out.x = fma(f.y, rb.y, out.x);
                out.y = fma(-f.x, rb.y, out.y); out.y = fma(-f.y, rb.x, out.y);
                Dn[r * TLD + lane] = out;
            }
            if (warp == p % NW) Dn[p * TLD + lane] = rb;
        }
        __syncthreads();
        cur ^= 1;
    }
    // undo the row exchanges: A^-1 = B P  <=>  column j of the result is column src[j] of B
    if (ts.swapped) {      // uniform: last written before the final barrier of the loop
        const int src = pivot_column_source(ts.piv, n, lane);
        if (lane < n)
            for (int i = warp; i < n; i += NW) ts.D[cur ^ 1][i][lane] = ts.D[cur][i][src];
        __syncthreads();
        if (tid == 0) ts.swapped = 0;
        cur ^= 1;
    }
    return cur;
}

// Compile-time option (off): measured, the polish costs 14 % of the factor phase at N = 2e4, and
// with the equilibration of the system (diag_scale_kernel) the plain Gauss-Jordan inverse is
// already good enough on both sides, see profiles/r2_summary.md.
#ifndef JPB_NEWTON
#define JPB_NEWTON 0
#endif
// One Newton-Schulz step X <- X + X (I - D X) on the inverted pivot tile.  Gauss-Jordan with
// partial pivoting leaves a good LEFT residual (X D - I, what Lt = F21 X and the transposed
// solves see) but a RIGHT residual (D X - I, what the application v_J = X t of the plain solves
// sees) ~40x above LAPACK's getri -- measured with a numpy replay of the schedule
// (profiles/r2_summary.md): that alone made the unrefined residual of a solve 50-100x the
// stand-in LU's and forced an unconditional refinement step (two more sweeps over L and U per
// solve).  The step squares both residuals; it costs 2 n^3 of the tile, ~10 % of the inversion.
// X: n x n inverse (leading dimension ldx), W: n x n scratch holding the ORIGINAL tile D on
// entry (leading dimension ldw, destroyed).  Both in shared memory.  A warp owns whole rows, so
// both products are formed in place; the caller synchronises the CTA before and after.
template <int NW>
__device__ __forceinline__ void newton_polish_tile(cd* X, int ldx, cd* W, int ldw, int n, int warp, int lane) {
    const int j = min(lane, n - 1);
    // W <- I - D X   (row i of W depends on row i of D only)
    for (int i = warp; i < n; i += NW) {
        cd acc = make_double2(i == j ? 1.0 : 0.0, 0.0);
#pragma unroll 4
        for (int p = 0; p < n; ++p) {
            const cd a = W[i * ldw + p], b = X[p * ldx + j];
            acc.x = fma(-a.x, b.x, acc.x); acc.x = fma(a.y, b.y, acc.x);
            acc.y = fma(-a.x, b.y, acc.y); acc.y = fma(-a.y, b.x, acc.y);
        }
        __syncwarp();
        if (lane < n) W[i * ldw + lane] = acc;
    }
    __syncthreads();
    // X <- X + X W   (row i of the result depends on row i of X only)
    for (int i = warp; i < n; i += NW) {
        cd acc = X[i * ldx + j];
#pragma unroll 4
        for (int p = 0; p < n; ++p) cfma(acc, X[i * ldx + p], W[p * ldw + j]);
        __syncwarp();
        if (lane < n) X[i * ldx + lane] = acc;
    }
}

// ------------------------------------------------------------------- numeric factorisation
constexpr int KMAX = 32;   // component matrices supported by the fused assembly

// One CTA per (front of this level, system): assemble the front (A entries + children's
// Schur complements), factor its k pivots, form the Schur complement of the rest.
//   FUSED = true : A entries are sum_k coef[b,k] * mats_f[k, q]  (mats_f in front-entry order),
//                  also stored to vals[b, q] for the residual kernels
//   FUSED = false: A entries are data[b, ent_src[q]]
#ifndef JPB_WN
#define JPB_WN 2     // measured: 16 x 16 warp tiles (no spills) beat 16 x 32 at cfg1 and tie at cfg4
#endif
#ifndef JPB_OCC
#define JPB_OCC 512
#endif
#ifndef JPB_STRIP_WN
#define JPB_STRIP_WN 2
#endif
//   TMA = true (NT = 256 only): the strip updates of every 64-column super tile and the Schur
//                  complement run through cta_gemm (operand panels staged in dynamic shared
//                  memory by bulk asynchronous copies); launch with GB_SMEM dynamic bytes
template <bool ENT_ORDER, int NT, bool TMA = false, bool SCHUR = true>
__global__ void __launch_bounds__(NT, JPB_OCC / NT)
factor_level_kernel(SymDev sd, const int* __restrict__ level_fronts, int nsys, cd* Fall,
                    const cd* __restrict__ data, int* status) {
    static_assert(!TMA || NT == 256, "the CTA-level GEMM is written for 8 warps");
    __shared__ TileSmem ts;
    extern __shared__ __align__(16) unsigned char gemm_smem[];
    GemmPipe gp{nullptr, nullptr, 0u};
    if (TMA) {
        gp.buf = reinterpret_cast<cd*>(gemm_smem);
        gp.full = reinterpret_cast<uint64_t*>(gemm_smem + 2 * GB_STAGE * sizeof(cd));
        if (threadIdx.x == 0) {
            mbar_init(gp.full, 1);
            mbar_init(gp.full + 1, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        // the first __syncthreads below (after the zero fill) publishes the initialised barriers
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    // warp tile: 16 rows x WC columns.  JPB_WN = 0: 16 x 32 for the 256-thread class (operand reuse
    // on the large fronts), 16 x 16 for the narrower ones (no register spills, more tiles per warp)
    constexpr int WN = JPB_WN ? JPB_WN : (NT >= 256 ? 4 : 2), WC = 8 * WN;
    constexpr int FACTOR_THREADS = NT;
    const int b = blockIdx.x % nsys;
    const FrontDesc fd = sd.fr[level_fronts[blockIdx.x / nsys]];
    cd* Fsys = Fall + (int64_t)b * sd.front_slots;
    cd* F = Fsys + fd.off;
    const int s = fd.s, k = fd.k, ld = fd.ld;

    __shared__ ChildInfo chs[MAXCH];
    if (tid == 0) { ts.sing = 0; ts.swapped = 0; }
    if (tid < fd.nchild && tid < MAXCH) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + tid]];
        chs[tid].cb = fc.off + (int64_t)fc.k * fc.ld + fc.k;
        chs[tid].ld = fc.ld;
        chs[tid].invoff = fc.invoff;
    }
    // ---- zero, then A entries (every entry owns its slot), then the children one by one
    // Only the pivot rows and pivot columns are assembled by zero + scatter; the contribution
    // block (rows and columns >= k) is written exactly once, by the Schur phase, which gathers
    // the children's blocks through the inverse maps while it stores (store_tile_cb).
    {
        const cd zero = make_double2(0.0, 0.0);
        for (int i = warp; i < s; i += NW) {
            cd* row = F + (int64_t)i * ld;
            const int lim = i < k ? s : k;
            for (int j = lane; j < lim; j += 32) row[j] = zero;
        }
    }
    __syncthreads();
    // A entries: ENT_ORDER = true : data[b, q] already in front-entry order (written by the
    // assembly kernel K1, coalesced here); false: data[b, ent_src[q]] in the caller's CSC order
    {
        const cd* db = data + (int64_t)b * sd.nnz;
        for (int q = tid; q < fd.nent; q += FACTOR_THREADS) {
            const int64_t gq = fd.entptr + q;
            F[sd.ent_slot[gq]] = ENT_ORDER ? db[gq] : db[sd.ent_src[gq]];
        }
    }
    __syncthreads();
    for (int ci = 0; ci < fd.nchild; ++ci) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
        const int* rel = sd.relmap + fc.reloff;
        const int mc = fc.s - fc.k, pc = fc.npan;
        const cd* Fc = Fsys + fc.off + (int64_t)fc.k * fc.ld + fc.k;
        // the part of the child's block that lands on pivot rows or pivot columns: its first pc
        // rows in full, then the first pc columns of the other rows.  Flattened over all
        // threads, eight independent elements in flight per thread.
        const int top = pc * mc, tot = top + (mc - pc) * pc;
        const float inv_mc = 1.0f / (float)mc, inv_pc = 1.0f / (float)max(pc, 1);
        constexpr int MLP = 8;
        for (int e0 = tid; e0 < tot; e0 += MLP * FACTOR_THREADS) {
            cd v[MLP];
            cd* p[MLP];
#pragma unroll
            for (int q = 0; q < MLP; ++q) {
                const int e = e0 + q * FACTOR_THREADS;
                if (e < tot) {
                    int i, j;
                    if (e < top) { i = fast_div(e, mc, inv_mc); j = e - i * mc; }
                    else { const int e2 = e - top; i = fast_div(e2, pc, inv_pc); j = e2 - i * pc; i += pc; }
                    v[q] = Fc[(int64_t)i * fc.ld + j];
                    p[q] = F + (int64_t)rel[i] * ld + rel[j];
                } else {
                    p[q] = nullptr;
                }
            }
            cd o[MLP];
#pragma unroll
            for (int q = 0; q < MLP; ++q)
                if (p[q]) o[q] = *p[q];
#pragma unroll
            for (int q = 0; q < MLP; ++q)
                if (p[q]) *p[q] = make_double2(o[q].x + v[q].x, o[q].y + v[q].y);
        }
        __syncthreads();
    }

    // ---- left-looking block factorisation of the pivot tiles
    const int KT = (k + TW - 1) / TW;
    for (int J = 0; J < KT; ++J) {
        const int j0 = J * TW;
        const int jw = min(TW, k - j0);
        const int j1 = j0 + jw;
        // ks0: first column of the pivot range the warp-level strip update below still has to
        // apply.  TMA: the range [0, j0s) of the 64-column super tile [j0s, j1s) this pivot tile
        // belongs to is applied to both tiles at once by the CTA-level GEMM when the super tile
        // opens; only the other tile of the super tile (32 columns) is left for the warp loops.
        int ks0 = 0;
        if (TMA) {
            const int j0s = j0 & ~(GB_N - 1);
            ks0 = j0s;
            if (j0 == j0s && j0s > 0) {
                const int j1s = min(j0s + GB_N, k);
                auto sub = [&](cd* C) {
                    return [=](Acc<2, 4>& acc, int r0, int c0, int mr, int nc, int ln) {
                        store_tile_sub_blockwise<2, 4>(acc, C + (int64_t)r0 * ld + c0, ld, mr, nc, ln);
                    };
                };
                // column strip rows [j0s, s) x cols [j0s, j1s), then row strip rows [j0s, j1s) x cols [j1s, s)
                cta_gemm(gp, F + (int64_t)j0s * ld, ld, s - j0s, F + j0s, ld, j1s - j0s, j0s, tid,
                         sub(F + (int64_t)j0s * ld + j0s));
                if (s > j1s)
                    cta_gemm(gp, F + (int64_t)j0s * ld, ld, j1s - j0s, F + j1s, ld, s - j1s, j0s, tid,
                             sub(F + (int64_t)j0s * ld + j1s));
            }
        }
        if (j0 > ks0) {
            // column strip rows [j0,s) x cols [j0,j1) and row strip rows [j0,j1) x cols [j1,s)
            // warp tiles of 16 x 16 here (SN = 2 accumulator blocks across): with the live state
            // of the tile loop around it the 16 x 32 tile made ptxas spill inside the k-loop
            constexpr int SN = JPB_STRIP_WN, SC = 8 * SN, CT = TW / SC;
            const int nct = ((s - j0 + 15) >> 4) * CT;
            const int nrr = (jw + 15) >> 4;
            const int nrc = (s - j1 + SC - 1) / SC;
            const int ntiles = nct + nrr * nrc;
            for (int tile = warp; tile < ntiles; tile += NW) {
                int r, c, mr, nc;
                if (tile < nct) {
                    const int rt = tile / CT, ct = tile % CT;
                    r = j0 + 16 * rt; c = j0 + SC * ct; mr = min(16, s - r); nc = min(SC, j1 - c);
                } else {
                    int q = tile - nct;
                    r = j0 + 16 * (q % nrr); c = j1 + SC * (q / nrr);
                    mr = min(16, j1 - r); nc = min(SC, s - c);
                }
                if (nc <= 0) continue;
                Acc<2, SN> acc;
                acc.zero();
                prefetch_tile<SN>(F + (int64_t)r * ld + c, ld, mr, nc, lane);
                warp_gemm<2, SN>(acc, F + (int64_t)r * ld + ks0, ld, mr, F + (int64_t)ks0 * ld + c, ld, nc,
                                 j0 - ks0, lane);
                store_tile<2, SN, true>(acc, F + (int64_t)r * ld + c, ld, mr, nc, lane);
            }
            __syncthreads();
        }
        for (int e = tid; e < jw * TW; e += FACTOR_THREADS) {
            int i = e >> 5, j = e & 31;
            if (j < jw) ts.D[0][i][j] = F[(int64_t)(j0 + i) * ld + j0 + j];
        }
        __syncthreads();
        const int cur = invert_tile<NT>(ts, jw, tid);
        if (JPB_NEWTON && !ts.sing) {      // uniform; the front still holds the original tile (written back below)
            for (int e = tid; e < jw * TW; e += FACTOR_THREADS) {
                int i = e >> 5, j = e & 31;
                if (j < jw) ts.D[cur ^ 1][i][j] = F[(int64_t)(j0 + i) * ld + j0 + j];
            }
            __syncthreads();
            newton_polish_tile<NW>(&ts.D[cur][0][0], TLD, &ts.D[cur ^ 1][0][0], TLD, jw, warp, lane);
            __syncthreads();
        }
        const cd (*Dinv)[TLD] = ts.D[cur];
        for (int e = tid; e < jw * TW; e += FACTOR_THREADS) {
            int i = e >> 5, j = e & 31;
            if (j < jw) F[(int64_t)(j0 + i) * ld + j0 + j] = Dinv[i][j];
        }
        // Lt strip: rows [j1,s) x cols [j0,j1)  <-  F[rows, j0:j1] * D^-1
        // (a warp owns whole rows of the strip: it reads them all before it stores any)
        const int nlt = (s - j1 + 15) >> 4;
        for (int tile = warp; tile < nlt; tile += NW) {
            int r = j1 + 16 * tile, mr = min(16, s - r);
            Acc<2, WN> acc[TW / WC];
#pragma unroll
            for (int h = 0; h < TW / WC; ++h) {
                acc[h].zero();
                if (WC * h < jw)
                    warp_gemm<2, WN>(acc[h], F + (int64_t)r * ld + j0, ld, mr, &Dinv[0][WC * h], TLD,
                                     jw - WC * h, jw, lane);
            }
#pragma unroll
            for (int h = 0; h < TW / WC; ++h)
                if (WC * h < jw)
                    store_tile<2, WN, false>(acc[h], F + (int64_t)r * ld + j0 + WC * h, ld, mr, jw - WC * h, lane);
        }
        __syncthreads();
    }
    // ---- contribution block: C = sum over children of their blocks (gathered) - Lt * U
    // (SCHUR = false: left to schur_band_kernel, which spreads it over several CTAs per front)
    if constexpr (SCHUR) {
    const int m = s - k;
    if (TMA && m > 0) {
        cd* C = F + (int64_t)k * ld + k;
        cta_gemm(gp, F + (int64_t)k * ld, ld, m, F + k, ld, m, k, tid,
                 [&](Acc<2, 4>& acc, int r0, int c0, int mr, int nc, int ln) {
                     store_tile_cb_inplace<2, 4>(acc, C + (int64_t)r0 * ld + c0, ld, mr, nc, r0, c0, sd, fd, chs,
                                                 Fsys, ln);
                 });
    } else if (m > 0) {
        const int ntr = (m + 15) >> 4, ntc = (m + WC - 1) / WC;
        for (int tile = warp; tile < ntr * ntc; tile += NW) {
            const int rc = 16 * (tile % ntr), cc = WC * (tile / ntr);      // within the block
            const int r = k + rc, c = k + cc;
            const int mr = min(16, s - r), nc = min(WC, s - c);
            Acc<2, WN> acc;
            acc.zero();
            warp_gemm<2, WN>(acc, F + (int64_t)r * ld, ld, mr, F + c, ld, nc, k, lane);
            store_tile_cb<2, WN>(acc, F + (int64_t)r * ld + c, ld, mr, nc, rc, cc, sd, fd, chs, Fsys, lane);
        }
    }
    }
    if (tid == 0 && ts.sing && status) atomicOr(status + b, 1);
}

// Contribution blocks of large fronts as a launch of their own: C = gathered children - Lt * U on
// `nbands` CTAs per (front, system), each taking a band of SCHUR_BAND columns.  Why: with one CTA per
// front, 296 fronts of ~10 MB are in flight and the m x k panel of Lt that every 32-column block of C
// re-reads comes from DRAM each time (cfg4: 3-4x the compulsory traffic, the launches run at 3.6 TB/s).
// The CTAs of one front sit next to each other in the grid, run at the same time and walk Lt in the
// same order, so the panel is fetched from DRAM once and served to the other bands by L2.
constexpr int SCHUR_BAND = 128;
template <int NT>
__global__ void __launch_bounds__(NT, JPB_OCC / NT)
schur_band_kernel(SymDev sd, const int* __restrict__ level_fronts, int nsys, cd* Fall, int nbands) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32, WN = 4, WC = 8 * WN;
    int bid = blockIdx.x;
    const int band = bid % nbands;
    bid /= nbands;
    const int b = bid % nsys;
    const FrontDesc fd = sd.fr[level_fronts[bid / nsys]];
    const int s = fd.s, k = fd.k, ld = fd.ld, m = s - k;
    const int c0 = band * SCHUR_BAND;
    if (c0 >= m) return;
    cd* Fsys = Fall + (int64_t)b * sd.front_slots;
    cd* F = Fsys + fd.off;
    __shared__ ChildInfo chs[MAXCH];
    if (tid < fd.nchild && tid < MAXCH) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + tid]];
        chs[tid].cb = fc.off + (int64_t)fc.k * fc.ld + fc.k;
        chs[tid].ld = fc.ld;
        chs[tid].invoff = fc.invoff;
    }
    __syncthreads();
    const int bw = min(SCHUR_BAND, m - c0);
    const int ntr = (m + 15) >> 4, ntc = (bw + WC - 1) / WC;
    for (int tile = warp; tile < ntr * ntc; tile += NW) {
        const int rc = 16 * (tile % ntr), cc = c0 + WC * (tile / ntr);      // within the block
        const int r = k + rc, c = k + cc;
        const int mr = min(16, s - r), nc = min(WC, s - c);
        Acc<2, WN> acc;
        acc.zero();
        warp_gemm<2, WN>(acc, F + (int64_t)r * ld, ld, mr, F + c, ld, nc, k, lane);
        store_tile_cb<2, WN>(acc, F + (int64_t)r * ld + c, ld, mr, nc, rc, cc, sd, fd, chs, Fsys, lane);
    }
}

// ------------------------------------------------- shared-memory-resident fronts (K2s)
// Fronts whose s x s block fits the SM's shared memory never touch HBM between assembly and
// the final store: zero fill, A entries, extend-add, pivot-tile inversion, the DMMA strip /
// Lt / Schur updates all run on the shared copy, and the finished front (factors + Schur
// block) leaves with one coalesced pass.  HBM traffic = the compulsory bytes only (entries
// and child Schur blocks in, the front out) and every dependent step of the pivot chain
// costs a shared-memory round trip instead of an L2 one.
//
// Shared leading dimension: lds = 2 (mod 4), which makes the DMMA B-fragment loads (4 rows
// x 2 adjacent complex per quarter warp) bank-conflict free.
__host__ __device__ __forceinline__ int smem_ld(int s) { return s + ((2 - (s & 3)) & 3); }

struct PivotSmem {
    int piv[TW];
    int src[TW];
    int sing;
    int swapped;
};

// In-place Gauss-Jordan inversion of the n x n block at T (leading dimension ldt, shared
// memory) with partial pivoting restricted to the block; same arithmetic and thread mapping
// as invert_tile.  Before the update of step p every thread holds what other warps are about
// to overwrite (its column of the normalised pivot row, of the old row p, and the old
// T[p][p]) in registers, so a step costs two barriers and no staging buffer.  On return T
// holds the inverse (row exchanges undone).
template <int NT>
__device__ __forceinline__ void invert_tile_inplace(cd* T, int ldt, int n, int tid, PivotSmem& ps) {
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    constexpr int R = TW / NW;        // rows of a thread: warp, warp + NW, ... (fixed over the pivots)
    const int jl = min(lane, n - 1);
    int roff[R];
#pragma unroll
    for (int u = 0; u < R; ++u) roff[u] = min(warp + u * NW, n - 1) * ldt;
    for (int p = 0; p < n; ++p) {
        const unsigned key = pivot_search(T, ldt, n, p, lane);
        const bool ok = (key >> 5) != 0u;
        const int r = ok ? 31 - (int)(key & 31u) : p;
        if (tid == 0) {
            ps.piv[p] = r;
            if (!ok) ps.sing = 1;
            if (r != p) ps.swapped = 1;
        }
        const cd d = ok ? pivot_reciprocal(T[r * ldt + p], key) : make_double2(1.0, 0.0);
        const cd rb = (lane == p) ? d : cmul(T[r * ldt + jl], d);
        const cd pr = T[p * ldt + jl];
        __syncthreads();
        // Rows p and r trade places: the owner warp of row r first drops the old row p into it
        // (everybody holds what they need of the old rows p and r in registers), after which
        // every row but p takes the same update; row p is written last, by its owner warp.
        // No per-element selects on the row index and no index arithmetic are left in the loop.
        if (r != p && warp == r % NW) {
            if (lane < n) T[r * ldt + lane] = pr;
            __syncwarp();
        }
        const cd* Tp = T + p;
        cd* Tl = T + jl;
#pragma unroll
        for (int ub = 0; ub < R; ub += 4) {      // four rows per trip (see invert_tile)
            if (warp + ub * NW < n) {
                cd f[4], a[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (ub + u < R) {
                        f[u] = Tp[roff[ub + u]];
                        a[u] = Tl[roff[ub + u]];
                    }
                }
                __syncwarp();      // every lane has read T[i][p] before lane p overwrites it
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (ub + u < R) {
                        cd out = a[u];
                        if (lane == p) out = make_double2(0.0, 0.0);
                        out.x = fma(-f[u].x, rb.x, out.x); out.x = fma(f[u].y, rb.y, out.x);
                        out.y = fma(-f[u].x, rb.y, out.y); out.y = fma(-f[u].y, rb.x, out.y);
                        if (warp + (ub + u) * NW < n && lane < n) Tl[roff[ub + u]] = out;
                    }
                }
            }
        }
        if (warp == p % NW) {      // same warp that just stored row p's (meaningless) update
            __syncwarp();
            if (lane < n) T[p * ldt + lane] = rb;
        }
        __syncthreads();
    }
    if (ps.swapped) {      // uniform: last written before the final barrier of the loop
        // A^-1 = B P: column j of the result is column src[j] of B
        const int src = pivot_column_source(ps.piv, n, lane);
        for (int i = warp; i < n; i += NW) {     // a row belongs to one warp
            cd v = make_double2(0.0, 0.0);
            if (lane < n) v = T[i * ldt + src];
            __syncwarp();
            if (lane < n) T[i * ldt + lane] = v;
        }
        __syncthreads();
        if (tid == 0) ps.swapped = 0;
    }
}

// One CTA per (front of this launch group, system); dynamic shared memory holds the front
// (s x lds complex).  `data` is [nsys, nnz] in front-entry order (ENT_ORDER) or the caller's
// CSC order.
#ifndef JPB_SMEM_OCC
#define JPB_SMEM_OCC 512
#endif
template <bool ENT_ORDER, int NT>
__global__ void __launch_bounds__(NT, (NT >= 256 ? 1 : JPB_SMEM_OCC / NT))
factor_smem_kernel(SymDev sd, const int* __restrict__ level_fronts, int nsys, cd* Fall,
                   const cd* __restrict__ data, int* status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cd* Fs = reinterpret_cast<cd*>(smem_raw);
    __shared__ PivotSmem ps;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    const int b = blockIdx.x % nsys;
    const FrontDesc fd = sd.fr[level_fronts[blockIdx.x / nsys]];
    cd* Fsys = Fall + (int64_t)b * sd.front_slots;
    cd* Fg = Fsys + fd.off;
    const int s = fd.s, k = fd.k, ld = fd.ld;
    const int lds = smem_ld(s);

    if (tid == 0) { ps.sing = 0; ps.swapped = 0; }
    {
        const cd zero = make_double2(0.0, 0.0);
        const int tot = s * lds;
        for (int e = tid; e < tot; e += NT) Fs[e] = zero;
    }
    __syncthreads();
    // A entries: every entry owns its slot.  ent_slot is in the global layout (i * ld + j).
    {
        const cd* db = data + (int64_t)b * sd.nnz;
        const float inv_ld = 1.0f / (float)ld;
        constexpr int MLP = 4;
        for (int q0 = tid; q0 < fd.nent; q0 += MLP * NT) {
            cd v[MLP];
            int sl[MLP];
#pragma unroll
            for (int u = 0; u < MLP; ++u) {
                const int q = q0 + u * NT;
                if (q < fd.nent) {
                    const int64_t gq = fd.entptr + q;
                    sl[u] = __ldg(sd.ent_slot + gq);
                    v[u] = ENT_ORDER ? db[gq] : db[__ldg(sd.ent_src + gq)];
                } else {
                    sl[u] = -1;
                }
            }
#pragma unroll
            for (int u = 0; u < MLP; ++u)
                if (sl[u] >= 0) {
                    const int i = __float2int_rz(((float)sl[u] + 0.5f) * inv_ld);
                    Fs[sl[u] + i * (lds - ld)] = v[u];
                }
        }
    }
    __syncthreads();
    // extend-add: the children's Schur blocks stream from HBM, the scatter-add hits shared
    for (int ci = 0; ci < fd.nchild; ++ci) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
        const int* rel = sd.relmap + fc.reloff;
        const int mc = fc.s - fc.k;
        const cd* Fc = Fsys + fc.off + (int64_t)fc.k * fc.ld + fc.k;
        const int tot = mc * mc;
        const float inv_mc = 1.0f / (float)mc;
        constexpr int MLP = 8;
        for (int e0 = tid; e0 < tot; e0 += MLP * NT) {
            cd v[MLP];
            int dst[MLP];
#pragma unroll
            for (int q = 0; q < MLP; ++q) {
                const int e = e0 + q * NT;
                if (e < tot) {
                    const int i = __float2int_rz(((float)e + 0.5f) * inv_mc), j = e - i * mc;
                    v[q] = Fc[i * fc.ld + j];
                    dst[q] = __ldg(rel + i) * lds + __ldg(rel + j);
                } else {
                    dst[q] = -1;
                }
            }
#pragma unroll
            for (int q = 0; q < MLP; ++q)
                if (dst[q] >= 0) {
                    cd o = Fs[dst[q]];
                    o.x += v[q].x; o.y += v[q].y;
                    Fs[dst[q]] = o;
                }
        }
        __syncthreads();
    }

    // left-looking block factorisation of the pivot tiles, all operands in shared memory
    const int KT = (k + TW - 1) / TW;
    for (int J = 0; J < KT; ++J) {
        const int j0 = J * TW;
        const int jw = min(TW, k - j0);
        const int j1 = j0 + jw;
        if (J > 0) {
            const int nct = (s - j0 + 15) >> 4;
            const int nrr = (jw + 15) >> 4;
            const int nrc = (s - j1 + 31) >> 5;
            const int ntiles = nct + nrr * nrc;
            for (int tile = warp; tile < ntiles; tile += NW) {
                int r, c, mr, nc;
                if (tile < nct) {
                    r = j0 + 16 * tile; c = j0; mr = min(16, s - r); nc = jw;
                } else {
                    int q = tile - nct;
                    r = j0 + 16 * (q % nrr); c = j1 + 32 * (q / nrr);
                    mr = min(16, j1 - r); nc = min(32, s - c);
                }
                Acc<2, 4> acc;
                acc.zero();
                warp_gemm_shared<2, 4>(acc, Fs + r * lds, lds, mr, Fs + c, lds, nc, j0, lane);
                store_tile<2, 4, true>(acc, Fs + r * lds + c, lds, mr, nc, lane);
            }
            __syncthreads();
        }
        cd* Wt = Fs + s * lds;      // jw x jw scratch behind the front (sized by the host per launch group)
        if (JPB_NEWTON)
            for (int e = tid; e < jw * jw; e += NT) Wt[e] = Fs[(j0 + e / jw) * lds + j0 + e % jw];
        invert_tile_inplace<NT>(Fs + j0 * lds + j0, lds, jw, tid, ps);      // starts and ends with a CTA barrier
        if (JPB_NEWTON && !ps.sing) {
            newton_polish_tile<NW>(Fs + j0 * lds + j0, lds, Wt, jw, jw, warp, lane);
            __syncthreads();
        }
        const int nlt = (s - j1 + 15) >> 4;
        for (int tile = warp; tile < nlt; tile += NW) {
            int r = j1 + 16 * tile, mr = min(16, s - r);
            Acc<2, 4> acc;
            acc.zero();
            warp_gemm_shared<2, 4>(acc, Fs + r * lds + j0, lds, mr, Fs + j0 * lds + j0, lds, jw, jw, lane);
            store_tile<2, 4, false>(acc, Fs + r * lds + j0, lds, mr, jw, lane);
        }
        __syncthreads();
    }
    const int m = s - k;
    if (m > 0 && k > 0) {
        const int ntr = (m + 15) >> 4, ntc = (m + 31) >> 5;
        for (int tile = warp; tile < ntr * ntc; tile += NW) {
            int r = k + 16 * (tile % ntr), c = k + 32 * (tile / ntr);
            int mr = min(16, s - r), nc = min(32, s - c);
            Acc<2, 4> acc;
            acc.zero();
            warp_gemm_shared<2, 4>(acc, Fs + r * lds, lds, mr, Fs + c, lds, nc, k, lane);
            store_tile<2, 4, true>(acc, Fs + r * lds + c, lds, mr, nc, lane);
        }
        __syncthreads();
    }
    // the finished front leaves in one pass: warp per row, lanes along the row
    for (int j = lane; j < s; j += 32) {
        for (int i0 = warp; i0 < s; i0 += 4 * NW) {      // four rows in flight per lane
            cd v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = Fs[min(i0 + u * NW, s - 1) * lds + j];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (i0 + u * NW < s) Fg[(int64_t)(i0 + u * NW) * ld + j] = v[u];
        }
    }
    if (tid == 0 && ps.sing && status) atomicOr(status + b, 1);
}

// ------------------------------------------- pivot panels in shared memory (K2p)
// Fronts whose PIVOT PANELS fit the SM's shared memory -- the k pivot rows (k x s: pivot block
// and U12) and the pivot columns of the other rows (m x k: L21), (2 s - k) k elements instead of
// the s^2 of factor_smem_kernel -- are factorised entirely out of shared memory: zero fill,
// A entries, the pivot part of the children's contribution blocks, the pivot-tile inversion and
// every DMMA operand (strips, Lt, Schur complement) hit the panels; the contribution block is
// never resident: its 16 x 32 tiles are formed in registers from the panels, the children's
// blocks are gathered into them through the inverse maps (as in factor_level_kernel) and each
// element is written to HBM exactly once.  Compared with factor_smem_kernel a front of order 45
// with 10 pivots needs 13 KB instead of 33 KB (2.6x the resident CTAs); compared with
// factor_level_kernel a front of order 86 with 25 pivots reads its DMMA operands at ~30 cycles
// instead of an L1/L2 round trip and never scatters into HBM.  HBM traffic = the compulsory
// bytes: entries and child blocks in, panels and contribution block out.
//
// Layout: Up[k][ldu] (ldu = smem_ld(s)) followed by Lp[m][ldl] (ldl = smem_ld(k)); both leading
// dimensions are 2 (mod 4), which keeps the DMMA B-fragment loads bank-conflict free.
__host__ __device__ __forceinline__ int panel_elems(int s, int k) { return k * smem_ld(s) + (s - k) * smem_ld(k); }

template <bool ENT_ORDER, int NT>
__global__ void __launch_bounds__(NT, 512 / NT)
factor_panel_kernel(SymDev sd, const int* __restrict__ level_fronts, int nsys, cd* Fall,
                    const cd* __restrict__ data, int* status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ PivotSmem ps;
    __shared__ ChildInfo chs[MAXCH];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NT / 32;
    const int b = blockIdx.x % nsys;
    const FrontDesc fd = sd.fr[level_fronts[blockIdx.x / nsys]];
    cd* Fsys = Fall + (int64_t)b * sd.front_slots;
    cd* Fg = Fsys + fd.off;
    const int s = fd.s, k = fd.k, ld = fd.ld, m = s - k;
    const int ldu = smem_ld(s), ldl = smem_ld(k);
    cd* Up = reinterpret_cast<cd*>(smem_raw);
    cd* Lp = Up + k * ldu;

    if (tid == 0) { ps.sing = 0; ps.swapped = 0; }
    if (tid < fd.nchild && tid < MAXCH) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + tid]];
        chs[tid].cb = fc.off + (int64_t)fc.k * fc.ld + fc.k;
        chs[tid].ld = fc.ld;
        chs[tid].invoff = fc.invoff;
    }
    {
        const cd zero = make_double2(0.0, 0.0);
        const int tot = k * ldu + m * ldl;
        for (int e = tid; e < tot; e += NT) Up[e] = zero;
    }
    __syncthreads();
    // A entries: every entry of the front has its row or its column among the pivots
    {
        const cd* db = data + (int64_t)b * sd.nnz;
        const float inv_ld = 1.0f / (float)ld;
        constexpr int MLP = 4;
        for (int q0 = tid; q0 < fd.nent; q0 += MLP * NT) {
            cd v[MLP];
            int sl[MLP];
#pragma unroll
            for (int u = 0; u < MLP; ++u) {
                const int q = q0 + u * NT;
                if (q < fd.nent) {
                    const int64_t gq = fd.entptr + q;
                    sl[u] = __ldg(sd.ent_slot + gq);
                    v[u] = ENT_ORDER ? db[gq] : db[__ldg(sd.ent_src + gq)];
                } else {
                    sl[u] = -1;
                }
            }
#pragma unroll
            for (int u = 0; u < MLP; ++u)
                if (sl[u] >= 0) {
                    const int i = __float2int_rz(((float)sl[u] + 0.5f) * inv_ld), j = sl[u] - i * ld;
                    if (i < k) Up[i * ldu + j] = v[u];
                    else Lp[(i - k) * ldl + j] = v[u];
                }
        }
    }
    __syncthreads();
    // children: the part of their contribution blocks that lands on pivot rows or pivot columns
    // (their first npan rows in full, then the first npan columns of the other rows)
    for (int ci = 0; ci < fd.nchild; ++ci) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
        const int* rel = sd.relmap + fc.reloff;
        const int mc = fc.s - fc.k, pc = fc.npan;
        const cd* Fc = Fsys + fc.off + (int64_t)fc.k * fc.ld + fc.k;
        const int top = pc * mc, tot = top + (mc - pc) * pc;
        const float inv_mc = 1.0f / (float)mc, inv_pc = 1.0f / (float)max(pc, 1);
        constexpr int MLP = 8;
        for (int e0 = tid; e0 < tot; e0 += MLP * NT) {
            cd v[MLP];
            int dst[MLP];
#pragma unroll
            for (int q = 0; q < MLP; ++q) {
                const int e = e0 + q * NT;
                if (e < tot) {
                    int i, j;
                    if (e < top) { i = fast_div(e, mc, inv_mc); j = e - i * mc; }
                    else { const int e2 = e - top; i = fast_div(e2, pc, inv_pc); j = e2 - i * pc; i += pc; }
                    v[q] = Fc[(int64_t)i * fc.ld + j];
                    const int pi = __ldg(rel + i), pj = __ldg(rel + j);
                    dst[q] = pi < k ? pi * ldu + pj : k * ldu + (pi - k) * ldl + pj;      // offset from Up
                } else {
                    dst[q] = -1;
                }
            }
#pragma unroll
            for (int q = 0; q < MLP; ++q)
                if (dst[q] >= 0) {
                    cd o = Up[dst[q]];
                    o.x += v[q].x; o.y += v[q].y;
                    Up[dst[q]] = o;
                }
        }
        __syncthreads();
    }

    // left-looking block factorisation of the pivot tiles, operands in the panels
    const int KT = (k + TW - 1) / TW;
    const int nlp = (m + 15) >> 4;                      // 16-row tiles of the L panel
    for (int J = 0; J < KT; ++J) {
        const int j0 = J * TW;
        const int jw = min(TW, k - j0);
        const int j1 = j0 + jw;
        if (J > 0) {
            // (a) rows [j0, k) x cols [j0, j1) in Up, (b) rows [k, s) x cols [j0, j1) in Lp,
            // (c) rows [j0, j1) x cols [j1, s) in Up; K = j0, B operand = Up[0:j0, cols]
            const int na = (k - j0 + 15) >> 4;
            const int nrr = (jw + 15) >> 4, nrc = (s - j1 + 31) >> 5;
            const int ntiles = na + nlp + nrr * nrc;
            for (int tile = warp; tile < ntiles; tile += NW) {
                cd* Crow;
                int ldc, c, mr, nc;
                if (tile < na) {
                    const int r = j0 + 16 * tile;
                    Crow = Up + r * ldu; ldc = ldu; c = j0; mr = min(16, k - r); nc = jw;
                } else if (tile < na + nlp) {
                    const int r = 16 * (tile - na);
                    Crow = Lp + r * ldl; ldc = ldl; c = j0; mr = min(16, m - r); nc = jw;
                } else {
                    const int q = tile - na - nlp;
                    const int r = j0 + 16 * (q % nrr);
                    c = j1 + 32 * (q / nrr);
                    Crow = Up + r * ldu; ldc = ldu; mr = min(16, j1 - r); nc = min(32, s - c);
                }
                Acc<2, 4> acc;
                acc.zero();
                warp_gemm_shared<2, 4>(acc, Crow, ldc, mr, Up + c, ldu, nc, j0, lane);
                store_tile<2, 4, true>(acc, Crow + c, ldc, mr, nc, lane);
            }
            __syncthreads();
        }
        invert_tile_inplace<NT>(Up + j0 * ldu + j0, ldu, jw, tid, ps);
        // Lt: rows (j1, k) of Up and all rows of Lp, columns [j0, j1)  <-  F[rows, j0:j1] * D^-1
        const int nlu = (k - j1 + 15) >> 4;
        for (int tile = warp; tile < nlu + nlp; tile += NW) {
            cd* Arow;
            int lda, mr;
            if (tile < nlu) { const int r = j1 + 16 * tile; Arow = Up + r * ldu + j0; lda = ldu; mr = min(16, k - r); }
            else { const int r = 16 * (tile - nlu); Arow = Lp + r * ldl + j0; lda = ldl; mr = min(16, m - r); }
            Acc<2, 4> acc;
            acc.zero();
            warp_gemm_shared<2, 4>(acc, Arow, lda, mr, Up + j0 * ldu + j0, ldu, jw, jw, lane);
            store_tile<2, 4, false>(acc, Arow, lda, mr, jw, lane);
        }
        __syncthreads();
    }
    // contribution block: C = sum over children (gathered) - L21 * U12, written once to HBM
    if (m > 0) {
        cd* C = Fg + (int64_t)k * ld + k;
        const int ntr = (m + 15) >> 4, ntc = (m + 31) >> 5;
        for (int tile = warp; tile < ntr * ntc; tile += NW) {
            const int rc = 16 * (tile % ntr), cc = 32 * (tile / ntr);
            const int mr = min(16, m - rc), nc = min(32, m - cc);
            Acc<2, 4> acc;
            acc.zero();
            warp_gemm_shared<2, 4>(acc, Lp + rc * ldl, ldl, mr, Up + k + cc, ldu, nc, k, lane);
            store_tile_cb_inplace<2, 4>(acc, C + (int64_t)rc * ld + cc, ld, mr, nc, rc, cc, sd, fd, chs, Fsys, lane);
        }
    }
    // the panels leave in one pass: pivot rows in full, then the pivot columns of the other rows
    for (int i0 = warp; i0 < k; i0 += NW)
        for (int j = lane; j < s; j += 32) Fg[(int64_t)i0 * ld + j] = Up[i0 * ldu + j];
    for (int i0 = warp; i0 < m; i0 += NW)
        for (int j = lane; j < k; j += 32) Fg[(int64_t)(k + i0) * ld + j] = Lp[i0 * ldl + j];
    if (tid == 0 && ps.sing && status) atomicOr(status + b, 1);
}

// ------------------------------------------------------------------------ triangular solves

// transposing warp reductions (see K3w below for their use)
__device__ __forceinline__ int row8(int lane) {
    return ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
}
__device__ __forceinline__ cd shfl_xor_cd(cd v, int m) {
    return make_double2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
}
// Generalisation to lane groups of CW = 2^LB lanes (LB = 3, 4, 5): the 32 / CW groups of a
// warp reduce independently; within a group the three upper lane bits select which of the 8
// values a lane ends up with (sel8g) and the remaining LB - 3 bits are summed plainly.
template <int LB>
__device__ __forceinline__ int sel8g(int lane) {
    return ((lane >> (LB - 1)) & 1) * 4 + ((lane >> (LB - 2)) & 1) * 2 + ((lane >> (LB - 3)) & 1);
}
template <int LB>
__device__ __forceinline__ cd reduce8g(const cd (&p)[8], int lane) {
    cd q[4], r[2];
    const bool b4 = lane & (1 << (LB - 1)), b3 = lane & (1 << (LB - 2)), b2 = lane & (1 << (LB - 3));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const cd send = b4 ? p[i] : p[i + 4], keep = b4 ? p[i + 4] : p[i];
        const cd got = shfl_xor_cd(send, 1 << (LB - 1));
        q[i] = make_double2(keep.x + got.x, keep.y + got.y);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const cd send = b3 ? q[i] : q[i + 2], keep = b3 ? q[i + 2] : q[i];
        const cd got = shfl_xor_cd(send, 1 << (LB - 2));
        r[i] = make_double2(keep.x + got.x, keep.y + got.y);
    }
    cd t;
    {
        const cd send = b2 ? r[0] : r[1], keep = b2 ? r[1] : r[0];
        const cd got = shfl_xor_cd(send, 1 << (LB - 3));
        t = make_double2(keep.x + got.x, keep.y + got.y);
    }
#pragma unroll
    for (int o = (1 << (LB - 3)) >> 1; o > 0; o >>= 1) {
        const cd got = shfl_xor_cd(t, o);
        t.x += got.x; t.y += got.y;
    }
    return t;
}
__device__ __forceinline__ cd reduce8(const cd (&p)[8], int lane) { return reduce8g<5>(p, lane); }

// rhs / solution addressing: system b, right-hand side j
struct VecView {
    const cd* rhs;      // rhs(b, j) = rhs + (j * rhs_sj + b * rhs_sb) * N
    int64_t rhs_sb, rhs_sj;
    cd* x;              // x(b, j)   = x + (j * nsys + b) * N
    cd* V;              // front-local vectors, (j * nsys + b) * vec_slots
    int nsys, nrhs;
    // adaptive refinement: when non-null, vectors with active[j * nsys + b] == 0 are skipped
    // (their blocks / warps exit at once), so a late refinement step costs only the systems
    // that still need it
    const int* active;
};

// Forward sweep over one level.  TRANS = false: Lt (unit block lower);  TRANS = true:
// Ubar^T (block lower with D_J^T on the diagonal) -- the first half of A^T x = b.
template <bool TRANS, int NT>
__global__ void __launch_bounds__(NT)
fwd_level_kernel(SymDev sd, const int* __restrict__ level_fronts, const cd* __restrict__ Fall,
                 VecView vv) {
    extern __shared__ cd v[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int SOLVE_THREADS = NT;
    constexpr int NW = NT / 32;
    int bid = blockIdx.x;
    const int b = bid % vv.nsys;
    bid /= vv.nsys;
    const int jr = bid % vv.nrhs;
    if (vv.active && !vv.active[jr * vv.nsys + b]) return;
    const FrontDesc fd = sd.fr[level_fronts[bid / vv.nrhs]];
    const cd* F = Fall + (int64_t)b * sd.front_slots + fd.off;
    cd* V = vv.V + ((int64_t)jr * vv.nsys + b) * sd.vec_slots;
    const cd* rhs = vv.rhs + ((int64_t)jr * vv.rhs_sj + (int64_t)b * vv.rhs_sb) * sd.N;
    const int s = fd.s, k = fd.k, ld = fd.ld;
    const int* rows = sd.rowidx_orig + fd.rowptr;

    for (int i = tid; i < s; i += SOLVE_THREADS)
        v[i] = i < k ? rhs[rows[i]] : make_double2(0.0, 0.0);
    __syncthreads();
    for (int ci = 0; ci < fd.nchild; ++ci) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
        const int* rel = sd.relmap + fc.reloff;
        const int mc = fc.s - fc.k;
        const cd* Vc = V + fc.voff + fc.k;
        for (int i = tid; i < mc; i += SOLVE_THREADS) {
            cd a = Vc[i];
            cd* p = v + rel[i];
            p->x += a.x;
            p->y += a.y;
        }
        __syncthreads();
    }
    const int KT = (k + TW - 1) / TW;
    for (int J = 0; J < KT; ++J) {
        const int j0 = J * TW, jw = min(TW, k - j0), j1 = j0 + jw;
        if (!TRANS) {
            // v[i] -= Lt[i, J] . v_J  for rows below the tile: eight rows per warp trip, lane per
            // column, one transposing reduction per eight rows
            const cd vj = lane < jw ? v[j0 + lane] : make_double2(0.0, 0.0);
            const cd* col = F + j0 + min(lane, jw - 1);
            for (int i0 = j1 + 8 * warp; i0 < s; i0 += 8 * NW) {
                cd p[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) p[q] = col[(int64_t)min(i0 + q, s - 1) * ld];
#pragma unroll
                for (int q = 0; q < 8; ++q) p[q] = cmul(p[q], vj);
                const cd t = reduce8(p, lane);
                const int i = i0 + row8(lane);
                if ((lane & 3) == 0 && i < s) { v[i].x -= t.x; v[i].y -= t.y; }
            }
            __syncthreads();
        } else {
            // z_J = D_J^-T v_J
            if (warp == 0) {
                cd z = make_double2(0.0, 0.0);
                if (lane < jw) {
#pragma unroll 8
                    for (int p = 0; p < jw; ++p) cfma(z, F[(int64_t)(j0 + p) * ld + j0 + lane], v[j0 + p]);
                }
                __syncwarp();
                if (lane < jw) v[j0 + lane] = z;
            }
            __syncthreads();
            // v[j] -= sum_p U[J_p, j] z_p  for columns right of the tile: thread per column
            for (int j = j1 + tid; j < s; j += SOLVE_THREADS) {
                cd acc = make_double2(0.0, 0.0);
#pragma unroll 8
                for (int p = 0; p < jw; ++p) cfma(acc, F[(int64_t)(j0 + p) * ld + j], v[j0 + p]);
                v[j].x -= acc.x;
                v[j].y -= acc.y;
            }
            __syncthreads();
        }
    }
    cd* Vf = V + fd.voff;
    for (int i = tid; i < s; i += SOLVE_THREADS) Vf[i] = v[i];
}

#ifndef JPB_BWD_UNROLL
#define JPB_BWD_UNROLL 8
#endif
// rows in flight per lane in the transposed backward sweep (lane per column, warps stride the rows).  Measured:
// 8 instead of 4 takes the adjoint solves from 32.9 to 30.1 ms at cfg4 and from 11.3 to 10.8 ms at cfg1; the same
// change in the plain sweep (warp per row) costs 5 %, so that loop stays at 4.
constexpr int BWD_UNROLL = JPB_BWD_UNROLL;
// Backward sweep over one level.  TRANS = false: Ubar ; TRANS = true: Lt^T.
template <bool TRANS, int NT>
__global__ void __launch_bounds__(NT)
bwd_level_kernel(SymDev sd, const int* __restrict__ level_fronts, const cd* __restrict__ Fall,
                 VecView vv) {
    extern __shared__ cd v[];
    constexpr int SOLVE_THREADS = NT;
    constexpr int NW = NT / 32;
    __shared__ cd part[NW][TW];
    __shared__ cd tbuf[TW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int bid = blockIdx.x;
    const int b = bid % vv.nsys;
    bid /= vv.nsys;
    const int jr = bid % vv.nrhs;
    if (vv.active && !vv.active[jr * vv.nsys + b]) return;
    const FrontDesc fd = sd.fr[level_fronts[bid / vv.nrhs]];
    const cd* F = Fall + (int64_t)b * sd.front_slots + fd.off;
    const cd* Vf = vv.V + ((int64_t)jr * vv.nsys + b) * sd.vec_slots + fd.voff;
    cd* x = vv.x + ((int64_t)jr * vv.nsys + b) * sd.N;
    const int s = fd.s, k = fd.k, ld = fd.ld;
    const int* rows = sd.rowidx_orig + fd.rowptr;

    for (int i = tid; i < s; i += SOLVE_THREADS) v[i] = i < k ? Vf[i] : x[rows[i]];
    __syncthreads();
    const int KT = (k + TW - 1) / TW;
    for (int J = KT - 1; J >= 0; --J) {
        const int j0 = J * TW, jw = min(TW, k - j0), j1 = j0 + jw;
        if (!TRANS) {
            // t_p = v[j0+p] - U[J_p, j1:] . v[j1:]   warp per row, lanes stride the columns
            for (int p = warp; p < jw; p += NW) {
                const cd* row = F + (int64_t)(j0 + p) * ld;
                cd acc = make_double2(0.0, 0.0);
#pragma unroll 4
                for (int j = j1 + lane; j < s; j += 32) cfma(acc, row[j], v[j]);
                acc = warp_sum(acc);
                if (lane == 0) tbuf[p] = make_double2(v[j0 + p].x - acc.x, v[j0 + p].y - acc.y);
            }
            __syncthreads();
            // v_J = D_J^-1 t
            for (int q = warp; q < jw; q += NW) {
                cd acc = make_double2(0.0, 0.0);
                if (lane < jw) acc = cmul(F[(int64_t)(j0 + q) * ld + j0 + lane], tbuf[lane]);
                acc = warp_sum(acc);
                if (lane == 0) v[j0 + q] = acc;
            }
            __syncthreads();
        } else {
            // v[j0+q] -= sum_{i >= j1} Lt[i, j0+q] v[i] : lane per column, warps stride the rows
            cd acc = make_double2(0.0, 0.0);
            if (lane < jw) {
#pragma unroll BWD_UNROLL
                for (int i = j1 + warp; i < s; i += NW) cfma(acc, F[(int64_t)i * ld + j0 + lane], v[i]);
            }
            part[warp][lane] = acc;
            __syncthreads();
            if (tid < jw) {
                cd sum = make_double2(0.0, 0.0);
#pragma unroll
                for (int w = 0; w < NW; ++w) { sum.x += part[w][tid].x; sum.y += part[w][tid].y; }
                v[j0 + tid].x -= sum.x;
                v[j0 + tid].y -= sum.y;
            }
            __syncthreads();
        }
    }
    for (int i = tid; i < k; i += SOLVE_THREADS) x[rows[i]] = v[i];
}

// ------------------------------------------- triangular solves, one warp per system (K3w)
// Small and mid-size fronts hold only a few KB of factor data each, so a CTA per (front,
// system) spends its life fetching descriptors and index lists.  Here a CTA takes one front
// for NW systems, one WARP per system: descriptors, row lists and extend-add maps are
// fetched once per CTA (L1 serves the other warps), the front-local vector of a system lives
// in that warp's slice of shared memory, and every synchronisation is a __syncwarp.
//
// reduce8: every lane holds 8 partial sums (rows 0..7 of a block); on return the lanes with
// (lane & 3) == 0 hold the total of row row8(lane).  A transposing butterfly: 36 32-bit
// shuffles per 8 rows instead of 160 for eight plain warp reductions.
// v[i] -= Lt[i, J] . v_J for the rows i in [j1, s) of a pivot tile of width jw <= 2^LB: a warp
// load covers 32 / CW rows x CW columns (CW = 2^LB), eight loads are in flight per lane, so a
// narrow strip (jw <= 8: 32 rows) is one memory round trip.  v in shared memory.
template <int LB>
__device__ __forceinline__ void strip_fwd(const cd* F, int ld, int s, int j0, int jw, int j1, cd* v,
                                          int lane) {
    constexpr int CW = 1 << LB, RPL = 32 / CW;
    const int c = lane & (CW - 1), ro = lane >> LB;
    const cd zero = make_double2(0.0, 0.0);
    const cd vj = c < jw ? v[j0 + c] : zero;
    const cd* col = F + j0 + min(c, jw - 1);
    for (int i0 = j1; i0 < s; i0 += 8 * RPL) {
        cd p[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) p[q] = col[(int64_t)min(i0 + q * RPL + ro, s - 1) * ld];
#pragma unroll
        for (int q = 0; q < 8; ++q) p[q] = cmul(p[q], vj);
        const cd t = reduce8g<LB>(p, lane);
        const int i = i0 + sel8g<LB>(lane) * RPL + ro;
        if ((lane & ((CW >> 3) - 1)) == 0 && i < s) { v[i].x -= t.x; v[i].y -= t.y; }
    }
}
// v[j0+c] -= sum_{i >= j1} Lt[i, j0+c] v[i]  (the transposed use of the same strip)
template <int LB>
__device__ __forceinline__ void strip_bwd_t(const cd* F, int ld, int s, int j0, int jw, int j1, cd* v,
                                            int lane) {
    constexpr int CW = 1 << LB, RPL = 32 / CW;
    const int c = lane & (CW - 1), ro = lane >> LB;
    cd acc = make_double2(0.0, 0.0);
    const cd* col = F + j0 + min(c, jw - 1);
    for (int i0 = j1; i0 < s; i0 += 8 * RPL) {
        cd a[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) a[q] = col[(int64_t)min(i0 + q * RPL + ro, s - 1) * ld];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int i = i0 + q * RPL + ro;
            if (i < s) cfma(acc, a[q], v[i]);
        }
    }
#pragma unroll
    for (int o = CW; o < 32; o <<= 1) {
        const cd got = shfl_xor_cd(acc, o);
        acc.x += got.x; acc.y += got.y;
    }
    __syncwarp();
    if (lane < jw) { v[j0 + lane].x -= acc.x; v[j0 + lane].y -= acc.y; }
}

// per-warp shared slice: [vstride] front-local vector followed by [TW] scratch
template <bool TRANS, int NW>
__global__ void __launch_bounds__(NW * 32)
fwd_warp_kernel(SymDev sd, const int* __restrict__ level_fronts, const cd* __restrict__ Fall,
                VecView vv, int vstride) {
    extern __shared__ cd vsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = vv.nsys * vv.nrhs;
    const int groups = (nvec + NW - 1) / NW;
    const int vec = (blockIdx.x % groups) * NW + warp;
    if (vec >= nvec || (vv.active && !vv.active[vec])) return;          // no block-wide barrier below
    const int b = vec % vv.nsys, jr = vec / vv.nsys;
    const FrontDesc fd = sd.fr[level_fronts[blockIdx.x / groups]];
    const cd* F = Fall + (int64_t)b * sd.front_slots + fd.off;
    cd* V = vv.V + (int64_t)vec * sd.vec_slots;
    const cd* rhs = vv.rhs + ((int64_t)jr * vv.rhs_sj + (int64_t)b * vv.rhs_sb) * sd.N;
    const int s = fd.s, k = fd.k, ld = fd.ld;
    const int* rows = sd.rowidx_orig + fd.rowptr;
    cd* v = vsm + warp * (vstride + TW);
    const cd zero = make_double2(0.0, 0.0);

    for (int i = lane; i < s; i += 32) v[i] = i < k ? rhs[__ldg(rows + i)] : zero;
    __syncwarp();
    for (int ci = 0; ci < fd.nchild; ++ci) {
        const FrontDesc fc = sd.fr[sd.children[fd.childptr + ci]];
        const int* rel = sd.relmap + fc.reloff;
        const int mc = fc.s - fc.k;
        const cd* Vc = V + fc.voff + fc.k;
        for (int i = lane; i < mc; i += 32) {
            const cd a = Vc[i];
            cd* p = v + __ldg(rel + i);
            p->x += a.x;
            p->y += a.y;
        }
        __syncwarp();
    }
    const int KT = (k + TW - 1) / TW;
    for (int J = 0; J < KT; ++J) {
        const int j0 = J * TW, jw = min(TW, k - j0), j1 = j0 + jw;
        if (!TRANS) {
            if (jw <= 8) strip_fwd<3>(F, ld, s, j0, jw, j1, v, lane);
            else if (jw <= 16) strip_fwd<4>(F, ld, s, j0, jw, j1, v, lane);
            else strip_fwd<5>(F, ld, s, j0, jw, j1, v, lane);
            __syncwarp();
        } else {
            // z_J = D_J^-T v_J, then v[j] -= sum_p U[J_p, j] z_p right of the tile: lane = column
            cd z = zero;
            {
                const cd* col = F + (int64_t)j0 * ld + j0 + min(lane, jw - 1);
#pragma unroll 8
                for (int p = 0; p < jw; ++p) cfma(z, col[(int64_t)p * ld], v[j0 + p]);
            }
            __syncwarp();
            if (lane < jw) v[j0 + lane] = z;
            __syncwarp();
            for (int j = j1 + lane; j < s; j += 32) {
                const cd* col = F + (int64_t)j0 * ld + j;
                cd acc = zero;
#pragma unroll 8
                for (int p = 0; p < jw; ++p) cfma(acc, col[(int64_t)p * ld], v[j0 + p]);
                v[j].x -= acc.x;
                v[j].y -= acc.y;
            }
            __syncwarp();
        }
    }
    cd* Vf = V + fd.voff;
    for (int i = lane; i < s; i += 32) Vf[i] = v[i];
}

#ifndef JPB_BWD_OCC
#define JPB_BWD_OCC 4
#endif
template <bool TRANS, int NW>
__global__ void __launch_bounds__(NW * 32, TRANS ? 4 : JPB_BWD_OCC)
bwd_warp_kernel(SymDev sd, const int* __restrict__ level_fronts, const cd* __restrict__ Fall,
                VecView vv, int vstride) {
    extern __shared__ cd vsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = vv.nsys * vv.nrhs;
    const int groups = (nvec + NW - 1) / NW;
    const int vec = (blockIdx.x % groups) * NW + warp;
    if (vec >= nvec || (vv.active && !vv.active[vec])) return;
    const int b = vec % vv.nsys;
    const FrontDesc fd = sd.fr[level_fronts[blockIdx.x / groups]];
    const cd* F = Fall + (int64_t)b * sd.front_slots + fd.off;
    const cd* Vf = vv.V + (int64_t)vec * sd.vec_slots + fd.voff;
    cd* x = vv.x + (int64_t)vec * sd.N;
    const int s = fd.s, k = fd.k, ld = fd.ld;
    const int* rows = sd.rowidx_orig + fd.rowptr;
    cd* v = vsm + warp * (vstride + TW);
    cd* tbuf = v + vstride;
    const cd zero = make_double2(0.0, 0.0);

    for (int i = lane; i < s; i += 32) v[i] = i < k ? Vf[i] : x[__ldg(rows + i)];
    __syncwarp();
    const int KT = (k + TW - 1) / TW;
    for (int J = KT - 1; J >= 0; --J) {
        const int j0 = J * TW, jw = min(TW, k - j0), j1 = j0 + jw;
        if (!TRANS) {
            // t_p = v[j0+p] - U[J_p, j1:] . v[j1:], eight rows of U per trip, lanes along the row
            for (int p0 = 0; p0 < jw; p0 += 8) {
                cd acc[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) acc[q] = zero;
                const cd* row0 = F + (int64_t)(j0 + p0) * ld;
                for (int j = j1 + lane; j < s; j += 32) {
                    const cd vj = v[j];
                    cd a[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) a[q] = row0[(int64_t)min(q, jw - 1 - p0) * ld + j];
#pragma unroll
                    for (int q = 0; q < 8; ++q) cfma(acc[q], a[q], vj);
                }
                const cd t = reduce8(acc, lane);
                const int p = p0 + row8(lane);
                if ((lane & 3) == 0 && p < jw) tbuf[p] = make_double2(v[j0 + p].x - t.x, v[j0 + p].y - t.y);
            }
            __syncwarp();
            // v_J = D_J^-1 t
            const cd tl = lane < jw ? tbuf[lane] : zero;
            const cd* col = F + (int64_t)j0 * ld + j0 + min(lane, jw - 1);
            for (int q0 = 0; q0 < jw; q0 += 8) {
                cd p[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) p[q] = col[(int64_t)min(q0 + q, jw - 1) * ld];
#pragma unroll
                for (int q = 0; q < 8; ++q) p[q] = cmul(p[q], tl);
                const cd t = reduce8(p, lane);
                const int qq = q0 + row8(lane);
                if ((lane & 3) == 0 && qq < jw) v[j0 + qq] = t;
            }
            __syncwarp();
        } else {
            if (jw <= 8) strip_bwd_t<3>(F, ld, s, j0, jw, j1, v, lane);
            else if (jw <= 16) strip_bwd_t<4>(F, ld, s, j0, jw, j1, v, lane);
            else strip_bwd_t<5>(F, ld, s, j0, jw, j1, v, lane);
            __syncwarp();
        }
    }
    for (int i = lane; i < k; i += 32) x[__ldg(rows + i)] = v[i];
}

// ------------------------------------------------------------------------ sparse products
// y(b, j) = [rhs(b, j) -] sum_{q in ptr[i]..ptr[i+1]} vals[b, pos[q]] * x(b, j)[idx[q]]
// (ptr, idx, pos) is the CSR view for A x and the CSC view for A^T x; pos == nullptr
// means identity.  grid.y = system * nrhs, one thread per row.
template <bool NORMS>
__global__ void spmv_kernel(int N, int64_t nnz, const int* __restrict__ ptr,
                            const int* __restrict__ idx, const int* __restrict__ pos,
                            const cd* __restrict__ vals, int64_t vals_sb,
                            const cd* __restrict__ x, int64_t x_sb, int64_t x_sj,
                            const cd* __restrict__ rhs, int64_t rhs_sb, int64_t rhs_sj,
                            cd* y, int nsys, double* norms, const int* active) {
    if (active && !active[blockIdx.y]) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y % nsys, j = blockIdx.y / nsys;
    // residual check: |r_i|^2 and t_i^2 with t_i = sum_e |a_e| |x_e| + |rhs_i| (1-norm moduli),
    // i.e. the ingredients of the normwise backward error ||r|| / || |A||x| + |b| ||
    double nr = 0.0, nb2 = 0.0, nbb = 0.0, tacc = 0.0;
    auto m1 = [](cd v) { return fabs(v.x) + fabs(v.y); };
    if (i < N) {
    const cd* vb = vals + (int64_t)b * vals_sb * nnz;
    const cd* xb = x + ((int64_t)j * x_sj + (int64_t)b * x_sb) * N;
    cd acc = make_double2(0.0, 0.0);
    // eight, then four entries per trip: the index loads, then the value / vector gathers, go
    // out together (a row holds ~24 entries, so the dependent round trips per row drop to 3-4)
    const int q1 = ptr[i + 1];
    int q = ptr[i];
    for (; q + 8 <= q1; q += 8) {
        int pq[8], jq[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { pq[u] = pos ? __ldg(pos + q + u) : q + u; jq[u] = __ldg(idx + q + u); }
        cd a[8], xv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { a[u] = vb[pq[u]]; xv[u] = xb[jq[u]]; }
#pragma unroll
        for (int u = 0; u < 8; ++u) cfma(acc, a[u], xv[u]);
        if (NORMS) {
#pragma unroll
            for (int u = 0; u < 8; ++u) tacc = fma(m1(a[u]), m1(xv[u]), tacc);
        }
    }
    for (; q + 4 <= q1; q += 4) {
        int pq[4], jq[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { pq[u] = pos ? __ldg(pos + q + u) : q + u; jq[u] = __ldg(idx + q + u); }
        cd a[4], xv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) { a[u] = vb[pq[u]]; xv[u] = xb[jq[u]]; }
#pragma unroll
        for (int u = 0; u < 4; ++u) cfma(acc, a[u], xv[u]);
        if (NORMS) {
#pragma unroll
            for (int u = 0; u < 4; ++u) tacc = fma(m1(a[u]), m1(xv[u]), tacc);
        }
    }
    for (; q < q1; ++q) {
        cd a = vb[pos ? pos[q] : q];
        const cd xv = xb[idx[q]];
        cfma(acc, a, xv);
        if (NORMS) tacc = fma(m1(a), m1(xv), tacc);
    }
    cd* yb = y + ((int64_t)j * nsys + b) * N;
    if (rhs) {
        cd r = rhs[((int64_t)j * rhs_sj + (int64_t)b * rhs_sb) * N + i];
        acc.x = r.x - acc.x;
        acc.y = r.y - acc.y;
        if (NORMS) {
            nr = acc.x * acc.x + acc.y * acc.y;
            nbb = r.x * r.x + r.y * r.y;
            tacc += m1(r);
            nb2 = tacc * tacc;
        }
    }
    yb[i] = acc;
    }
    // norms[(j * nsys + b) * 3 + {0, 1, 2}] += sum_i {|r_i|^2, t_i^2, |rhs_i|^2}: feeds the per-system
    // residual tests (refinement decision, status bit 2); the summation order is irrelevant for
    // a threshold test
    if (NORMS) {
        __shared__ double red[3][32];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            nr += __shfl_xor_sync(0xffffffffu, nr, o);
            nb2 += __shfl_xor_sync(0xffffffffu, nb2, o);
            nbb += __shfl_xor_sync(0xffffffffu, nbb, o);
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) { red[0][warp] = nr; red[1][warp] = nb2; red[2][warp] = nbb; }
        __syncthreads();
        if (warp == 0) {
            const int nw = blockDim.x >> 5;
            double a = lane < nw ? red[0][lane] : 0.0, c = lane < nw ? red[1][lane] : 0.0,
                   e = lane < nw ? red[2][lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a += __shfl_xor_sync(0xffffffffu, a, o);
                c += __shfl_xor_sync(0xffffffffu, c, o);
                e += __shfl_xor_sync(0xffffffffu, e, o);
            }
            if (lane == 0) {
                atomicAdd(norms + 3 * (int64_t)blockIdx.y, a);
                atomicAdd(norms + 3 * (int64_t)blockIdx.y + 1, c);
                atomicAdd(norms + 3 * (int64_t)blockIdx.y + 2, e);
            }
        }
    }
}

// status[vec % nsys] |= 2 where ||r||^2 > tol2 * || |A||x| + |b| ||^2 (or the residual is not finite)
__global__ void residual_flag_kernel(int nvec, int nsys, const double* __restrict__ norms, double tol2,
                                     int* status) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvec) return;
    const double nr = norms[3 * v], nb = norms[3 * v + 1];
    if (!(nr <= tol2 * nb) && !(nr == 0.0)) atomicOr(status + v % nsys, 2);
}

// Residual-driven refinement: a vector gets a refinement step (active[v] = 1) when its normwise
// backward error omega_v exceeds refine_omega OR its residual relative to the right-hand side,
// ||r|| / ||b||, exceeds refine_resid (the systems next to a resonance, where the solution is
// amplified by 1e3 ... 1e6 and the same backward error leaves a larger residual);
// status[v % nsys] |= 2 where omega_v > tol0 or is not finite (a step is not expected to rescue
// that).  norms[3 v + {0, 1, 2}] = ||r||^2, || |A||x| + |b| ||^2, ||b||^2.  counters[0] counts the
// refined vectors, counters[1] all of them (instrumentation).
__global__ void omega_active_kernel(int nvec, int nsys, const double* __restrict__ norms, double ro2, double rr2,
                                    double tol2, int* active, int* status, unsigned long long* counters) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvec) return;
    const double nr = norms[3 * v], nb = norms[3 * v + 1], nbb = norms[3 * v + 2];
    const bool finite = (nr == nr) && (nb == nb) && nr < 1e300 && nb < 1e300;
    const int act = (finite && (nr > ro2 * nb || nr > rr2 * nbb)) ? 1 : 0;
    active[v] = act;
    if (counters) {
        if (act) atomicAdd(counters, 1ull);
        atomicAdd(counters + 1, 1ull);
    }
    if (status && !(nr <= tol2 * nb) && !(nr == 0.0)) atomicOr(status + v % nsys, 2);
}

// ------------------------------------------------------------------------- equilibration
// Symmetric diagonal scaling A~ = S A S with s_i = 2^-floor(log2|a_ii| / 2): a power of two, so
// the scaled entries are exact and the scaled system is the same problem in other units.  It
// removes the 1e7 spread between the in-plane and the bending blocks of the plate matrices
// from the pivot tiles; with it the block factorisation with tile-local pivoting reaches the
// backward error of a CPU sparse LU in BOTH the plain and the transposed solve without a
// refinement step (unscaled: 10-40x larger residual in the transposed solve; measured with the
// numpy replay of the schedule, profiles/r2_summary.md).  s = 1 where the diagonal entry is
// absent, zero or not finite.
__global__ void diag_scale_kernel(int N, int64_t nnz, const int* __restrict__ diagpos,
                                  const cd* __restrict__ vals, double* s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int b = blockIdx.y;
    double out = 1.0;
    const int q = diagpos[i];
    if (q >= 0) {
        const cd d = vals[(int64_t)b * nnz + q];
        const double m = fmax(fabs(d.x), fabs(d.y));
        if (m > 0.0 && m < 1e300) out = scalbn(1.0, -(ilogb(m) >> 1));
    }
    s[(int64_t)b * N + i] = out;
}
// vals[b, q] *= s[b, row q] * s[b, col q]
__global__ void scale_values_kernel(int64_t nnz, int N, const int* __restrict__ row, const int* __restrict__ col,
                                    const double* __restrict__ s, cd* vals) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nnz) return;
    const int b = blockIdx.y;
    const double* sb = s + (int64_t)b * N;
    const double f = sb[row[q]] * sb[col[q]];
    cd v = vals[(int64_t)b * nnz + q];
    v.x *= f; v.y *= f;
    vals[(int64_t)b * nnz + q] = v;
}
// v[vec, i] *= s[vec % nsys, i]
__global__ void scale_vec_kernel(int N, int nsys, const double* __restrict__ s, cd* v) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int vec = blockIdx.y;
    const double f = s[(int64_t)(vec % nsys) * N + i];
    cd a = v[(int64_t)vec * N + i];
    a.x *= f; a.y *= f;
    v[(int64_t)vec * N + i] = a;
}
// dst[vec, i] = src[i]
__global__ void broadcast_vec_kernel(int N, const cd* __restrict__ src, cd* dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) dst[(int64_t)blockIdx.y * N + i] = src[i];
}

// x += d  (vector v = blockIdx.y; skipped when inactive)
__global__ void axpy_kernel(int N, cd* x, const cd* __restrict__ d, const int* __restrict__ active) {
    const int v = blockIdx.y;
    if (active && !active[v]) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) {
        const int64_t q = (int64_t)v * N + i;
        cd a = x[q], c = d[q];
        a.x += c.x; a.y += c.y;
        x[q] = a;
    }
}

// Convergence test of the refinement: one CTA per vector; the vector stays active while the
// correction just applied is larger than tol * ||x|| (max-norms), otherwise it is retired.
__global__ void refine_active_kernel(int N, const cd* __restrict__ x, const cd* __restrict__ d, double tol,
                                     int* active, int first) {
    __shared__ double red[2][32];
    const int v = blockIdx.x;
    if (!first && !active[v]) return;
    double md = 0.0, mx = 0.0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const cd a = d[(int64_t)v * N + i], b = x[(int64_t)v * N + i];
        md = fmax(md, fmax(fabs(a.x), fabs(a.y)));
        mx = fmax(mx, fmax(fabs(b.x), fabs(b.y)));
    }
    for (int o = 16; o > 0; o >>= 1) {
        md = fmax(md, __shfl_xor_sync(0xffffffffu, md, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = md; red[1][threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { md = fmax(md, red[0][w]); mx = fmax(mx, red[1][w]); }
        // NaN-safe: a non-finite correction retires the vector (the residual test reports it)
        active[v] = (md > tol * mx) && (md == md) && (md < 1e300) ? 1 : 0;
    }
}

// data[b, e] = sum_k coef[b,k] * mats[k,e]   (K1; HBM-bound: reads K*8 B, writes 16 B per entry)
// Each thread keeps one entry of all K matrices in registers and loops over the systems of
// its batch slice, so mats is read once per slice.
// SCALE: the equilibration factors s[b, row e] * s[b, col e] (powers of two, from
// diag_scale_coef_kernel) are applied as the entry is formed, so the scaled matrix costs no
// extra pass over the values.
template <int KK, bool SCALE = false>
__global__ void assemble_kernel(int64_t nnz, int K, int B, const double* __restrict__ mats,
                                const cd* __restrict__ coef, cd* data, int bchunk,
                                const int* __restrict__ row = nullptr, const int* __restrict__ col = nullptr,
                                const double* __restrict__ scal = nullptr, int N = 0) {
    extern __shared__ cd csh[];   // [bchunk][K]
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int b0 = blockIdx.y * bchunk;
    const int nb = min(bchunk, B - b0);
    for (int q = threadIdx.x; q < nb * K; q += blockDim.x) csh[q] = coef[(int64_t)b0 * K + q];
    __syncthreads();
    if (e >= nnz) return;
    double m[KK];
#pragma unroll
    for (int kk = 0; kk < KK; ++kk) m[kk] = kk < K ? mats[(int64_t)kk * nnz + e] : 0.0;
    int er = 0, ec = 0;
    if (SCALE) { er = row[e]; ec = col[e]; }
    for (int bb = 0; bb < nb; ++bb) {
        cd v = make_double2(0.0, 0.0);
#pragma unroll
        for (int kk = 0; kk < KK; ++kk) {
            if (kk < K) {
                cd c = csh[bb * K + kk];
                v.x = fma(c.x, m[kk], v.x);
                v.y = fma(c.y, m[kk], v.y);
            }
        }
        if (SCALE) {
            const double* sb = scal + (int64_t)(b0 + bb) * N;
            const double f = sb[er] * sb[ec];
            v.x *= f; v.y *= f;
        }
        data[(int64_t)(b0 + bb) * nnz + e] = v;
    }
}

// s[b, i] of the equilibration from the diagonal of A_b = sum_k coef[b, k] M_k, formed from the
// K diagonals of the component matrices (mats_diag [K][N], extracted at registration) -- same
// exponent rule as diag_scale_kernel
template <int KK>
__global__ void diag_scale_coef_kernel(int N, int K, const double* __restrict__ mats_diag,
                                       const cd* __restrict__ coef, double* s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const int b = blockIdx.y;
    cd v = make_double2(0.0, 0.0);
#pragma unroll
    for (int kk = 0; kk < KK; ++kk)
        if (kk < K) {
            const cd c = coef[(int64_t)b * K + kk];
            const double m = mats_diag[(int64_t)kk * N + i];
            v.x = fma(c.x, m, v.x);
            v.y = fma(c.y, m, v.y);
        }
    const double mm = fmax(fabs(v.x), fabs(v.y));
    s[(int64_t)b * N + i] = (mm > 0.0 && mm < 1e300) ? scalbn(1.0, -(ilogb(mm) >> 1)) : 1.0;
}

// ------------------------------------------------------------------- read-out / loss / grad
// out[b, r] = sum_i Q[r, i] * u[b, i]   (Q real), R <= 4; one CTA per system
__global__ void readout_kernel(int N, int R, const double* __restrict__ Q,
                               const cd* __restrict__ u, cd* out) {
    __shared__ cd red[4][32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const cd* ub = u + (int64_t)b * N;
    cd acc[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) acc[r] = make_double2(0.0, 0.0);
    for (int i = tid; i < N; i += blockDim.x) {
        cd uv = ub[i];
#pragma unroll
        for (int r = 0; r < 4; ++r)
            if (r < R) {
                double q = Q[(int64_t)r * N + i];
                acc[r].x = fma(q, uv.x, acc[r].x);
                acc[r].y = fma(q, uv.y, acc[r].y);
            }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        acc[r] = warp_sum(acc[r]);
        if (lane == 0) red[r][warp] = acc[r];
    }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            cd v = lane < nw ? red[r][lane] : make_double2(0.0, 0.0);
            v = warp_sum(v);
            if (lane == 0 && r < R) out[(int64_t)b * R + r] = v;
        }
    }
}

// fr_b = sqrt(sum_r w_r |y_br|^2); loss term; ybar[b,r] = dL/dfr * w_r * conj(y_br) / fr_b
// (JAX cotangent convention).  One thread per system.
__global__ void loss_kernel(int B, int R, const cd* __restrict__ y, const double* __restrict__ w,
                            const double* __restrict__ ref, int loss_type, double inv_F,
                            double* fr, double* loss_term, cd* ybar) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double s2 = 0.0;
    for (int r = 0; r < R; ++r) {
        cd v = y[(int64_t)b * R + r];
        s2 += w[r] * (v.x * v.x + v.y * v.y);
    }
    const double f = sqrt(s2), rf = (ref && loss_type < 5) ? ref[b] : 1.0;
    double term, dfr;
    if (loss_type == 4) {            // identity: the cotangent of fr itself (Jacobian rows)
        term = f; dfr = 1.0; inv_F = 1.0;
    } else if (loss_type == 0) {     // MSE
        term = (f - rf) * (f - rf); dfr = 2.0 * (f - rf);
    } else if (loss_type == 1) {     // RMSE
        double q = (f - rf) / rf; term = q * q; dfr = 2.0 * (f - rf) / (rf * rf);
    } else if (loss_type == 2) {     // MSE_AFC
        double q = fabs(f) - fabs(rf); term = q * q; dfr = 2.0 * q * (f >= 0.0 ? 1.0 : -1.0);
    } else if (loss_type == 3) {     // MSE_LOG_AFC
        double q = log(fabs(f)) - log(fabs(rf)); term = q * q; dfr = 2.0 * q / f;
    } else {
        // complex read-out (R = 1, mid-plane symmetric path, Problem.py:415 with the losses of :948-960):
        // 5: |y - ref|^2, 6: |(y - ref) / ref|^2 with a COMPLEX reference (ref = [B] interleaved
        // re, im); cotangent ybar = 2 conj(y - ref) [/ |ref|^2]
        const cd rc = reinterpret_cast<const cd*>(ref)[b];
        const cd v = y[b];
        const double dx = v.x - rc.x, dy = v.y - rc.y;
        const double sc = loss_type == 6 ? 1.0 / (rc.x * rc.x + rc.y * rc.y) : 1.0;
        fr[b] = f;
        loss_term[b] = (dx * dx + dy * dy) * sc;
        const double c = 2.0 * sc * inv_F;
        ybar[b] = make_double2(c * dx, -c * dy);
        return;
    }
    fr[b] = f;
    loss_term[b] = term;
    dfr *= inv_F;
    for (int r = 0; r < R; ++r) {
        cd v = y[(int64_t)b * R + r];
        double c = dfr * w[r] / f;
        ybar[(int64_t)b * R + r] = make_double2(c * v.x, -c * v.y);
    }
}

// ubar[b, i] = sum_r ybar[b, r] * Q[r, i]
__global__ void cotangent_kernel(int N, int R, const double* __restrict__ Q,
                                 const cd* __restrict__ ybar, cd* ubar) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= N) return;
    cd acc = make_double2(0.0, 0.0);
    for (int r = 0; r < R; ++r) {
        double q = Q[(int64_t)r * N + i];
        cd yb = ybar[(int64_t)b * R + r];
        acc.x = fma(q, yb.x, acc.x);
        acc.y = fma(q, yb.y, acc.y);
    }
    ubar[(int64_t)b * N + i] = acc;
}

// b[b, i] = scale[b] * rhs_vec[i]
__global__ void scale_rhs_kernel(int N, const double* __restrict__ rhs_vec,
                                 const cd* __restrict__ scale, cd* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= N) return;
    double r = rhs_vec[i];
    cd sc = scale[b];
    out[(int64_t)b * N + i] = make_double2(sc.x * r, sc.y * r);
}

// K5: partial[chunk, b, k] = - sum_{e in chunk} mats[k, e] * lam[b, row e] * u[b, col e]
// This is the GEMM  C[B x K] = W[B x nnz] * M^T[nnz x K]  with W never materialised
// (w_be = lam_b[row e] * u_b[col e] is formed in registers as the A fragment), run on the
// FP64 tensor cores: per warp 8 systems x 4 entries x 32 components per DMMA group.
// A block stages GRAD_E entries of all K matrices (+ their row/col indices) in shared
// memory and serves 8 warps x 8 systems with them; GRAD_SUB such slices per block.
constexpr int GRAD_THREADS = 256;
constexpr int GRAD_E = 256;
constexpr int GRAD_SUB = 8;
constexpr int GRAD_CHUNK = GRAD_E * GRAD_SUB;      // entries per block
constexpr int GRAD_SYS = 8 * (GRAD_THREADS / 32);  // systems per block
constexpr int GRAD_LD = GRAD_E + 4;
constexpr size_t GRAD_SMEM = (size_t)KMAX * GRAD_LD * sizeof(double) + 2 * GRAD_E * sizeof(int);

__global__ void __launch_bounds__(GRAD_THREADS)
grad_kernel(int64_t nnz, int N, int K, const int* __restrict__ row, const int* __restrict__ col,
            const double* __restrict__ mats, const cd* __restrict__ lam, const cd* __restrict__ u,
            cd* partial, int nsys, int ngroups) {
    extern __shared__ double gsm[];
    double (*ms)[GRAD_LD] = reinterpret_cast<double (*)[GRAD_LD]>(gsm);
    int* rs = reinterpret_cast<int*>(gsm + (size_t)KMAX * GRAD_LD);
    int* cs = rs + GRAD_E;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int group = blockIdx.x % ngroups, chunk = blockIdx.x / ngroups;
    const int b = group * GRAD_SYS + warp * 8 + g;
    const bool valid = b < nsys;
    const cd* lb = lam + (int64_t)(valid ? b : 0) * N;
    const cd* ub = u + (int64_t)(valid ? b : 0) * N;
    double re[KMAX / 8][2], im[KMAX / 8][2];
#pragma unroll
    for (int nb = 0; nb < KMAX / 8; ++nb) { re[nb][0] = re[nb][1] = im[nb][0] = im[nb][1] = 0.0; }
    for (int sub = 0; sub < GRAD_SUB; ++sub) {
        const int64_t e0 = (int64_t)chunk * GRAD_CHUNK + (int64_t)sub * GRAD_E;
        if (e0 >= nnz) break;
        __syncthreads();
        for (int idx = tid; idx < KMAX * GRAD_E; idx += GRAD_THREADS) {
            const int kk = idx / GRAD_E, e = idx % GRAD_E;
            ms[kk][e] = (kk < K && e0 + e < nnz) ? mats[(int64_t)kk * nnz + e0 + e] : 0.0;
        }
        for (int e = tid; e < GRAD_E; e += GRAD_THREADS) {
            const bool in = e0 + e < nnz;
            rs[e] = in ? row[e0 + e] : 0;
            cs[e] = in ? col[e0 + e] : 0;
        }
        __syncthreads();
#pragma unroll 2
        for (int e4 = 0; e4 < GRAD_E; e4 += 4) {
            const int e = e4 + t;
            cd w = cmul(lb[rs[e]], ub[cs[e]]);
            if (!valid) w = make_double2(0.0, 0.0);
#pragma unroll
            for (int nb = 0; nb < KMAX / 8; ++nb) {
                const double m = ms[8 * nb + g][e];
                dmma(re[nb][0], re[nb][1], w.x, m);
                dmma(im[nb][0], im[nb][1], w.y, m);
            }
        }
    }
    if (valid) {
        cd* out = partial + ((int64_t)chunk * nsys + b) * K;
#pragma unroll
        for (int nb = 0; nb < KMAX / 8; ++nb)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int kk = 8 * nb + 2 * t + e;
                if (kk < K) out[kk] = make_double2(-re[nb][e], -im[nb][e]);
            }
    }
}

// out[b, k] = sum_chunk partial[chunk, b, k]   (fixed order => deterministic)
__global__ void grad_reduce_kernel(int nchunks, int64_t nbk, const cd* __restrict__ partial, cd* out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nbk) return;
    cd v = make_double2(0.0, 0.0);
    for (int c = 0; c < nchunks; ++c) { cd p = partial[(int64_t)c * nbk + i]; v.x += p.x; v.y += p.y; }
    out[i] = v;
}

// out[g] = sum of v[b] over the systems with seg[b] == g (seg == nullptr: one segment);
// one CTA per segment, fixed order => deterministic
__global__ void seg_sum_kernel(int n, const double* __restrict__ v, const int* __restrict__ seg, double* out) {
    __shared__ double red[32];
    const int g = blockIdx.x;
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        if (!seg || seg[i] == g) acc += v[i];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double a = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (threadIdx.x == 0) out[g] = a;
    }
}

// sum of a double vector by one CTA (fixed order)
__global__ void sum_kernel(int n, const double* __restrict__ v, double* out) {
    __shared__ double red[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += v[i];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double a = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (threadIdx.x == 0) *out += a;
    }
}

// ------------------------------------------------------------------------ instrumentation
// register-resident DMMA loop: 16 independent accumulators per warp
__global__ void dmma_peak_kernel(int iters, double* sink) {
    double acc[16][2];
#pragma unroll
    for (int i = 0; i < 16; ++i) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) dmma(acc[i][0], acc[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i][0] + acc[i][1];
    if (s == 12345.678) sink[0] = s;
}

// warp_gemm against a scalar loop on one 16 x 32 tile: err = max |P - P_ref| / max |P_ref|
__global__ void selftest_gemm_kernel(const cd* A, int lda, int mrows, const cd* B, int ldb, int ncols,
                                     int K, cd* C, cd* Cref) {
    const int lane = threadIdx.x;
    Acc<2, 4> acc;
    acc.zero();
    warp_gemm<2, 4>(acc, A, lda, mrows, B, ldb, ncols, K, lane);
    store_tile<2, 4, false>(acc, C, 32, mrows, ncols, lane);
    for (int e = lane; e < 16 * 32; e += 32) {
        int i = e / 32, j = e % 32;
        if (i >= mrows || j >= ncols) continue;
        cd v = make_double2(0.0, 0.0);
        for (int p = 0; p < K; ++p) cfma(v, A[(int64_t)i * lda + p], B[(int64_t)p * ldb + j]);
        Cref[i * 32 + j] = v;
    }
}

// cta_gemm against scalar loops: C0 - A B (two back-to-back calls through one pipe, so that the
// stage / parity bookkeeping across calls is exercised); Cref receives the scalar result
__global__ void __launch_bounds__(256)
selftest_cta_gemm_kernel(const cd* A, int lda, int m, const cd* B, int ldb, int n, int K, cd* C, cd* Cref, int ldc) {
    extern __shared__ __align__(16) unsigned char gemm_smem[];
    GemmPipe gp{reinterpret_cast<cd*>(gemm_smem), reinterpret_cast<uint64_t*>(gemm_smem + 2 * GB_STAGE * sizeof(cd)), 0u};
    const int tid = threadIdx.x;
    if (tid == 0) {
        mbar_init(gp.full, 1);
        mbar_init(gp.full + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto sub = [=](Acc<2, 4>& acc, int r0, int c0, int mr, int nc, int ln) {
        store_tile<2, 4, true>(acc, C + (int64_t)r0 * ldc + c0, ldc, mr, nc, ln);
    };
    const int K1 = K / 2;
    if (K1 > 0) cta_gemm(gp, A, lda, m, B, ldb, n, K1, tid, sub);
    cta_gemm(gp, A + K1, lda, m, B + (int64_t)K1 * ldb, ldb, n, K - K1, tid, sub);
    for (int e = tid; e < m * n; e += 256) {
        const int i = e / n, j = e % n;
        cd v = Cref[(int64_t)i * ldc + j];
        for (int p = 0; p < K; ++p) {
            const cd a = A[(int64_t)i * lda + p], b = B[(int64_t)p * ldb + j];
            v.x -= a.x * b.x - a.y * b.y;
            v.y -= a.x * b.y + a.y * b.x;
        }
        Cref[(int64_t)i * ldc + j] = v;
    }
}

}  // namespace jpb
