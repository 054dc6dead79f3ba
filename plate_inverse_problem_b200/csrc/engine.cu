// Host side of libjaxplate_b200.so: pattern registry, wave scheduling, kernel launches and
// the C-ABI declared in include/jax_plate_b200.h.  Replaces jax_plate_lib's InnerState
// (/root/reference/source/jax_plate_lib/include/InnerState.h:86-471) and the UMFPACK calls
// behind it; see the header for the entry-point-by-entry-point mapping.
#include "../../include/jax_plate_b200.h"

#include <algorithm>
#include <array>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include <math_constants.h>

#include "kernels.cuh"
#include "symbolic.h"

using namespace jpb;

namespace {

struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            char buf_[512];                                                                   \
            snprintf(buf_, sizeof buf_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                     __FILE__, __LINE__);                                                     \
            throw CudaError(buf_);                                                            \
        }                                                                                     \
    } while (0)

struct InvalidArg : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct StateError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// grow-only device buffer
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* get(size_t bytes) {
        if (bytes > cap) {
            if (p) cudaFree(p);
            p = nullptr;
            cap = 0;
            cudaError_t e = cudaMalloc(&p, bytes);
            if (e != cudaSuccess) {
                p = nullptr;
                char buf[256];
                snprintf(buf, sizeof buf, "cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
                throw CudaError(buf);
            }
            cap = bytes;
        }
        return p;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    ~DevBuf() { release(); }
};

// page-locked host mirror of a device staging block: the caller's (pageable) arrays are
// copied through it so the transfers are true asynchronous DMA, one per direction
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    void* get(size_t bytes) {
        if (bytes > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr;
            cap = 0;
            if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
                p = nullptr;
                throw CudaError("cudaHostAlloc of the staging mirror failed");
            }
            cap = bytes;
        }
        return p;
    }
    ~PinBuf() { if (p) cudaFreeHost(p); }
};


template <class T>
T* upload(const std::vector<T>& v) {
    T* d = nullptr;
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    CK(cudaMalloc(&d, bytes));
    if (!v.empty()) CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

struct Pattern {
    Symbolic S;
    std::vector<int> levelptr;
    // fronts of a level are sorted by order, largest first; cls[l] = {#big, #mid} (rest small)
    std::vector<std::array<int, 2>> cls;
    // launch groups of the numeric factorisation, per level: a run of the (size-sorted)
    // front list handled by one kernel instantiation
    struct FactorGroup {
        int first, count;   // into the level's FACTOR front list (d_flf + flevelptr[l])
        int smem;           // dynamic shared bytes (0 for the global-memory kernel)
        int nt;
        int kind;           // 0: factor_level_kernel, 1: factor_smem_kernel, 2: factor_panel_kernel
        int maxm;           // largest contribution-block order in the group (bands of schur_band_kernel)
    };
    std::vector<int> flevelptr;
    int* d_flf = nullptr;
    std::vector<std::vector<FactorGroup>> fgroups;
    // triangular solves: per level {#fronts above wsolve_max (CTA per front), order of the
    // largest remaining front (one warp per system)}
    std::vector<std::array<int, 2>> wsolve;
    // symbolic data on the device
    FrontDesc* d_fr = nullptr;
    int *d_rowidx_orig = nullptr, *d_children = nullptr, *d_relmap = nullptr, *d_ent_src = nullptr,
        *d_ent_slot = nullptr, *d_levelfronts = nullptr, *d_cbinv = nullptr;
    // sparse views: CSC (as given) and CSR (rebuilt); *_pos_ent map into front-entry order
    int *d_csc_ptr = nullptr, *d_csc_row = nullptr, *d_csc_col = nullptr, *d_csc_pos_ent = nullptr;
    int *d_csr_ptr = nullptr, *d_csr_col = nullptr, *d_csr_pos = nullptr, *d_csr_pos_ent = nullptr;
    // equilibration: position of every diagonal entry (CSC order / front-entry order, -1 if absent)
    // and (row, col) of the entries in front-entry order
    int *d_diag_csc = nullptr, *d_diag_ent = nullptr, *d_ent_row = nullptr, *d_ent_col = nullptr;
    // registered components
    int K = 0, R = 0;
    double *d_mats = nullptr, *d_mats_f = nullptr, *d_rhs_vec = nullptr, *d_Q = nullptr;
    double* d_mats_diag = nullptr;     // [K][N] diagonals of the component matrices (equilibration)
    SymDev sd{};
    int max_s = 0;

    ~Pattern() {
        void* ptrs[] = {d_fr, d_rowidx_orig, d_children, d_relmap, d_ent_src, d_ent_slot,
                        d_levelfronts, d_cbinv, d_csc_ptr, d_csc_row, d_csc_col, d_csc_pos_ent, d_csr_ptr,
                        d_csr_col, d_csr_pos, d_csr_pos_ent, d_mats, d_mats_f, d_rhs_vec, d_Q,
                        d_diag_csc, d_diag_ent, d_ent_row, d_ent_col, d_mats_diag, d_flf};
        for (void* p : ptrs)
            if (p) cudaFree(p);
    }
};

enum Phase { PH_FACTOR = 0, PH_FWD, PH_ADJ, PH_SPMV, PH_GRAD, PH_READOUT, PH_H2D, PH_D2H, PH_COUNT };
const char* const PHASE_NAMES[PH_COUNT] = {"jpb:factor", "jpb:fwd_solve", "jpb:adj_solve", "jpb:spmv",
                                           "jpb:grad", "jpb:readout_loss", "jpb:h2d", "jpb:d2h"};

struct TimedSpan {
    int phase;
    cudaEvent_t a, b;
};

}  // namespace

struct jpb_state {
    int device = 0;
    cudaStream_t stream = nullptr;     // the state's own stream
    std::vector<std::unique_ptr<Pattern>> pats;
    std::string err;
    int64_t budget = 0;
    // refinement policy of every triangular solve: `refine` unconditional steps (default 1); with
    // refine = 0 a residual pass decides per system whether ONE step runs (backward error above
    // refine_omega or relative residual above refine_resid), then up to refine_extra
    // correction-driven steps
    int refine = 1;
    double refine_omega = 1e-15;
    double refine_resid = 1e300;   // ... or whose residual relative to the right-hand side is above this (off)
    unsigned long long* counters = nullptr;      // device: {vectors refined, vectors solved} of the last call
    int equilibrate = 1;           // symmetric power-of-two diagonal scaling before the factorisation
    int sm_count = 148;
    int refine_extra = 0;          // adaptive steps on top, only for vectors not yet converged (off by default:
                                   // measured, they do not move the plate systems)
    double refine_tol = 1e-11;     // ... i.e. last correction > refine_tol * ||x|| (max-norms)
    // residual acceptance (status bit 2) on the normwise backward error
    // omega = ||b - A x|| / || |A||x| + |b| ||: jpb_solve checks omega of what it returns against
    // resid_tol; the fused path checks omega of the UNREFINED solution against resid_tol0 (the
    // refinement step that follows contracts it by ~cond(A) * omega) at no extra pass over A
    double resid_tol = 1e-10, resid_tol0 = 1e-6;
    // workspaces
    DevBuf F, V, vals, data, vb, vx, vr, vd, vlam, partial, small, hstage, status, norms, lterm, active, scal;
    PinBuf pin;
    // instrumentation
    std::vector<TimedSpan> spans;
    std::vector<cudaEvent_t> pool;
    size_t pool_used = 0;
    double launches = 0;
    double wave = 0;
    bool timing = true;
    int nstreams = 1;                  // sub-waves of a fused evaluation run concurrently
    std::vector<cudaStream_t> aux;
    std::vector<cudaStream_t> lvl;     // side streams of LevelFork (independent launches of one tree level)

    void ensure_level_streams(int n) {
        while ((int)lvl.size() < n) {
            cudaStream_t q;
            if (cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking) != cudaSuccess)
                throw std::runtime_error("cudaStreamCreate failed");
            lvl.push_back(q);
        }
    }
    void ensure_streams(int n) {
        while ((int)aux.size() < n) {
            cudaStream_t q;
            if (cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking) != cudaSuccess)
                throw std::runtime_error("cudaStreamCreate failed");
            aux.push_back(q);
        }
    }
    ~jpb_state() {
        for (auto q : aux) cudaStreamDestroy(q);
        for (auto q : lvl) cudaStreamDestroy(q);
        for (auto e : pool) cudaEventDestroy(e);
        if (stream) cudaStreamDestroy(stream);
        if (counters) cudaFree(counters);
    }
    cudaEvent_t ev() {
        if (pool_used == pool.size()) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            pool.push_back(e);
        }
        return pool[pool_used++];
    }
    void reset_timing() {
        spans.clear();
        pool_used = 0;
        launches = 0;
        if (counters) cudaMemsetAsync(counters, 0, 16, stream);
    }
};

namespace {

struct Span {
    jpb_state* st;
    cudaStream_t s;
    TimedSpan t;
    bool on;
    Span(jpb_state* st_, cudaStream_t s_, int phase) : st(st_), s(s_), on(st_->timing) {
        nvtxRangePushA(PHASE_NAMES[phase]);     // NVTX range per phase (host-side enqueue span)
        if (!on) return;
        t.phase = phase;
        t.a = st->ev();
        t.b = st->ev();
        cudaEventRecord(t.a, s);
    }
    ~Span() {
        nvtxRangePop();
        if (!on) return;
        cudaEventRecord(t.b, s);
        st->spans.push_back(t);
    }
};

int fail(jpb_state* st, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (st) st->err = buf;
    return code;
}

// status bits of a system: 1 = zero / non-finite pivot block, 2 = residual above tolerance
const char* status_message(int bits) {
    if (bits & 1) return "matrix is singular (zero pivot block)";
    return "matrix is singular to working precision or badly scaled for static pivoting: the "
           "backward error (residual) of a system exceeds the tolerance (jpb_set_residual_tolerance)";
}

#define JPB_TRY(st) try {
#define JPB_CATCH(st)                                                    \
    }                                                                    \
    catch (const InvalidArg& e) { return fail(st, JPB_EINVAL, "%s", e.what()); } \
    catch (const StateError& e) { return fail(st, JPB_ESTATE, "%s", e.what()); } \
    catch (const CudaError& e) { return fail(st, JPB_ECUDA, "%s", e.what()); }   \
    catch (const std::bad_alloc&) { return fail(st, JPB_ENOMEM, "out of host memory"); } \
    catch (const std::exception& e) { return fail(st, JPB_EINVAL, "%s", e.what()); }

Pattern& pattern(jpb_state* st, int solver_num) {
    if (!st) throw InvalidArg("null state");
    if (solver_num < 0 || solver_num >= (int)st->pats.size())
        throw InvalidArg("solver_num out of range (no such pattern registered)");
    return *st->pats[solver_num];
}

int64_t budget_bytes(jpb_state* st) {
    if (st->budget > 0) return st->budget;
    size_t fr = 0, tot = 0;
    CK(cudaMemGetInfo(&fr, &tot));
    // memory already held by our grow-only workspaces is reusable
    size_t held = st->F.cap + st->V.cap + st->vals.cap + st->data.cap + st->vb.cap + st->vx.cap +
                  st->vr.cap + st->vd.cap + st->vlam.cap + st->partial.cap;
    return (int64_t)((fr + held) * 0.8);
}

// ---------------------------------------------------------------------------- launches
void launch_check(jpb_state* st) {
    st->launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw CudaError(std::string("kernel launch failed: ") + cudaGetErrorString(e));
}

// size classes of fronts: CTA width follows the front order.  Defaults can be overridden
// for tuning runs with JPB_FRONT_BIG / JPB_FRONT_MID / JPB_NT_BIG / JPB_NT_MID / JPB_NT_SMALL.
struct Tuning {
    int front_big = 128, front_mid = 64;
    int nt[3] = {256, 128, 64};
    // shared-memory-resident fronts: order bound and CTA width of each class, ascending
    int smem_max = 64;
    int wsolve_max = 160;   // fronts up to this order: triangular solves with one warp per system
    // ... launch groups of `smem_step` consecutive orders (the group's largest front sizes the
    // shared memory, so finer groups mean more resident CTAs for the smaller fronts);
    // CTA width 64 up to order smem_nt64, 128 up to smem_nt128, 256 above
    int smem_step = 4, smem_nt64 = 48, smem_nt128 = 80;
    // 256-thread class: strip updates and Schur complement through the TMA-fed CTA-level GEMM
    // (default off: measured slower than the per-warp loops, profiles/r2_summary.md)
    int tma = 0;
    // fronts whose pivot panels ((2 s - k) k elements) fit this many KB of shared memory go to
    // factor_panel_kernel (0: class off); CTA width 64 up to order panel_nt64, 128 up to panel_nt128
    // (measured at cfg1, factor phase: off 50.6 ms, 48 KB 50.3, 72 KB 52.1, 110 KB 54.0; DRAM traffic of
    // the classes it replaces 45 -> 40 GB: these fronts are bound by their pivot chain, not by memory)
    int panel_max = 48, panel_nt64 = 48, panel_nt128 = 96;
    // launches of one elimination-tree level (one per size class / kernel kind) are independent: run
    // them on side streams so that the narrow ones fill the SMs the wide ones leave idle.  Default off:
    // measured (profiles/r2_summary.md section 11) the factor phase gains 2 % at cfg1 (50.3 -> 49.3 ms), 1 % at
    // cfg4, the solves nothing, and the ~400 extra event calls per evaluation make the device-resident
    // arm host-bound (78.7 -> 90.8 ms per step)
    int fork = 0;
    // a launch of the 256-thread class with no more CTAs than SMs (the one- and two-front levels at the
    // top of the elimination tree of a wave that memory limits to ~one system per SM) runs with 512
    // threads per front instead: one CTA per SM either way, twice the warps to hide its latencies
    int wide = 1;
    // 256-thread class: groups whose largest contribution block has at least this order compute their
    // contribution blocks in schur_band_kernel (several CTAs per front) instead of inside the factor
    // kernel.  0 = never, the default: measured at cfg4 (profiles/r2_summary.md section 13) the split
    // halves the DRAM traffic of the large-front levels (level 12: 65 -> 33 GB) and leaves their time
    // where it was (18.6 -> 19.9 ms): the contribution-block GEMM already runs at ~75 % of the DMMA peak
    int schur_split = 0;
    Tuning() {
        auto geti = [](const char* name, int def) {
            const char* v = getenv(name);
            return v ? atoi(v) : def;
        };
        front_big = geti("JPB_FRONT_BIG", front_big);
        front_mid = geti("JPB_FRONT_MID", front_mid);
        nt[0] = geti("JPB_NT_BIG", nt[0]);
        nt[1] = geti("JPB_NT_MID", nt[1]);
        nt[2] = geti("JPB_NT_SMALL", nt[2]);
        for (int& v : nt)
            if (v != 64 && v != 128 && v != 256) v = 128;
        smem_max = geti("JPB_SMEM_MAX", smem_max);
        wsolve_max = std::min(320, geti("JPB_WSOLVE_MAX", wsolve_max));
        if (smem_max > 112) smem_max = 112;
        smem_step = std::max(1, geti("JPB_SMEM_STEP", smem_step));
        smem_nt64 = geti("JPB_SMEM_NT64", smem_nt64);
        smem_nt128 = geti("JPB_SMEM_NT128", smem_nt128);
        tma = geti("JPB_TMA", tma);
        panel_max = std::min(200, geti("JPB_PANEL_MAX", panel_max));
        panel_nt64 = geti("JPB_PANEL_NT64", panel_nt64);
        panel_nt128 = geti("JPB_PANEL_NT128", panel_nt128);
        fork = geti("JPB_FORK", fork);
        wide = geti("JPB_WIDE", wide);
        schur_split = geti("JPB_SCHUR_SPLIT", schur_split);
    }
};
// symbolic options, overridable for tuning runs (JPB_LEAF, JPB_RELAX_SMALL, ...)
SymbolicOptions symbolic_options() {
    SymbolicOptions o;
    auto geti = [](const char* name, int def) { const char* v = getenv(name); return v ? atoi(v) : def; };
    auto getd = [](const char* name, double def) { const char* v = getenv(name); return v ? atof(v) : def; };
    o.leaf_size = geti("JPB_LEAF", o.leaf_size);
    o.relax_small = geti("JPB_RELAX_SMALL", o.relax_small);
    o.relax_mid = geti("JPB_RELAX_MID", o.relax_mid);
    o.relax_mid_frac = getd("JPB_RELAX_MID_FRAC", o.relax_mid_frac);
    o.relax_big = geti("JPB_RELAX_BIG", o.relax_big);
    o.relax_big_frac = getd("JPB_RELAX_BIG_FRAC", o.relax_big_frac);
    o.relax_any_frac = getd("JPB_RELAX_ANY_FRAC", o.relax_any_frac);
    return o;
}
const Tuning& tuning() {
    static Tuning t;
    return t;
}
#define FRONT_BIG (tuning().front_big)
#define FRONT_MID (tuning().front_mid)

// K1: vals[b, q] = sum_k coef[b, k] * mats_f[k, q]  (front-entry order), mats read once per
// slice of ASM_BCHUNK systems
constexpr int ASM_BCHUNK = 32;
// With `scal` (equilibration on) the factors s come from the assembled diagonal first and the
// entries leave the assembly already scaled.  Returns scal or nullptr.
const double* assemble_entries(jpb_state* st, cudaStream_t s, Pattern& P, int nsys, const cd* coef, cd* vals,
                               double* scal) {
    dim3 grid((unsigned)((P.S.nnz + 255) / 256), (unsigned)((nsys + ASM_BCHUNK - 1) / ASM_BCHUNK));
    if (st->equilibrate && scal) {
        dim3 g1((P.S.N + 255) / 256, nsys);
        diag_scale_coef_kernel<KMAX><<<g1, 256, 0, s>>>(P.S.N, P.K, P.d_mats_diag, coef, scal);
        launch_check(st);
        assemble_kernel<KMAX, true><<<grid, 256, (size_t)ASM_BCHUNK * P.K * 16, s>>>(
            P.S.nnz, P.K, nsys, P.d_mats_f, coef, vals, ASM_BCHUNK, P.d_ent_row, P.d_ent_col, scal, P.S.N);
        launch_check(st);
        return scal;
    }
    assemble_kernel<KMAX><<<grid, 256, (size_t)ASM_BCHUNK * P.K * 16, s>>>(P.S.nnz, P.K, nsys, P.d_mats_f, coef,
                                                                         vals, ASM_BCHUNK);
    launch_check(st);
    return nullptr;
}

// The launches of one level of the elimination tree touch disjoint fronts: the first goes to the
// caller's stream, the others to side streams that wait for the level to open (everything the
// caller's stream held before) and are joined before the next level.  Same kernels, same
// results; only the tails of the narrow launches overlap with the wide ones.
struct LevelFork {
    static constexpr int NSIDE = 3;
    jpb_state* st;
    cudaStream_t s;
    bool on;
    int k = 0;
    bool touched[NSIDE] = {false, false, false};
    LevelFork(jpb_state* st_, cudaStream_t s_, int nlaunches) : st(st_), s(s_), on(tuning().fork && nlaunches > 1) {
        if (on) {
            st->ensure_level_streams(NSIDE);
            cudaEvent_t open = st->ev();
            CK(cudaEventRecord(open, s));
            for (int i = 0; i < NSIDE && i < nlaunches - 1; ++i) {
                CK(cudaStreamWaitEvent(st->lvl[i], open, 0));
                touched[i] = true;
            }
        }
    }
    cudaStream_t next() {
        const int i = k++;
        if (!on || i == 0) return s;
        return st->lvl[(i - 1) % NSIDE];
    }
    void join() {
        if (!on) return;
        for (int i = 0; i < NSIDE; ++i)
            if (touched[i]) {
                cudaEvent_t done = st->ev();
                CK(cudaEventRecord(done, st->lvl[i]));
                CK(cudaStreamWaitEvent(s, done, 0));
            }
    }
};

// numeric factorisation of `nsys` systems into F.  ent_order: `data` is [nsys, nnz] in
// front-entry order (from assemble_entries) instead of the caller's CSC order.
void factor(jpb_state* st, cudaStream_t s0, Pattern& P, int nsys, cd* F, const cd* data, bool ent_order,
            int* status) {
    Span span(st, s0, PH_FACTOR);
    for (int l = 0; l < P.S.nlevels; ++l) {
        const int* lfl = P.d_flf + P.flevelptr[l];
        LevelFork lf_fork(st, s0, (int)P.fgroups[l].size());
        for (const Pattern::FactorGroup& g : P.fgroups[l]) {
            cudaStream_t s = lf_fork.next();
            const int64_t grid64 = (int64_t)g.count * nsys;
            if (grid64 > 0x7fffffff) throw InvalidArg("wave too large for one launch");
            const unsigned grid = (unsigned)grid64;
            const int* lf = lfl + g.first;
#define JPB_FACTOR(KERNEL, NT)                                                                       \
    if (ent_order) KERNEL<true, NT><<<grid, NT, g.smem, s>>>(P.sd, lf, nsys, F, data, status);       \
    else KERNEL<false, NT><<<grid, NT, g.smem, s>>>(P.sd, lf, nsys, F, data, status);
            if (g.kind == 2) {
                if (g.nt == 256) { JPB_FACTOR(factor_panel_kernel, 256) }
                else if (g.nt == 128) { JPB_FACTOR(factor_panel_kernel, 128) }
                else { JPB_FACTOR(factor_panel_kernel, 64) }
            } else if (g.kind == 1) {
                if (g.nt == 256) { JPB_FACTOR(factor_smem_kernel, 256) }
                else if (g.nt == 128) { JPB_FACTOR(factor_smem_kernel, 128) }
                else { JPB_FACTOR(factor_smem_kernel, 64) }
            } else {
                if (g.nt == 256 && tuning().tma) {
                    if (ent_order) factor_level_kernel<true, 256, true><<<grid, 256, GB_SMEM, s>>>(P.sd, lf, nsys, F, data, status);
                    else factor_level_kernel<false, 256, true><<<grid, 256, GB_SMEM, s>>>(P.sd, lf, nsys, F, data, status);
                }
                else if (g.nt == 256 && tuning().schur_split > 0 && g.maxm >= tuning().schur_split) {
                    // pivot panels here, contribution blocks in a launch of their own (several CTAs per front)
                    const bool wide = tuning().wide && grid <= (unsigned)st->sm_count;
                    if (wide) {
                        if (ent_order) factor_level_kernel<true, 512, false, false><<<grid, 512, 0, s>>>(P.sd, lf, nsys, F, data, status);
                        else factor_level_kernel<false, 512, false, false><<<grid, 512, 0, s>>>(P.sd, lf, nsys, F, data, status);
                    } else {
                        if (ent_order) factor_level_kernel<true, 256, false, false><<<grid, 256, 0, s>>>(P.sd, lf, nsys, F, data, status);
                        else factor_level_kernel<false, 256, false, false><<<grid, 256, 0, s>>>(P.sd, lf, nsys, F, data, status);
                    }
                    launch_check(st);
                    const int nbands = (g.maxm + SCHUR_BAND - 1) / SCHUR_BAND;
                    const int64_t g2 = grid64 * nbands;
                    if (g2 > 0x7fffffff) throw InvalidArg("wave too large for one launch");
                    schur_band_kernel<256><<<(unsigned)g2, 256, 0, s>>>(P.sd, lf, nsys, F, nbands);
                }
                else if (g.nt == 256 && tuning().wide && grid <= (unsigned)st->sm_count) { JPB_FACTOR(factor_level_kernel, 512) }
                else if (g.nt == 256) { JPB_FACTOR(factor_level_kernel, 256) }
                else if (g.nt == 128) { JPB_FACTOR(factor_level_kernel, 128) }
                else { JPB_FACTOR(factor_level_kernel, 64) }
            }
#undef JPB_FACTOR
            launch_check(st);
        }
        lf_fork.join();
    }
}

// equilibration helpers: s from the diagonal of the assembled values, values scaled in place,
// vectors scaled on the way into / out of the scaled system
double* equilibrate_values(jpb_state* st, cudaStream_t s, Pattern& P, int nsys, cd* vals, bool ent_order,
                           double* scal) {
    if (!st->equilibrate) return nullptr;
    const int N = P.S.N;
    const int64_t nnz = P.S.nnz;
    dim3 g1((N + 255) / 256, nsys);
    diag_scale_kernel<<<g1, 256, 0, s>>>(N, nnz, ent_order ? P.d_diag_ent : P.d_diag_csc, vals, scal);
    launch_check(st);
    dim3 g2((unsigned)((nnz + 255) / 256), nsys);
    scale_values_kernel<<<g2, 256, 0, s>>>(nnz, N, ent_order ? P.d_ent_row : P.d_csc_row,
                                           ent_order ? P.d_ent_col : P.d_csc_col, scal, vals);
    launch_check(st);
    return scal;
}
void scale_vectors(jpb_state* st, cudaStream_t s, Pattern& P, const double* scal, cd* v, int nsys, int nvec) {
    if (!scal) return;
    dim3 g((P.S.N + 255) / 256, nvec);
    scale_vec_kernel<<<g, 256, 0, s>>>(P.S.N, nsys, scal, v);
    launch_check(st);
}

// x(b, j) = A_b^-1 rhs(b, j)  (or A_b^-T)
constexpr int WSOLVE_NW = 8;     // systems (warps) per CTA of the warp-per-system solve kernels
void tri_solve(jpb_state* st, cudaStream_t s0, Pattern& P, const cd* F, bool trans, VecView vv, int phase) {
    Span span(st, s0, phase);
    const size_t smem = (size_t)P.max_s * sizeof(cd);
    const int64_t mult = (int64_t)vv.nsys * vv.nrhs;
    const int64_t wgroups = (mult + WSOLVE_NW - 1) / WSOLVE_NW;
    // fronts of a level are sorted largest first: the first wsolve[l][0] go to the CTA-per-front
    // kernels (by size class), the rest to the warp-per-system kernels
    auto level = [&](int l, bool fwd) {
        const int n = P.levelptr[l + 1] - P.levelptr[l];
        const int nabove = P.wsolve[l][0];
        const int* lf = P.d_levelfronts + P.levelptr[l];
        const int nbig = std::min(P.cls[l][0], nabove), nmid = std::min(P.cls[l][1], nabove - nbig);
        const int cnt[3] = {nbig, nmid, nabove - nbig - nmid};
        LevelFork lf_fork(st, s0, (cnt[0] > 0) + (cnt[1] > 0) + (cnt[2] > 0) + (n > nabove));
        int off = 0;
        for (int c = 0; c < 3; ++c) {
            if (!cnt[c]) continue;
            cudaStream_t s = lf_fork.next();
            const int64_t grid64 = (int64_t)cnt[c] * mult;
            if (grid64 > 0x7fffffff) throw InvalidArg("wave too large for one launch");
            const unsigned grid = (unsigned)grid64;
            const int* lfc = lf + off;
            const int nt = tuning().nt[c];
#define JPB_SOLVE(KERNEL, NT)                                                       \
    if (trans) KERNEL<true, NT><<<grid, NT, smem, s>>>(P.sd, lfc, F, vv);           \
    else KERNEL<false, NT><<<grid, NT, smem, s>>>(P.sd, lfc, F, vv);
            if (fwd) {
                if (nt == 256) { JPB_SOLVE(fwd_level_kernel, 256) }
                else if (nt == 128) { JPB_SOLVE(fwd_level_kernel, 128) }
                else { JPB_SOLVE(fwd_level_kernel, 64) }
            } else {
                if (nt == 256) { JPB_SOLVE(bwd_level_kernel, 256) }
                else if (nt == 128) { JPB_SOLVE(bwd_level_kernel, 128) }
                else { JPB_SOLVE(bwd_level_kernel, 64) }
            }
#undef JPB_SOLVE
            launch_check(st);
            off += cnt[c];
        }
        if (n > nabove) {
            const int64_t grid64 = (int64_t)(n - nabove) * wgroups;
            if (grid64 > 0x7fffffff) throw InvalidArg("wave too large for one launch");
            const unsigned grid = (unsigned)grid64;
            const int vstride = (P.wsolve[l][1] + 1) & ~1;
            const size_t wsmem = (size_t)WSOLVE_NW * (vstride + TW) * sizeof(cd);
            const int* lfw = lf + nabove;
            cudaStream_t s = lf_fork.next();
            if (fwd) {
                if (trans) fwd_warp_kernel<true, WSOLVE_NW><<<grid, WSOLVE_NW * 32, wsmem, s>>>(P.sd, lfw, F, vv, vstride);
                else fwd_warp_kernel<false, WSOLVE_NW><<<grid, WSOLVE_NW * 32, wsmem, s>>>(P.sd, lfw, F, vv, vstride);
            } else {
                if (trans) bwd_warp_kernel<true, WSOLVE_NW><<<grid, WSOLVE_NW * 32, wsmem, s>>>(P.sd, lfw, F, vv, vstride);
                else bwd_warp_kernel<false, WSOLVE_NW><<<grid, WSOLVE_NW * 32, wsmem, s>>>(P.sd, lfw, F, vv, vstride);
            }
            launch_check(st);
        }
        lf_fork.join();
    };
    for (int l = 0; l < P.S.nlevels; ++l) level(l, true);
    for (int l = P.S.nlevels - 1; l >= 0; --l) level(l, false);
}

// y(b,j) = [rhs(b,j) -] op(A_b) x(b,j)
void spmv(jpb_state* st, cudaStream_t s, Pattern& P, bool trans, const cd* vals, int64_t vals_sb,
          bool vals_ent_order, const cd* x, int64_t x_sb, int64_t x_sj, const cd* rhs, int64_t rhs_sb,
          int64_t rhs_sj, cd* y, int nsys, int nrhs, double* norms = nullptr, const int* active = nullptr) {
    Span span(st, s, PH_SPMV);
    const int* ptr = trans ? P.d_csc_ptr : P.d_csr_ptr;
    const int* idx = trans ? P.d_csc_row : P.d_csr_col;
    const int* pos = trans ? (vals_ent_order ? P.d_csc_pos_ent : nullptr)
                           : (vals_ent_order ? P.d_csr_pos_ent : P.d_csr_pos);
    dim3 grid((P.S.N + 127) / 128, (unsigned)(nsys * nrhs));
    if ((int64_t)nsys * nrhs > 65535) throw InvalidArg("too many vectors for one spmv launch");
    if (norms)
        spmv_kernel<true><<<grid, 128, 0, s>>>(P.S.N, P.S.nnz, ptr, idx, pos, vals, vals_sb, x, x_sb, x_sj, rhs,
                                               rhs_sb, rhs_sj, y, nsys, norms, active);
    else
        spmv_kernel<false><<<grid, 128, 0, s>>>(P.S.N, P.S.nnz, ptr, idx, pos, vals, vals_sb, x, x_sb, x_sj, rhs,
                                                rhs_sb, rhs_sj, y, nsys, norms, active);
    launch_check(st);
}

// solve + iterative refinement.  x receives the solution in the dense (j*nsys + b) layout.
void solve_refined(jpb_state* st, cudaStream_t s, Pattern& P, const cd* F, bool trans, const cd* vals,
                   int64_t vals_sb, bool vals_ent_order, const cd* rhs, int64_t rhs_sb, int64_t rhs_sj,
                   cd* x, cd* V, cd* r, cd* d, int nsys, int nrhs, int phase, int* status,
                   bool final_check) {
    VecView vv{rhs, rhs_sb, rhs_sj, x, V, nsys, nrhs, nullptr};
    tri_solve(st, s, P, F, trans, vv, phase);
    const int nvec = nsys * nrhs;
    const int N = P.S.N;
    double* norms = (double*)st->norms.get((size_t)nvec * 24);
    auto residual = [&](double* nrm, const int* act) {       // r = rhs - op(A) x  [+ backward-error norms]
        if (nrm) CK(cudaMemsetAsync(nrm, 0, (size_t)nvec * 24, s));
        spmv(st, s, P, trans, vals, vals_sb, vals_ent_order, x, 1, nsys, rhs, rhs_sb, rhs_sj, r, nsys, nrhs, nrm, act);
    };
    auto correct = [&](const int* act) {                     // x += op(A)^-1 r
        VecView v2{r, 1, nsys, d, V, nsys, nrhs, act};
        tri_solve(st, s, P, F, trans, v2, phase);
        dim3 grid((N + 255) / 256, nvec);
        axpy_kernel<<<grid, 256, 0, s>>>(N, x, d, act);
        launch_check(st);
    };
    // (1) unconditional steps (jpb_set_refinement; 1 by default).  The residual of the unrefined
    //     solution doubles as the acceptance test (status bit 2 above resid_tol0): no extra pass.
    for (int it = 0; it < st->refine; ++it) {
        const bool test = it == 0 && status && !final_check;
        residual(test ? norms : nullptr, nullptr);
        if (test) {
            residual_flag_kernel<<<(nvec + 127) / 128, 128, 0, s>>>(nvec, nsys, norms, st->resid_tol0 * st->resid_tol0, status);
            launch_check(st);
        }
        correct(nullptr);
    }
    // (2) without unconditional steps the residual of what we have is the acceptance test and
    //     decides, per system, whether ONE refinement step runs (jpb_set_refinement_threshold);
    //     systems that pass leave their blocks at once in the launches of that step
    if (st->refine == 0 && st->refine_omega > 0.0) {
        int* active = (int*)st->active.get((size_t)nvec * sizeof(int));
        residual(norms, nullptr);
        omega_active_kernel<<<(nvec + 127) / 128, 128, 0, s>>>(
            nvec, nsys, norms, st->refine_omega * st->refine_omega, st->refine_resid * st->refine_resid,
            st->resid_tol0 * st->resid_tol0, active, final_check ? nullptr : status, st->counters);
        launch_check(st);
        correct(active);
        // (3) optional correction-driven steps for the systems still moving (jpb_set_adaptive_refinement)
        for (int e = 0; e < st->refine_extra; ++e) {
            refine_active_kernel<<<nvec, 256, 0, s>>>(N, x, d, st->refine_tol, active, 0);
            launch_check(st);
            residual(nullptr, active);
            correct(active);
        }
    }
    if (status && final_check) {      // backward error of what is returned (jpb_solve)
        residual(norms, nullptr);
        residual_flag_kernel<<<(nvec + 127) / 128, 128, 0, s>>>(nvec, nsys, norms, st->resid_tol * st->resid_tol, status);
        launch_check(st);
    }
}

struct ModeShape {
    int64_t nsys, nrhs;        // factorisations, right-hand sides per factorisation
    int64_t rhs_sb, rhs_sj;    // strides (in vectors) of b for (system, rhs)
};

ModeShape decode_mode(int mode, int64_t nb0, int64_t nb1) {
    switch (mode) {
        case 0: return {1, 1, 0, 0};
        case 1: if (nb0 < 1) break; return {nb0, 1, 0, 0};
        case 2: if (nb0 < 1) break; return {1, nb0, 0, 1};
        case 3: if (nb0 < 1) break; return {nb0, 1, 1, 0};
        case 4: if (nb0 < 1 || nb1 < 1) break; return {nb1, nb0, 1, nb1};
        default: throw InvalidArg("Invalid `mode` argument in InnerState::solve().");
    }
    throw InvalidArg("batch sizes must be positive");
}

void summarise_timings(const jpb_state* st, double* t) {
    for (int i = 0; i < 10; ++i) t[i] = 0.0;
    for (const TimedSpan& sp : st->spans) {
        cudaEventSynchronize(sp.b);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) t[sp.phase] += ms;
    }
    t[8] = st->launches;
    t[9] = st->wave;
}

// core of the fused evaluation; every pointer is a device pointer
void eval_core(jpb_state* st, cudaStream_t s, Pattern& P, int64_t B, const cd* coef, const cd* scale,
               const double* ref, const double* w, int loss_type, int64_t F_total, cd* u_out, cd* y,
               double* fr, double* loss_sum, cd* ct_coef, cd* ct_s, int* status, bool adjoint,
               const int* seg = nullptr, int nseg = 1) {
    if (P.K == 0) throw StateError("components not registered for this pattern (jpb_register_components)");
    const int N = P.S.N;
    const int64_t nnz = P.S.nnz;
    const int K = P.K, R = P.R;
    const int64_t nchunks = (nnz + GRAD_CHUNK - 1) / GRAD_CHUNK;
    // per-system workspace bytes
    int64_t per = (P.S.front_slots + nnz + P.S.vec_slots + 5 * (int64_t)N) * 16 + nchunks * K * 16 + 256 + 8 * (int64_t)N;
    int64_t W = std::max<int64_t>(1, std::min<int64_t>(B, budget_bytes(st) / per));
    W = std::min<int64_t>(W, 65535);
    {   // Wave shaping.  Near the root of the elimination tree a level holds one or two fronts
        // per system, i.e. W or 2 W CTAs of the 256-thread class (2 resident per SM): a wave of
        // 160 systems puts a second CTA on 12 of the 148 SMs and the whole level waits for them,
        // so a wave that exceeds the SM count is cut to a multiple of it (measured at N = 1e5:
        // 1.52 ms per system at W = 148 against 1.61 at W = 160).  Then equal waves: 250 systems
        // with room for 148 run as 125 + 125, not 148 + 102 (the narrow launches cost the same).
        if (W > st->sm_count && P.S.max_s > 256) W = (W / st->sm_count) * st->sm_count;
        const int64_t nwaves = (B + W - 1) / W;
        W = std::min<int64_t>(W, (((B + nwaves - 1) / nwaves) + 7) & ~(int64_t)7);
    }
    st->wave = (double)W;
    cd* F = (cd*)st->F.get((size_t)W * P.S.front_slots * 16);
    cd* vals = (cd*)st->vals.get((size_t)W * nnz * 16);
    cd* V = (cd*)st->V.get((size_t)W * P.S.vec_slots * 16);
    cd* vb = (cd*)st->vb.get((size_t)W * N * 16);
    cd* vx = (cd*)st->vx.get((size_t)W * N * 16);
    cd* vr = (cd*)st->vr.get((size_t)W * N * 16);
    cd* vd = (cd*)st->vd.get((size_t)W * N * 16);
    cd* vlam = (cd*)st->vlam.get((size_t)W * N * 16);
    double* scal_all = (double*)st->scal.get((size_t)W * N * 8);
    cd* partial = adjoint ? (cd*)st->partial.get((size_t)nchunks * W * K * 16) : nullptr;
    // small per-wave array ybar [W,R]; loss terms of the whole batch [B] (summed per segment
    // = per parameter candidate at the end)
    cd* ybar = (cd*)st->small.get((size_t)W * R * 16 + 64);
    double* loss_term = adjoint ? (double*)st->lterm.get((size_t)B * 8) : nullptr;
    if (status) CK(cudaMemsetAsync(status, 0, sizeof(int) * B, s));

    // One pipeline (assemble -> factor -> solve -> read-out [-> loss -> adjoint solve -> gradient])
    // over the systems [g0, g0 + n) of the batch, which sit at [l0, l0 + n) of the wave buffers.
    auto run_chunk = [&](cudaStream_t sc, int64_t g0, int64_t l0, int n) {
        cd* Fc = F + l0 * P.S.front_slots;
        cd* valsc = vals + l0 * nnz;
        cd* Vc = V + l0 * P.S.vec_slots;
        cd *vbc = vb + l0 * N, *vxc = vx + l0 * N, *vrc = vr + l0 * N, *vdc = vd + l0 * N,
           *vlamc = vlam + l0 * N;
        const double* scal = nullptr;
        {
            Span span(st, sc, PH_FACTOR);
            scal = assemble_entries(st, sc, P, n, coef + g0 * K, valsc, scal_all + l0 * N);
        }
        factor(st, sc, P, n, Fc, valsc, true, status ? status + g0 : nullptr);
        {
            dim3 grid((N + 255) / 256, n);
            scale_rhs_kernel<<<grid, 256, 0, sc>>>(N, P.d_rhs_vec, scale + g0, vbc);
            launch_check(st);
            scale_vectors(st, sc, P, scal, vbc, n, n);
        }
        solve_refined(st, sc, P, Fc, false, valsc, 1, true, vbc, 1, n, vxc, Vc, vrc, vdc, n, 1, PH_FWD,
                      status ? status + g0 : nullptr, false);
        scale_vectors(st, sc, P, scal, vxc, n, n);      // back to the caller's units
        {
            Span span(st, sc, PH_READOUT);
            readout_kernel<<<n, 256, 0, sc>>>(N, R, P.d_Q, vxc, y + g0 * R);
            launch_check(st);
            if (u_out) CK(cudaMemcpyAsync(u_out + g0 * N, vxc, (size_t)n * N * 16, cudaMemcpyDeviceToDevice, sc));
        }
        if (!adjoint) return;
        {
            Span span(st, sc, PH_READOUT);
            loss_kernel<<<(n + 127) / 128, 128, 0, sc>>>(n, R, y + g0 * R, w,
                                                         ref ? ref + (loss_type >= 5 ? 2 : 1) * g0 : nullptr,
                                                         loss_type, 1.0 / (double)F_total, fr + g0,
                                                         loss_term + g0, ybar + l0 * R);
            launch_check(st);
            dim3 grid((N + 255) / 256, n);
            cotangent_kernel<<<grid, 256, 0, sc>>>(N, R, P.d_Q, ybar + l0 * R, vbc);
            launch_check(st);
            scale_vectors(st, sc, P, scal, vbc, n, n);
        }
        solve_refined(st, sc, P, Fc, true, valsc, 1, true, vbc, 1, n, vlamc, Vc, vrc, vdc, n, 1, PH_ADJ,
                      status ? status + g0 : nullptr, false);
        scale_vectors(st, sc, P, scal, vlamc, n, n);
        {
            Span span(st, sc, PH_GRAD);
            const int ngroups = (n + GRAD_SYS - 1) / GRAD_SYS;
            cd* part = partial + nchunks * l0 * K;
            grad_kernel<<<(unsigned)(nchunks * ngroups), GRAD_THREADS, GRAD_SMEM, sc>>>(
                nnz, N, K, P.d_csc_row, P.d_csc_col, P.d_mats, vlamc, vxc, part, n, ngroups);
            launch_check(st);
            int64_t nbk = (int64_t)n * K;
            grad_reduce_kernel<<<(unsigned)((nbk + 127) / 128), 128, 0, sc>>>((int)nchunks, nbk, part,
                                                                              ct_coef + g0 * K);
            launch_check(st);
            readout_kernel<<<n, 256, 0, sc>>>(N, 1, P.d_rhs_vec, vlamc, ct_s + g0);
            launch_check(st);
        }
    };

    // The systems of a wave are independent, so a wave is cut into `nstreams` sub-waves that run
    // the whole pipeline on their own streams: the narrow launches near the root of the
    // elimination tree (a few hundred CTAs, long pivot chains) and the bandwidth-bound solve
    // sweeps of one sub-wave fill the SMs the others leave idle.
    for (int64_t b0 = 0; b0 < B; b0 += W) {
        const int nw = (int)std::min<int64_t>(W, B - b0);
        const int ns = (int)std::max<int64_t>(1, std::min<int64_t>(st->nstreams, nw / 32));
        if (ns == 1) {
            run_chunk(s, b0, 0, nw);
        } else {
            st->ensure_streams(ns - 1);
            cudaEvent_t fork = st->ev();
            CK(cudaEventRecord(fork, s));
            const int per = (nw + ns - 1) / ns;
            for (int c = 0; c < ns; ++c) {
                const int l0 = c * per, n = std::min(per, nw - l0);
                if (n <= 0) break;
                cudaStream_t sc = c == 0 ? s : st->aux[c - 1];
                if (c > 0) CK(cudaStreamWaitEvent(sc, fork, 0));
                run_chunk(sc, b0 + l0, l0, n);
                if (c > 0) {
                    cudaEvent_t join = st->ev();
                    CK(cudaEventRecord(join, sc));
                    CK(cudaStreamWaitEvent(s, join, 0));
                }
            }
        }
    }
    if (adjoint && loss_sum) {
        seg_sum_kernel<<<nseg, 256, 0, s>>>((int)B, loss_term, seg, loss_sum);
        launch_check(st);
    }
}

}  // namespace

// ============================================================================== C-ABI
extern "C" {

int jpb_create(jpb_state** out, int device) {
    if (!out) return JPB_EINVAL;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0 || device < 0 || device >= n) return JPB_ECUDA;
    auto st = std::make_unique<jpb_state>();
    st->device = device;
    if (const char* v = getenv("JPB_STREAMS")) st->nstreams = std::max(1, std::min(16, atoi(v)));
    if (cudaSetDevice(device) != cudaSuccess) return JPB_ECUDA;
    {
        int n_sm = 0;
        if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && n_sm > 0)
            st->sm_count = n_sm;
    }
    if (cudaStreamCreateWithFlags(&st->stream, cudaStreamNonBlocking) != cudaSuccess) return JPB_ECUDA;
    if (cudaMalloc(&st->counters, 16) != cudaSuccess || cudaMemset(st->counters, 0, 16) != cudaSuccess) return JPB_ECUDA;
    if (cudaFuncSetAttribute(grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GRAD_SMEM) !=
        cudaSuccess)
        return JPB_ECUDA;
    {
        const int smem_cap = (112 * smem_ld(112) + TW * TW) * (int)sizeof(cd);
        bool ok = true;
        auto set = [&](auto kernel) {
            ok &= cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap) == cudaSuccess;
            ok &= cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       cudaSharedmemCarveoutMaxShared) == cudaSuccess;
        };
        set(factor_smem_kernel<true, 256>); set(factor_smem_kernel<false, 256>);
        set(factor_smem_kernel<true, 128>); set(factor_smem_kernel<false, 128>);
        set(factor_smem_kernel<true, 64>); set(factor_smem_kernel<false, 64>);
        if (!ok) return JPB_ECUDA;
        // the CTA-per-front solve kernels hold one front-local vector in dynamic shared memory;
        // the attribute is per function and device, so it is raised once to the cap add_mat
        // enforces (a later, smaller pattern must not lower it for an earlier one)
        auto setv = [&](auto kernel) {
            ok &= cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) == cudaSuccess;
        };
        auto setg = [&](auto kernel, int bytes) {
            ok &= cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess;
        };
        for (int bytes : {204 * 1024}) {
            setg(factor_panel_kernel<true, 256>, bytes); setg(factor_panel_kernel<false, 256>, bytes);
            setg(factor_panel_kernel<true, 128>, bytes); setg(factor_panel_kernel<false, 128>, bytes);
            setg(factor_panel_kernel<true, 64>, bytes); setg(factor_panel_kernel<false, 64>, bytes);
        }
        setg(factor_level_kernel<true, 256, true>, GB_SMEM);
        setg(factor_level_kernel<false, 256, true>, GB_SMEM);
        setg(selftest_cta_gemm_kernel, GB_SMEM);
        setv(fwd_level_kernel<false, 256>); setv(fwd_level_kernel<true, 256>);
        setv(bwd_level_kernel<false, 256>); setv(bwd_level_kernel<true, 256>);
        setv(fwd_level_kernel<false, 128>); setv(fwd_level_kernel<true, 128>);
        setv(bwd_level_kernel<false, 128>); setv(bwd_level_kernel<true, 128>);
        setv(fwd_level_kernel<false, 64>); setv(fwd_level_kernel<true, 64>);
        setv(bwd_level_kernel<false, 64>); setv(bwd_level_kernel<true, 64>);
        if (!ok) return JPB_ECUDA;
    }
    *out = st.release();
    return JPB_OK;
}

void jpb_destroy(jpb_state* st) {
    if (!st) return;
    cudaSetDevice(st->device);
    cudaDeviceSynchronize();
    delete st;
}

const char* jpb_last_error(const jpb_state* st) { return st ? st->err.c_str() : "null state"; }

int jpb_set_memory_budget(jpb_state* st, int64_t bytes) {
    if (!st || bytes < 0) return JPB_EINVAL;
    st->budget = bytes;
    return JPB_OK;
}

int jpb_set_streams(jpb_state* st, int n) {
    if (!st || n < 1 || n > 16) return fail(st, JPB_EINVAL, "streams must be in [1, 16]");
    st->nstreams = n;
    return JPB_OK;
}

int jpb_set_refinement(jpb_state* st, int steps) {
    if (!st || steps < 0 || steps > 8) return fail(st, JPB_EINVAL, "refinement steps must be in [0, 8]");
    st->refine = steps;
    return JPB_OK;
}

int jpb_set_equilibration(jpb_state* st, int on) {
    if (!st) return JPB_EINVAL;
    st->equilibrate = on ? 1 : 0;
    return JPB_OK;
}

int jpb_set_refinement_threshold(jpb_state* st, double omega, double rel_residual) {
    if (!st || !(omega >= 0.0) || !(rel_residual >= 0.0))
        return fail(st, JPB_EINVAL, "refinement thresholds must be >= 0");
    st->refine_omega = omega;
    st->refine_resid = rel_residual > 0.0 ? rel_residual : 1e300;
    return JPB_OK;
}

int jpb_last_refined(jpb_state* st, int64_t* refined, int64_t* solved) {
    JPB_TRY(st)
    if (!st || !refined || !solved) throw InvalidArg("null argument");
    CK(cudaSetDevice(st->device));
    unsigned long long h[2] = {0, 0};
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h, st->counters, 16, cudaMemcpyDeviceToHost));
    *refined = (int64_t)h[0];
    *solved = (int64_t)h[1];
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_set_adaptive_refinement(jpb_state* st, int extra_steps, double tol) {
    if (!st || extra_steps < 0 || extra_steps > 8 || !(tol >= 0.0))
        return fail(st, JPB_EINVAL, "adaptive refinement: steps in [0, 8], tol >= 0");
    st->refine_extra = extra_steps;
    st->refine_tol = tol;
    return JPB_OK;
}

int jpb_set_residual_tolerance(jpb_state* st, double tol_final, double tol_unrefined) {
    if (!st || !(tol_final > 0.0) || !(tol_unrefined > 0.0))
        return fail(st, JPB_EINVAL, "residual tolerances must be positive");
    st->resid_tol = tol_final;
    st->resid_tol0 = tol_unrefined;
    return JPB_OK;
}

int jpb_add_mat(jpb_state* st, int64_t N, int64_t nnz, const void* indices, const void* indptr,
                const void* indices_T, const void* indptr_T, const void* perm, int index_bytes,
                int* solver_num) {
    (void)indptr_T;   // the reference passes the CSC indptr here (Sparse.py:34); never used
    JPB_TRY(st)
    if (!st) throw InvalidArg("null state");
    if (N <= 0 || N > INT32_MAX) throw InvalidArg("add_mat: N out of range");
    if (nnz <= 0 || nnz > INT32_MAX) throw InvalidArg("add_mat: nnz out of range");
    if (!indices || !indptr) throw InvalidArg("add_mat: null index arrays");
    if (index_bytes != 4 && index_bytes != 8) throw InvalidArg("add_mat: index width must be 4 or 8 bytes");
    CK(cudaSetDevice(st->device));
    auto get = [&](const void* p, int64_t i) -> int64_t {
        return index_bytes == 4 ? (int64_t)((const int32_t*)p)[i] : ((const int64_t*)p)[i];
    };
    std::vector<int64_t> ap(N + 1), ai(nnz);
    for (int64_t i = 0; i <= N; ++i) ap[i] = get(indptr, i);
    for (int64_t i = 0; i < nnz; ++i) ai[i] = get(indices, i);
    if (ap[0] != 0 || ap[N] != nnz) throw InvalidArg("add_mat: indptr does not match nnz");
    for (int64_t j = 0; j < N; ++j) {
        if (ap[j + 1] < ap[j]) throw InvalidArg("add_mat: indptr not monotone");
        for (int64_t e = ap[j]; e < ap[j + 1]; ++e) {
            if (ai[e] < 0 || ai[e] >= N) throw InvalidArg("add_mat: row index out of range");
            if (e > ap[j] && ai[e] <= ai[e - 1]) throw InvalidArg("add_mat: row indices must be strictly increasing within a column");
        }
    }
    auto P = std::make_unique<Pattern>();
    analyze((int)N, ap.data(), ai.data(), symbolic_options(), P->S);
    Symbolic& S = P->S;

    // CSR view (exact transpose of the CSC pattern) with CSR->CSC positions
    std::vector<int> csc_ptr(N + 1), csc_row(nnz), csc_col(nnz), csr_ptr(N + 1, 0), csr_col(nnz), csr_pos(nnz);
    for (int64_t j = 0; j <= N; ++j) csc_ptr[j] = (int)ap[j];
    for (int64_t j = 0; j < N; ++j)
        for (int64_t e = ap[j]; e < ap[j + 1]; ++e) {
            csc_row[e] = (int)ai[e];
            csc_col[e] = (int)j;
            csr_ptr[ai[e] + 1]++;
        }
    for (int64_t i = 0; i < N; ++i) csr_ptr[i + 1] += csr_ptr[i];
    {
        std::vector<int> pos(csr_ptr.begin(), csr_ptr.end() - 1);
        for (int64_t j = 0; j < N; ++j)
            for (int64_t e = ap[j]; e < ap[j + 1]; ++e) {
                int q = pos[ai[e]]++;
                csr_col[q] = (int)j;
                csr_pos[q] = (int)e;
            }
    }
    if (perm)
        for (int64_t q = 0; q < nnz; ++q)
            if (get(perm, q) != csr_pos[q]) throw InvalidArg("add_mat: `permutation` is not the CSR->CSC map of the pattern");
    if (indices_T)
        for (int64_t q = 0; q < nnz; ++q)
            if (get(indices_T, q) != csr_col[q]) throw InvalidArg("add_mat: `indices_T` is not the CSR column index of the pattern");

    std::vector<int> entpos(nnz);
    for (int64_t q = 0; q < nnz; ++q) entpos[S.ent_src[q]] = (int)q;
    std::vector<int> csr_pos_ent(nnz), csc_pos_ent(nnz);
    for (int64_t q = 0; q < nnz; ++q) {
        csr_pos_ent[q] = entpos[csr_pos[q]];
        csc_pos_ent[q] = entpos[q];
    }

    std::vector<FrontDesc> fr(S.nfronts);
    for (int f = 0; f < S.nfronts; ++f) {
        FrontDesc& d = fr[f];
        d.off = S.f_off[f];
        d.entptr = S.f_entptr[f];
        d.s = S.f_s[f];
        d.k = S.f_k[f];
        d.ld = S.f_ld[f];
        d.nent = (int)(S.f_entptr[f + 1] - S.f_entptr[f]);
        d.rowptr = S.f_rowptr[f];
        d.voff = S.f_voff[f];
        d.childptr = S.f_childptr[f];
        d.nchild = S.f_childptr[f + 1] - S.f_childptr[f];
        d.reloff = (int)std::max<int64_t>(S.f_reloff[f], 0);
        d.parent = S.f_parent[f];
        d.npan = 0;
        d.invoff = 0;
    }
    if (!S.relmap.empty() && (int64_t)S.relmap.size() > INT32_MAX) throw InvalidArg("add_mat: pattern too large");
    P->levelptr = S.levelptr;
    P->max_s = S.max_s;
    P->cls.resize(S.nlevels);
    for (int l = 0; l < S.nlevels; ++l) {
        int nbig = 0, nmid = 0;
        for (int q = S.levelptr[l]; q < S.levelptr[l + 1]; ++q) {
            int sz = S.f_s[S.levelfronts[q]];
            if (sz > FRONT_BIG) ++nbig;
            else if (sz > FRONT_MID) ++nmid;
        }
        P->cls[l] = {nbig, nmid};
    }
    P->wsolve.resize(S.nlevels);
    for (int l = 0; l < S.nlevels; ++l) {
        int nabove = 0, smax = 0;
        for (int q = S.levelptr[l]; q < S.levelptr[l + 1]; ++q) {
            const int sz = S.f_s[S.levelfronts[q]];
            if (sz > tuning().wsolve_max) ++nabove;
            else smax = std::max(smax, sz);
        }
        P->wsolve[l] = {nabove, smax};
    }
    // Factor launch groups.  Every level gets its own ordering of the fronts for the numeric
    // phase (the solve kernels keep S.levelfronts, sorted by order): first the fronts handled in
    // global memory (largest first, split by CTA width), then the fronts whose whole s x s block
    // lives in shared memory (legacy class, JPB_PANEL_MAX=0), then the fronts whose PIVOT PANELS
    // live in shared memory, bucketed by their shared-memory footprint.
    P->fgroups.assign(S.nlevels, {});
    std::vector<int> flf;
    flf.reserve(S.nfronts);
    P->flevelptr.assign(S.nlevels + 1, 0);
    {
        const Tuning& tn = tuning();
        struct Item { int f, kind, nt, smem; };
        for (int l = 0; l < S.nlevels; ++l) {
            std::vector<Item> items;
            for (int q = S.levelptr[l]; q < S.levelptr[l + 1]; ++q) {
                const int f = S.levelfronts[q], sz = S.f_s[f], kk = S.f_k[f];
                const int64_t pbytes = (int64_t)panel_elems(sz, kk) * (int64_t)sizeof(cd);
                Item it{f, 0, 0, 0};
                if (tn.panel_max > 0 && pbytes <= (int64_t)tn.panel_max * 1024) {
                    it.kind = 2;
                    it.nt = sz <= tn.panel_nt64 ? 64 : (sz <= tn.panel_nt128 ? 128 : 256);
                    it.smem = (int)(((pbytes + 4095) / 4096) * 4096);        // 4 KB buckets
                } else if (sz <= tn.smem_max) {
                    it.kind = 1;
                    it.nt = sz <= tn.smem_nt64 ? 64 : (sz <= tn.smem_nt128 ? 128 : 256);
                    const int top = ((sz + tn.smem_step - 1) / tn.smem_step) * tn.smem_step;   // group's largest order
                    it.smem = (top * smem_ld(top) + TW * TW) * (int)sizeof(cd);
                } else {
                    const int c = sz > tn.front_big ? 0 : (sz > tn.front_mid ? 1 : 2);
                    it.kind = 0;
                    it.nt = tn.nt[c];
                }
                items.push_back(it);
            }
            std::stable_sort(items.begin(), items.end(), [&](const Item& a, const Item& b) {
                if (a.kind != b.kind) return a.kind < b.kind;
                if (a.kind == 0) return false;                       // keep "largest first"
                if (a.smem != b.smem) return a.smem > b.smem;
                return a.nt > b.nt;
            });
            P->flevelptr[l] = (int)flf.size();
            for (size_t q = 0; q < items.size(); ++q) {
                const Item& it = items[q];
                auto& groups = P->fgroups[l];
                if (groups.empty() || groups.back().kind != it.kind || groups.back().nt != it.nt ||
                    groups.back().smem != it.smem)
                    groups.push_back({(int)q, 0, it.smem, it.nt, it.kind, 0});
                groups.back().count++;
                groups.back().maxm = std::max(groups.back().maxm, S.f_s[it.f] - S.f_k[it.f]);
                flf.push_back(it.f);
            }
        }
        P->flevelptr[S.nlevels] = (int)flf.size();
    }
    P->d_flf = upload(flf);
    // inverse extend-add maps (parent contribution row -> child contribution row) and the
    // number of child rows that land on parent pivots; the factor kernel for large fronts
    // gathers the children's Schur blocks with them while it writes its own (no zero fill,
    // no scatter read-modify-write of the contribution block)
    std::vector<int> cbinv;
    for (int c = 0; c < S.nfronts; ++c) {
        const int p = S.f_parent[c];
        if (p < 0) continue;
        const int kp = S.f_k[p], mp = S.f_s[p] - kp, mc = S.f_s[c] - S.f_k[c];
        fr[c].invoff = (int)cbinv.size();
        cbinv.resize(cbinv.size() + mp, -1);
        int npan = 0;
        for (int i = 0; i < mc; ++i) {
            const int r = S.relmap[S.f_reloff[c] + i];
            if (r < kp) {
                if (npan != i) throw InvalidArg("add_mat: extend-add map is not pivot-rows-first");
                ++npan;
            }
            else cbinv[fr[c].invoff + (r - kp)] = i;
        }
        fr[c].npan = npan;
    }
    if (cbinv.size() > (size_t)INT32_MAX) throw InvalidArg("add_mat: pattern too large");
    P->d_cbinv = upload(cbinv);
    P->d_fr = upload(fr);
    P->d_rowidx_orig = upload(S.rowidx_orig);
    P->d_children = upload(S.children);
    P->d_relmap = upload(S.relmap);
    P->d_ent_src = upload(S.ent_src);
    P->d_ent_slot = upload(S.ent_slot);
    P->d_levelfronts = upload(S.levelfronts);
    P->d_csc_ptr = upload(csc_ptr);
    P->d_csc_row = upload(csc_row);
    P->d_csc_col = upload(csc_col);
    P->d_csc_pos_ent = upload(csc_pos_ent);
    P->d_csr_ptr = upload(csr_ptr);
    P->d_csr_col = upload(csr_col);
    P->d_csr_pos = upload(csr_pos);
    P->d_csr_pos_ent = upload(csr_pos_ent);
    {
        std::vector<int> diag_csc(N, -1), diag_ent(N, -1), ent_row(nnz), ent_col(nnz);
        for (int64_t e = 0; e < nnz; ++e)
            if (csc_row[e] == csc_col[e]) { diag_csc[csc_col[e]] = (int)e; diag_ent[csc_col[e]] = entpos[e]; }
        for (int64_t q = 0; q < nnz; ++q) { ent_row[q] = csc_row[S.ent_src[q]]; ent_col[q] = csc_col[S.ent_src[q]]; }
        P->d_diag_csc = upload(diag_csc);
        P->d_diag_ent = upload(diag_ent);
        P->d_ent_row = upload(ent_row);
        P->d_ent_col = upload(ent_col);
    }
    P->sd = SymDev{P->d_fr, P->d_rowidx_orig, P->d_children, P->d_relmap, P->d_ent_src, P->d_ent_slot, P->d_cbinv,
                   S.front_slots, S.vec_slots, S.nnz, S.N};
    // the dynamic shared memory of the solve kernels holds one front-local vector
    size_t smem = (size_t)S.max_s * sizeof(cd);
    if (smem > 200 * 1024) throw InvalidArg("add_mat: largest front exceeds the solve kernels' shared memory");
    st->pats.push_back(std::move(P));
    if (solver_num) *solver_num = (int)st->pats.size() - 1;
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_symbolic_info(const jpb_state* st, int solver_num, double* info) {
    jpb_state* s = const_cast<jpb_state*>(st);
    JPB_TRY(s)
    Pattern& P = pattern(s, solver_num);
    if (!info) throw InvalidArg("null output");
    const Symbolic& S = P.S;
    info[0] = S.N; info[1] = (double)S.nnz; info[2] = (double)S.nnz_lu; info[3] = S.flops;
    info[4] = (double)S.front_slots; info[5] = S.nfronts; info[6] = S.nlevels; info[7] = S.max_s;
    info[8] = S.max_k; info[9] = (double)S.vec_slots;
    return JPB_OK;
    JPB_CATCH(s)
}

int jpb_get_ordering(const jpb_state* st, int solver_num, int32_t* perm) {
    jpb_state* s = const_cast<jpb_state*>(st);
    JPB_TRY(s)
    Pattern& P = pattern(s, solver_num);
    if (!perm) throw InvalidArg("null output");
    for (int i = 0; i < P.S.N; ++i) perm[i] = P.S.perm[i];
    return JPB_OK;
    JPB_CATCH(s)
}

int jpb_solve(jpb_state* st, int solver_num, int mode, int transpose, const double* data,
              const double* b, int64_t nb0, int64_t nb1, double* out, int* status) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    if (!data || !b || !out) throw InvalidArg("solve: null array");
    ModeShape ms = decode_mode(mode, nb0, nb1);
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = st->stream;
    const int N = P.S.N;
    const int64_t nnz = P.S.nnz;
    const bool trans = transpose != 0;
    // rhs chunk: bound the front-local vector workspace
    const int64_t JC = std::min<int64_t>(ms.nrhs, 64);
    int64_t per = (P.S.front_slots + nnz) * 16 + JC * (P.S.vec_slots + 4 * (int64_t)N) * 16 + 64 + 8 * (int64_t)N;
    int64_t W = std::max<int64_t>(1, std::min<int64_t>(ms.nsys, budget_bytes(st) / per));
    while (W * JC > 65535) --W;
    st->wave = (double)W;
    cd* F = (cd*)st->F.get((size_t)W * P.S.front_slots * 16);
    cd* dat = (cd*)st->data.get((size_t)W * nnz * 16);
    cd* V = (cd*)st->V.get((size_t)W * JC * P.S.vec_slots * 16);
    cd* vb = (cd*)st->vb.get((size_t)W * JC * N * 16);
    cd* vx = (cd*)st->vx.get((size_t)W * JC * N * 16);
    cd* vr = (cd*)st->vr.get((size_t)W * JC * N * 16);
    cd* vd = (cd*)st->vd.get((size_t)W * JC * N * 16);
    double* scal_buf = (double*)st->scal.get((size_t)W * N * 8);
    int* dstat = (int*)st->status.get((size_t)ms.nsys * sizeof(int));
    CK(cudaMemsetAsync(dstat, 0, (size_t)ms.nsys * sizeof(int), s));
    const cd* hb = (const cd*)b;
    cd* hout = (cd*)out;
    for (int64_t i0 = 0; i0 < ms.nsys; i0 += W) {
        const int nw = (int)std::min<int64_t>(W, ms.nsys - i0);
        {
            Span span(st, s, PH_H2D);
            CK(cudaMemcpyAsync(dat, (const cd*)data + i0 * nnz, (size_t)nw * nnz * 16, cudaMemcpyDefault, s));
        }
        const double* scal = equilibrate_values(st, s, P, nw, dat, false, scal_buf);
        factor(st, s, P, nw, F, dat, false, dstat + i0);
        for (int64_t j0 = 0; j0 < ms.nrhs; j0 += JC) {
            const int nj = (int)std::min<int64_t>(JC, ms.nrhs - j0);
            // stage rhs(b, j) densely as vb[(j * nw + b) * N]
            int64_t sb, sj;
            {
                Span span(st, s, PH_H2D);
                if (ms.rhs_sb == 0 && ms.rhs_sj == 0 && !scal) {          // one shared vector (modes 0, 1)
                    CK(cudaMemcpyAsync(vb, hb, (size_t)N * 16, cudaMemcpyDefault, s));
                    sb = 0; sj = 0;
                } else if (ms.rhs_sb == 0 && ms.rhs_sj == 0) {
                    // ... which every system sees in its own scaled units: one copy per system
                    CK(cudaMemcpyAsync(vr, hb, (size_t)N * 16, cudaMemcpyDefault, s));
                    dim3 g((N + 255) / 256, nw);
                    broadcast_vec_kernel<<<g, 256, 0, s>>>(N, vr, vb);
                    launch_check(st);
                    sb = 1; sj = nw;
                } else {
                    for (int j = 0; j < nj; ++j) {
                        const cd* src = hb + ((j0 + j) * ms.rhs_sj + i0 * ms.rhs_sb) * N;
                        CK(cudaMemcpyAsync(vb + (int64_t)j * nw * N, src, (size_t)nw * N * 16,
                                           cudaMemcpyDefault, s));
                    }
                    sb = 1; sj = nw;
                    if (ms.rhs_sb == 0) sb = 0;
                }
            }
            scale_vectors(st, s, P, scal, vb, nw, nw * nj);
            solve_refined(st, s, P, F, trans, dat, 1, false, vb, sb, sj, vx, V, vr, vd, nw, nj,
                          trans ? PH_ADJ : PH_FWD, dstat + i0, true);
            scale_vectors(st, s, P, scal, vx, nw, nw * nj);
            {
                Span span(st, s, PH_D2H);
                for (int j = 0; j < nj; ++j)
                    CK(cudaMemcpyAsync(hout + ((j0 + j) * ms.nsys + i0) * N, vx + (int64_t)j * nw * N,
                                       (size_t)nw * N * 16, cudaMemcpyDefault, s));
            }
        }
    }
    std::vector<int> hstat(ms.nsys);
    CK(cudaMemcpyAsync(hstat.data(), dstat, (size_t)ms.nsys * sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    int bad = 0;
    for (int64_t i = 0; i < ms.nsys; ++i) {
        if (status) status[i] = hstat[i];
        bad |= hstat[i];
    }
    if (bad) return fail(st, JPB_ESINGULAR, "solve: %s", status_message(bad));
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_matvec(jpb_state* st, int solver_num, int mode, int transpose, const double* mat,
               const double* vec, int64_t nb0, int64_t nb1, double* out) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    if (!mat || !vec || !out) throw InvalidArg("matvec: null array");
    ModeShape ms = decode_mode(mode, nb0, nb1);
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = st->stream;
    const int N = P.S.N;
    const int64_t nnz = P.S.nnz;
    const int64_t JC = std::min<int64_t>(ms.nrhs, 64);
    int64_t per = nnz * 16 + JC * 2 * (int64_t)N * 16 + 64;
    int64_t W = std::max<int64_t>(1, std::min<int64_t>(ms.nsys, budget_bytes(st) / per));
    while (W * JC > 65535) --W;
    cd* dat = (cd*)st->data.get((size_t)W * nnz * 16);
    cd* vb = (cd*)st->vb.get((size_t)W * JC * N * 16);
    cd* vx = (cd*)st->vx.get((size_t)W * JC * N * 16);
    const cd* hv = (const cd*)vec;
    cd* hout = (cd*)out;
    for (int64_t i0 = 0; i0 < ms.nsys; i0 += W) {
        const int nw = (int)std::min<int64_t>(W, ms.nsys - i0);
        CK(cudaMemcpyAsync(dat, (const cd*)mat + i0 * nnz, (size_t)nw * nnz * 16, cudaMemcpyDefault, s));
        for (int64_t j0 = 0; j0 < ms.nrhs; j0 += JC) {
            const int nj = (int)std::min<int64_t>(JC, ms.nrhs - j0);
            int64_t sb, sj;
            if (ms.rhs_sb == 0 && ms.rhs_sj == 0) {
                CK(cudaMemcpyAsync(vb, hv, (size_t)N * 16, cudaMemcpyDefault, s));
                sb = 0; sj = 0;
            } else {
                for (int j = 0; j < nj; ++j)
                    CK(cudaMemcpyAsync(vb + (int64_t)j * nw * N, hv + ((j0 + j) * ms.rhs_sj + i0 * ms.rhs_sb) * N,
                                       (size_t)nw * N * 16, cudaMemcpyDefault, s));
                sb = ms.rhs_sb == 0 ? 0 : 1; sj = nw;
            }
            spmv(st, s, P, transpose != 0, dat, 1, false, vb, sb, sj, nullptr, 0, 0, vx, nw, nj);
            for (int j = 0; j < nj; ++j)
                CK(cudaMemcpyAsync(hout + ((j0 + j) * ms.nsys + i0) * N, vx + (int64_t)j * nw * N,
                                   (size_t)nw * N * 16, cudaMemcpyDefault, s));
        }
    }
    CK(cudaStreamSynchronize(s));
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_assemble(jpb_state* st, int64_t nnz, int K, int64_t B, const double* mats, const double* coef,
                 double* data) {
    JPB_TRY(st)
    if (!st) throw InvalidArg("null state");
    if (!mats || !coef || !data) throw InvalidArg("assemble: null array");
    if (nnz <= 0 || B <= 0 || K <= 0 || K > KMAX) throw InvalidArg("assemble: bad sizes (K must be <= 32)");
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = st->stream;
    double* dm = (double*)st->vals.get((size_t)K * nnz * 8);
    cd* dc = (cd*)st->small.get((size_t)B * K * 16);
    int64_t W = std::max<int64_t>(1, std::min<int64_t>(B, budget_bytes(st) / (nnz * 16)));
    cd* dd = (cd*)st->data.get((size_t)W * nnz * 16);
    CK(cudaMemcpyAsync(dm, mats, (size_t)K * nnz * 8, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(dc, coef, (size_t)B * K * 16, cudaMemcpyHostToDevice, s));
    const int bchunk = 16;
    for (int64_t b0 = 0; b0 < B; b0 += W) {
        const int nw = (int)std::min<int64_t>(W, B - b0);
        dim3 grid((unsigned)((nnz + 255) / 256), (unsigned)((nw + bchunk - 1) / bchunk));
        assemble_kernel<KMAX><<<grid, 256, (size_t)bchunk * K * 16, s>>>(nnz, K, nw, dm, dc + b0 * K, dd, bchunk);
        launch_check(st);
        CK(cudaMemcpyAsync((cd*)data + b0 * nnz, dd, (size_t)nw * nnz * 16, cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_register_components(jpb_state* st, int solver_num, int K, const double* mats,
                            const double* rhs_vec, int R, const double* Q) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    if (!mats || !rhs_vec || !Q) throw InvalidArg("register_components: null array");
    if (K <= 0 || K > KMAX) throw InvalidArg("register_components: K must be in [1, 32]");
    if (R <= 0 || R > 4) throw InvalidArg("register_components: R must be in [1, 4]");
    CK(cudaSetDevice(st->device));
    const int64_t nnz = P.S.nnz;
    const int N = P.S.N;
    for (double** p : {&P.d_mats, &P.d_mats_f, &P.d_rhs_vec, &P.d_Q, &P.d_mats_diag})
        if (*p) { cudaFree(*p); *p = nullptr; }
    P.K = 0;
    CK(cudaMalloc(&P.d_mats, (size_t)K * nnz * 8));
    CK(cudaMalloc(&P.d_mats_f, (size_t)K * nnz * 8));
    CK(cudaMalloc(&P.d_rhs_vec, (size_t)N * 8));
    CK(cudaMalloc(&P.d_Q, (size_t)R * N * 8));
    CK(cudaMemcpy(P.d_mats, mats, (size_t)K * nnz * 8, cudaMemcpyHostToDevice));
    std::vector<double> mf((size_t)K * nnz);
    for (int k = 0; k < K; ++k)
        for (int64_t q = 0; q < nnz; ++q) mf[(size_t)k * nnz + q] = mats[(size_t)k * nnz + P.S.ent_src[q]];
    CK(cudaMemcpy(P.d_mats_f, mf.data(), (size_t)K * nnz * 8, cudaMemcpyHostToDevice));
    {
        std::vector<int> diag_csc(N, -1);
        CK(cudaMemcpy(diag_csc.data(), P.d_diag_csc, (size_t)N * sizeof(int), cudaMemcpyDeviceToHost));
        std::vector<double> md((size_t)K * N, 0.0);
        for (int k = 0; k < K; ++k)
            for (int i = 0; i < N; ++i)
                if (diag_csc[i] >= 0) md[(size_t)k * N + i] = mats[(size_t)k * nnz + diag_csc[i]];
        CK(cudaMalloc(&P.d_mats_diag, (size_t)K * N * 8));
        CK(cudaMemcpy(P.d_mats_diag, md.data(), (size_t)K * N * 8, cudaMemcpyHostToDevice));
    }
    CK(cudaMemcpy(P.d_rhs_vec, rhs_vec, (size_t)N * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(P.d_Q, Q, (size_t)R * N * 8, cudaMemcpyHostToDevice));
    P.K = K;
    P.R = R;
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_eval_forward(jpb_state* st, int solver_num, int64_t B, const double* coef,
                     const double* rhs_scale, double* u, double* y, int* status) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    if (!coef || !rhs_scale || !y || B <= 0) throw InvalidArg("eval_forward: bad arguments");
    if (P.K == 0) throw StateError("components not registered for this pattern (jpb_register_components)");
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = st->stream;
    const int K = P.K, R = P.R, N = P.S.N;
    size_t bytes = (size_t)B * (K * 16 + 16 + R * 16 + 4) + 256;
    char* stage = (char*)st->hstage.get(bytes);
    cd* dcoef = (cd*)stage;
    cd* dscale = dcoef + B * K;
    cd* dy = dscale + B;
    int* dstat = (int*)(dy + B * R);
    cd* du = u ? (cd*)st->data.get((size_t)B * N * 16) : nullptr;
    {
        Span span(st, s, PH_H2D);
        CK(cudaMemcpyAsync(dcoef, coef, (size_t)B * K * 16, cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(dscale, rhs_scale, (size_t)B * 16, cudaMemcpyHostToDevice, s));
    }
    eval_core(st, s, P, B, dcoef, dscale, nullptr, nullptr, 0, B, du, dy, nullptr, nullptr, nullptr,
              nullptr, dstat, false);
    std::vector<int> hstat(B);
    {
        Span span(st, s, PH_D2H);
        CK(cudaMemcpyAsync(y, dy, (size_t)B * R * 16, cudaMemcpyDeviceToHost, s));
        if (u) CK(cudaMemcpyAsync(u, du, (size_t)B * N * 16, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(hstat.data(), dstat, (size_t)B * 4, cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    int bad = 0;
    for (int64_t i = 0; i < B; ++i) {
        if (status) status[i] = hstat[i];
        bad |= hstat[i];
    }
    if (bad) return fail(st, JPB_ESINGULAR, "eval_forward: %s", status_message(bad));
    return JPB_OK;
    JPB_CATCH(st)
}

namespace {

void check_eval_args(Pattern& P, int64_t B, const void* coef, const void* rhs_scale, const void* ref,
                     const void* w, int loss_type, int64_t F_total, const void* fr, const void* loss_sum,
                     const void* ct_coef, const void* ct_s, const int* seg, int nseg) {
    if (!coef || !rhs_scale || !w || !fr || !ct_coef || !ct_s || B <= 0)
        throw InvalidArg("eval_loss_grad: bad arguments");
    if (loss_type < 0 || loss_type > 6) throw InvalidArg("eval_loss_grad: unknown loss type");
    if (loss_type >= 5 && P.R != 1)
        throw InvalidArg("eval_loss_grad: the complex losses (types 5, 6) need a single read-out row");
    if (loss_type != 4 && (!ref || !loss_sum || F_total < 1)) throw InvalidArg("eval_loss_grad: bad arguments");
    if (nseg < 1 || (nseg > 1 && !seg)) throw InvalidArg("eval_loss_grad: bad candidate segments");
    if (B > INT32_MAX) throw InvalidArg("eval_loss_grad: batch too large");
    if (P.K == 0) throw StateError("components not registered for this pattern (jpb_register_components)");
}

// host-buffer evaluation: inputs are packed into the page-locked mirror and go to the device in
// one asynchronous copy, results come back the same way
int eval_host(jpb_state* st, int solver_num, int64_t B, const double* coef, const double* rhs_scale,
              const double* ref, const double* w, int loss_type, int64_t F_total, const int32_t* seg,
              int nseg, double* fr, double* loss_sum, double* ct_coef, double* ct_s, int* status) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    check_eval_args(P, B, coef, rhs_scale, ref, w, loss_type, F_total, fr, loss_sum, ct_coef, ct_s, seg, nseg);
    if (seg)
        for (int64_t i = 0; i < B; ++i)
            if (seg[i] < 0 || seg[i] >= nseg) throw InvalidArg("eval_candidates: candidate id out of range");
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = st->stream;
    const int K = P.K, R = P.R;
    // device staging block, inputs first and results last so that each direction is one copy:
    //   in : coef [B,K] c, scale [B] c, ref [B] d (c for the complex losses), w [4] d, seg [B] i
    //   -- : y [B,R] c (device only)
    //   out: ct_coef [B,K] c, ct_s [B] c, fr [B] d, loss [nseg] d, status [B] i
    const size_t in_bytes = ((size_t)B * (K * 16 + 16 + 16 + 4) + 32 + 15) & ~(size_t)15;
    const int ref_doubles = loss_type >= 5 ? 2 : 1;
    const size_t y_bytes = (size_t)B * R * 16;
    const size_t out_bytes = (size_t)B * (K * 16 + 16 + 8 + 4) + (size_t)nseg * 8 + 16;
    char* stage = (char*)st->hstage.get(in_bytes + y_bytes + out_bytes + 64);
    char* pin = (char*)st->pin.get(in_bytes + out_bytes);
    cd* dcoef = (cd*)stage;
    cd* dscale = dcoef + B * K;
    double* dref = (double*)(dscale + B);
    double* dw = dref + 2 * B;
    int* dseg = (int*)(dw + 4);
    cd* dy = (cd*)(stage + in_bytes);
    char* dout = stage + in_bytes + y_bytes;
    cd* dct = (cd*)dout;
    cd* dcts = dct + B * K;
    double* dfr = (double*)(dcts + B);
    double* dloss = dfr + B;
    int* dstat = (int*)(dloss + nseg);
    {
        Span span(st, s, PH_H2D);
        memcpy(pin, coef, (size_t)B * K * 16);
        memcpy(pin + ((char*)dscale - stage), rhs_scale, (size_t)B * 16);
        if (ref) memcpy(pin + ((char*)dref - stage), ref, (size_t)B * 8 * ref_doubles);
        memcpy(pin + ((char*)dw - stage), w, (size_t)R * 8);
        if (seg) memcpy(pin + ((char*)dseg - stage), seg, (size_t)B * 4);
        CK(cudaMemcpyAsync(stage, pin, in_bytes, cudaMemcpyHostToDevice, s));
    }
    eval_core(st, s, P, B, dcoef, dscale, ref ? dref : nullptr, dw, loss_type, std::max<int64_t>(F_total, 1),
              nullptr, dy, dfr, dloss, dct, dcts, dstat, true, seg ? dseg : nullptr, nseg);
    char* pout = pin + in_bytes;
    {
        Span span(st, s, PH_D2H);
        CK(cudaMemcpyAsync(pout, dout, out_bytes, cudaMemcpyDeviceToHost, s));
    }
    CK(cudaStreamSynchronize(s));
    memcpy(ct_coef, pout, (size_t)B * K * 16);
    memcpy(ct_s, pout + ((char*)dcts - dout), (size_t)B * 16);
    memcpy(fr, pout + ((char*)dfr - dout), (size_t)B * 8);
    if (loss_sum) memcpy(loss_sum, pout + ((char*)dloss - dout), (size_t)nseg * 8);
    const int* hstat = (const int*)(pout + ((char*)dstat - dout));
    int bad = 0;
    for (int64_t i = 0; i < B; ++i) {
        if (status) status[i] = hstat[i];
        bad |= hstat[i];
    }
    if (bad) return fail(st, JPB_ESINGULAR, "%s", status_message(bad));
    return JPB_OK;
    JPB_CATCH(st)
}

}  // namespace

int jpb_eval_candidates_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                            const double* rhs_scale, const double* ref, const double* w, int loss_type,
                            int64_t F_total, const int32_t* cand, int ncand, double* fr, double* loss_sum,
                            double* ct_coef, double* ct_s, int* status, void* stream) {
    JPB_TRY(st)
    Pattern& P = pattern(st, solver_num);
    check_eval_args(P, B, coef, rhs_scale, ref, w, loss_type, F_total, fr, loss_sum, ct_coef, ct_s, cand, ncand);
    CK(cudaSetDevice(st->device));
    st->reset_timing();
    cudaStream_t s = (cudaStream_t)stream;
    cd* dy = (cd*)st->hstage.get((size_t)B * P.R * 16 + 64);
    eval_core(st, s, P, B, (const cd*)coef, (const cd*)rhs_scale, ref, w, loss_type, std::max<int64_t>(F_total, 1),
              nullptr, dy, fr, loss_sum, (cd*)ct_coef, (cd*)ct_s, status, true, cand, ncand);
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_eval_loss_grad_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                           const double* rhs_scale, const double* ref, const double* w, int loss_type,
                           int64_t F_total, double* fr, double* loss_sum, double* ct_coef, double* ct_s,
                           int* status, void* stream) {
    if (loss_type == 4) return fail(st, JPB_EINVAL, "eval_loss_grad: unknown loss type");
    return jpb_eval_candidates_dev(st, solver_num, B, coef, rhs_scale, ref, w, loss_type, F_total, nullptr, 1,
                                   fr, loss_sum, ct_coef, ct_s, status, stream);
}

int jpb_eval_loss_grad(jpb_state* st, int solver_num, int64_t B, const double* coef,
                       const double* rhs_scale, const double* ref, const double* w, int loss_type,
                       int64_t F_total, double* fr, double* loss_sum, double* ct_coef, double* ct_s,
                       int* status) {
    if (loss_type == 4) return fail(st, JPB_EINVAL, "eval_loss_grad: unknown loss type");
    if (F_total < B) return fail(st, JPB_EINVAL, "eval_loss_grad: bad arguments");
    return eval_host(st, solver_num, B, coef, rhs_scale, ref, w, loss_type, F_total, nullptr, 1, fr, loss_sum,
                     ct_coef, ct_s, status);
}

int jpb_eval_candidates(jpb_state* st, int solver_num, int64_t B, const double* coef,
                        const double* rhs_scale, const double* ref, const double* w, int loss_type,
                        int64_t F_total, const int32_t* cand, int ncand, double* fr, double* loss_sum,
                        double* ct_coef, double* ct_s, int* status) {
    if (loss_type == 4) return fail(st, JPB_EINVAL, "eval_candidates: unknown loss type");
    if (!cand) return fail(st, JPB_EINVAL, "eval_candidates: null candidate ids");
    return eval_host(st, solver_num, B, coef, rhs_scale, ref, w, loss_type, F_total, cand, ncand, fr, loss_sum,
                     ct_coef, ct_s, status);
}

int jpb_eval_jacobian(jpb_state* st, int solver_num, int64_t B, const double* coef,
                      const double* rhs_scale, const double* w, double* fr, double* ct_coef, double* ct_s,
                      int* status) {
    return eval_host(st, solver_num, B, coef, rhs_scale, nullptr, w, 4, 1, nullptr, 1, fr, nullptr, ct_coef,
                     ct_s, status);
}

int jpb_eval_jacobian_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                          const double* rhs_scale, const double* w, double* fr, double* ct_coef,
                          double* ct_s, int* status, void* stream) {
    return jpb_eval_candidates_dev(st, solver_num, B, coef, rhs_scale, nullptr, w, 4, 1, nullptr, 1, fr,
                                   nullptr, ct_coef, ct_s, status, stream);
}

int jpb_last_timings(const jpb_state* st, double* t10) {
    if (!st || !t10) return JPB_EINVAL;
    cudaSetDevice(st->device);
    summarise_timings(st, t10);
    return JPB_OK;
}

int jpb_dmma_peak(jpb_state* st, double* tflops) {
    JPB_TRY(st)
    if (!st || !tflops) throw InvalidArg("null argument");
    CK(cudaSetDevice(st->device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, st->device));
    double* sink = (double*)st->small.get(64);
    const int blocks = prop.multiProcessorCount * 4, threads = 256, iters = 20000;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    dmma_peak_kernel<<<blocks, threads, 0, st->stream>>>(1000, sink);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(a, st->stream));
        dmma_peak_kernel<<<blocks, threads, 0, st->stream>>>(iters, sink);
        CK(cudaEventRecord(b, st->stream));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        best = std::min(best, ms);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    double flops = (double)blocks * (threads / 32) * (double)iters * 16.0 * (2.0 * 8 * 8 * 4);
    *tflops = flops / (best * 1e-3) / 1e12;
    return JPB_OK;
    JPB_CATCH(st)
}

int jpb_selftest(jpb_state* st, double* max_err) {
    JPB_TRY(st)
    if (!st || !max_err) throw InvalidArg("null argument");
    CK(cudaSetDevice(st->device));
    const int lda = 37, ldb = 41;
    std::vector<cd> A(16 * lda), B(40 * ldb);
    unsigned seed = 12345u;
    auto rnd = [&]() { seed = seed * 1664525u + 1013904223u; return (double)(seed >> 8) / (1 << 24) - 0.5; };
    for (auto& v : A) v = make_double2(rnd(), rnd());
    for (auto& v : B) v = make_double2(rnd(), rnd());
    cd *dA = upload(A), *dB = upload(B), *dC, *dR;
    CK(cudaMalloc(&dC, 16 * 32 * 16));
    CK(cudaMalloc(&dR, 16 * 32 * 16));
    double worst = 0.0;
    const int cases[][3] = {{16, 32, 32}, {16, 32, 35}, {13, 27, 30}, {5, 3, 2}, {16, 32, 4}, {9, 17, 1}};
    for (auto& c : cases) {
        CK(cudaMemset(dC, 0, 16 * 32 * 16));
        CK(cudaMemset(dR, 0, 16 * 32 * 16));
        selftest_gemm_kernel<<<1, 32, 0, st->stream>>>(dA, lda, c[0], dB, ldb, c[1], c[2], dC, dR);
        CK(cudaStreamSynchronize(st->stream));
        std::vector<cd> hC(16 * 32), hR(16 * 32);
        CK(cudaMemcpy(hC.data(), dC, 16 * 32 * 16, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(hR.data(), dR, 16 * 32 * 16, cudaMemcpyDeviceToHost));
        double num = 0, den = 1e-300;
        for (int i = 0; i < 16 * 32; ++i) {
            num = std::max(num, std::max(std::fabs(hC[i].x - hR[i].x), std::fabs(hC[i].y - hR[i].y)));
            den = std::max(den, std::max(std::fabs(hR[i].x), std::fabs(hR[i].y)));
        }
        worst = std::max(worst, num / den);
    }
    cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFree(dR);
    {   // CTA-level GEMM fed by bulk asynchronous copies, against scalar loops
        const int cases2[][3] = {{64, 64, 16}, {64, 64, 64}, {100, 70, 37}, {130, 200, 5}, {33, 17, 1},
                                 {256, 192, 203}, {16, 300, 64}, {300, 32, 96}};
        for (auto& c : cases2) {
            const int m = c[0], n = c[1], K = c[2], la = K + 3, lb = n + 5, lc = n + 1;
            std::vector<cd> hA((size_t)m * la), hB((size_t)K * lb), hC((size_t)m * lc);
            for (auto& v : hA) v = make_double2(rnd(), rnd());
            for (auto& v : hB) v = make_double2(rnd(), rnd());
            for (auto& v : hC) v = make_double2(rnd(), rnd());
            cd *gA = upload(hA), *gB = upload(hB), *gC = upload(hC), *gR = upload(hC);
            selftest_cta_gemm_kernel<<<1, 256, GB_SMEM, st->stream>>>(gA, la, m, gB, lb, n, K, gC, gR, lc);
            CK(cudaStreamSynchronize(st->stream));
            std::vector<cd> oC(hC.size()), oR(hC.size());
            CK(cudaMemcpy(oC.data(), gC, oC.size() * 16, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(oR.data(), gR, oR.size() * 16, cudaMemcpyDeviceToHost));
            double num = 0, den = 1e-300;
            for (size_t i = 0; i < oC.size(); ++i) {
                num = std::max(num, std::max(std::fabs(oC[i].x - oR[i].x), std::fabs(oC[i].y - oR[i].y)));
                den = std::max(den, std::max(std::fabs(oR[i].x), std::fabs(oR[i].y)));
            }
            worst = std::max(worst, num / den);
            cudaFree(gA); cudaFree(gB); cudaFree(gC); cudaFree(gR);
        }
    }
    *max_err = worst;
    return JPB_OK;
    JPB_CATCH(st)
}

struct jpb_host_symbolic {
    Symbolic S;
};

int jpb_host_analyze(int64_t N, int64_t nnz, const void* indices, const void* indptr, int index_bytes,
                     jpb_host_symbolic** out, char* errbuf, int errlen) {
    try {
        if (!out || !indices || !indptr || N <= 0 || nnz <= 0 || (index_bytes != 4 && index_bytes != 8))
            throw std::runtime_error("host_analyze: bad arguments");
        auto get = [&](const void* p, int64_t i) -> int64_t {
            return index_bytes == 4 ? (int64_t)((const int32_t*)p)[i] : ((const int64_t*)p)[i];
        };
        std::vector<int64_t> ap(N + 1), ai(nnz);
        for (int64_t i = 0; i <= N; ++i) ap[i] = get(indptr, i);
        for (int64_t i = 0; i < nnz; ++i) ai[i] = get(indices, i);
        if (ap[0] != 0 || ap[N] != nnz) throw std::runtime_error("host_analyze: indptr does not match nnz");
        auto h = std::make_unique<jpb_host_symbolic>();
        analyze((int)N, ap.data(), ai.data(), symbolic_options(), h->S);
        *out = h.release();
        return JPB_OK;
    } catch (const std::exception& e) {
        if (errbuf && errlen > 0) snprintf(errbuf, errlen, "%s", e.what());
        return JPB_EINVAL;
    }
}

int jpb_host_sizes(const jpb_host_symbolic* h, int64_t* z, double* flops) {
    if (!h || !z) return JPB_EINVAL;
    const Symbolic& S = h->S;
    z[0] = S.nfronts; z[1] = (int64_t)S.rowidx.size(); z[2] = (int64_t)S.relmap.size(); z[3] = S.nnz;
    z[4] = S.nlevels; z[5] = S.front_slots; z[6] = S.nnz_lu; z[7] = S.N;
    if (flops) *flops = S.flops;
    return JPB_OK;
}

int jpb_host_export(const jpb_host_symbolic* h, int32_t* perm, int32_t* f_s, int32_t* f_k, int32_t* f_ld,
                    int64_t* f_off, int32_t* f_rowptr, int32_t* rowidx, int32_t* f_parent,
                    int32_t* f_level, int64_t* f_reloff, int32_t* relmap, int64_t* f_entptr,
                    int32_t* ent_src, int32_t* ent_slot) {
    if (!h) return JPB_EINVAL;
    const Symbolic& S = h->S;
    auto cp = [](auto* dst, const auto& v) { if (dst) std::copy(v.begin(), v.end(), dst); };
    cp(perm, S.perm); cp(f_s, S.f_s); cp(f_k, S.f_k); cp(f_ld, S.f_ld); cp(f_off, S.f_off);
    cp(f_rowptr, S.f_rowptr); cp(rowidx, S.rowidx); cp(f_parent, S.f_parent); cp(f_level, S.f_level);
    cp(f_reloff, S.f_reloff); cp(relmap, S.relmap); cp(f_entptr, S.f_entptr); cp(ent_src, S.ent_src);
    cp(ent_slot, S.ent_slot);
    return JPB_OK;
}

void jpb_host_free(jpb_host_symbolic* h) { delete h; }

}  // extern "C"

// ------------------------------------------------------------ allocation-free residual check
// Scratch of jpb_residual_noalloc: the three squared norms spmv_kernel<true> accumulates per
// system.  File-scope device memory (96 KB on every device that loads the library, held until
// exit) so that the call itself never allocates.  The name below is the HOST-side stand-in of
// the symbol; kernels get the device address obtained with cudaGetSymbolAddress.
namespace {
constexpr int NOALLOC_MAX_SYSTEMS = JPB_NOALLOC_MAX_SYSTEMS;
static __device__ __align__(256) double g_noalloc_norms[3 * NOALLOC_MAX_SYSTEMS];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// device address of the scratch array on `device` (which must be current); resolved once per device
cudaError_t noalloc_scratch(int device, double** out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        double* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_norms);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

// omega[v] = ||r_v|| / || |A||x_v| + |b_v| ||  from the squared norms (0 for an all-zero system)
__global__ void noalloc_omega_kernel(int nsys, const double* __restrict__ norms, double* __restrict__ omega) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nsys) return;
    const double r2 = norms[3 * v], t2 = norms[3 * v + 1];
    omega[v] = t2 > 0.0 ? sqrt(r2 / t2) : (r2 > 0.0 ? CUDART_INF : 0.0);
}
}  // namespace

extern "C" int jpb_residual_noalloc(jpb_state* st, int solver_num, int transpose, int64_t nsys,
                                    const double* mat, const double* x, const double* b, double* r,
                                    double* omega, void* stream) {
    if (!st) return JPB_EINVAL;
    if (solver_num < 0 || solver_num >= (int)st->pats.size())
        return fail(st, JPB_EINVAL, "residual_noalloc: solver_num out of range (no such pattern registered)");
    if (!mat || !x || !b || !r || !omega) return fail(st, JPB_EINVAL, "residual_noalloc: null array");
    if (nsys <= 0 || nsys > NOALLOC_MAX_SYSTEMS)
        return fail(st, JPB_EINVAL, "residual_noalloc: %lld systems in one call, the static scratch holds at most %d "
                    "(JPB_NOALLOC_MAX_SYSTEMS); split the batch", (long long)nsys, NOALLOC_MAX_SYSTEMS);
    const Pattern& P = *st->pats[solver_num];
    cudaError_t e = cudaSetDevice(st->device);
    double* norms = nullptr;
    if (e == cudaSuccess) e = noalloc_scratch(st->device, &norms);
    if (e != cudaSuccess) return fail(st, JPB_ECUDA, "residual_noalloc: %s", cudaGetErrorString(e));
    cudaStream_t s = stream ? (cudaStream_t)stream : st->stream;
    // the array keeps what the previous call left: the accumulators start at zero
    e = cudaMemsetAsync(norms, 0, (size_t)nsys * 24, s);
    if (e != cudaSuccess) return fail(st, JPB_ECUDA, "residual_noalloc: %s", cudaGetErrorString(e));
    const bool trans = transpose != 0;
    const int N = P.S.N;
    dim3 grid((N + 127) / 128, (unsigned)nsys);
    spmv_kernel<true><<<grid, 128, 0, s>>>(N, P.S.nnz, trans ? P.d_csc_ptr : P.d_csr_ptr,
                                           trans ? P.d_csc_row : P.d_csr_col, trans ? nullptr : P.d_csr_pos,
                                           (const cd*)mat, 1, (const cd*)x, 1, 0, (const cd*)b, 1, 0, (cd*)r,
                                           (int)nsys, norms, nullptr);
    noalloc_omega_kernel<<<(int)((nsys + 127) / 128), 128, 0, s>>>((int)nsys, norms, omega);
    st->launches += 2;
    e = cudaGetLastError();
    if (e == cudaSuccess && !stream) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return fail(st, JPB_ECUDA, "residual_noalloc: %s", cudaGetErrorString(e));
    return JPB_OK;
}
