/* C-ABI of libjaxplate_b200.so -- the drop-in boundary for jax_plate's frequency-domain
 * forward / adjoint / gradient hot path on B200 (sm_100a).
 *
 * What each entry point replaces in the reference (paths relative to /root/reference):
 *
 *   jpb_add_mat        InnerState::add_mat<T,I>           source/jax_plate_lib/include/InnerState.h:120-162
 *                      (+ umfpack::symbolic               source/jax_plate_lib/include/umfpack_interface.h:42-52)
 *   jpb_solve          InnerState::solve<T,I>             InnerState.h:164-308   (mode table :187-232)
 *   jpb_matvec         InnerState::matvec<T,I>            InnerState.h:310-470, csc_matvec.h:31-66
 *   jpb_assemble       XLA-CPU assembly in Problem._solve source/jax_plate/Problem.py:440-445
 *   jpb_register_components / jpb_eval_forward / jpb_eval_loss_grad
 *                      the whole per-frequency body of Problem._solve + its reverse-mode
 *                      derivative: Problem.py:437-477 (assembly, rhs, solve, read-out),
 *                      Problem.py:948-975 (losses), Sparse.py:162-176,200-222 (AD rules)
 *
 * Conventions
 *   - complex128 arrays are passed as packed (re, im) pairs of doubles, like the reference
 *     passes them to UMFPACK (umfpack_interface.h:50-51: Az = NULL).
 *   - all array arguments are HOST pointers unless the function name ends in `_dev`;
 *     they are borrowed for the duration of the call; results are written into
 *     caller-allocated buffers.
 *   - every function returns 0 on success and a non-zero JPB_E* code otherwise;
 *     jpb_last_error() returns the message (the reference throws std::runtime_error,
 *     umfpack_interface.h:11-18; the Python layer turns codes into RuntimeError).
 *   - a state is bound to one CUDA device; calls on one state must not overlap.
 *   - there is no CPU fallback: without a usable CUDA device jpb_create fails.
 */
#ifndef JAX_PLATE_B200_H
#define JAX_PLATE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct jpb_state jpb_state;

enum {
    JPB_OK = 0,
    JPB_EINVAL = 1,    /* bad argument / shape / mode (reference: runtime_error "Invalid `mode`") */
    JPB_ECUDA = 2,     /* CUDA runtime failure */
    JPB_ESINGULAR = 3, /* a system was numerically singular, or its residual failed the
                          acceptance test (reference: UMFPACK status != 0) */
    JPB_ENOMEM = 4,
    JPB_ESTATE = 5     /* call sequence error (e.g. components not registered) */
};

/* loss types of Problem.getLossFunction (Problem.py:948-975) */
enum { JPB_LOSS_MSE = 0, JPB_LOSS_RMSE = 1, JPB_LOSS_MSE_AFC = 2, JPB_LOSS_MSE_LOG_AFC = 3,
       /* complex read-out of the mid-plane symmetric path (Problem.py:415: fr = c0 + c.u, one read-out row):
        * |fr - ref|^2 and |(fr - ref) / ref|^2 (Problem.py:948-960) with a COMPLEX reference, ref = [B]
        * interleaved (re, im) doubles; the `fr` output stays the amplitude |fr| */
       JPB_LOSS_MSE_COMPLEX = 5, JPB_LOSS_RMSE_COMPLEX = 6 };

/* ---- life cycle ------------------------------------------------------------------- */
int jpb_create(jpb_state** out, int device);
void jpb_destroy(jpb_state* st);
const char* jpb_last_error(const jpb_state* st);
/* upper bound of device memory (bytes) the state may use for fronts/workspaces;
 * 0 = 80 % of what is free at first use */
int jpb_set_memory_budget(jpb_state* st, int64_t bytes);
/* Iterative refinement of every triangular solve (UMFPACK's Control[UMFPACK_IRSTEP] analogue;
 * umfpack_zi_solve refines while the backward error improves, at most 2 steps by default):
 *   jpb_set_refinement(steps)          unconditional steps.  Default 1: the parity-grade setting --
 *                                      with it the forward error is at or below that of a CPU sparse
 *                                      LU on the same matrix.
 *   jpb_set_refinement_threshold(w, t) only used with steps = 0 ("fast mode", +14 % throughput at
 *                                      N = 2e4): a residual pass over A, then ONE step for exactly the
 *                                      systems whose normwise backward error
 *                                      omega = ||b - A x|| / || |A||x| + |b| || is above w (default
 *                                      1e-15) or whose residual relative to the right-hand side is
 *                                      above t (default 0 = no such test); w = 0 switches the residual
 *                                      pass off.  The equilibrated plate systems come out at
 *                                      omega ~ 1e-16 unrefined, like a CPU sparse LU, and skip the
 *                                      step; their forward error is then within ~10x of the stand-in
 *                                      LU's (both are cond(A) * eps). */
int jpb_set_refinement(jpb_state* st, int steps);
int jpb_set_refinement_threshold(jpb_state* st, double omega, double rel_residual);
/* how many of the vectors solved by the last call took the residual-driven step */
int jpb_last_refined(jpb_state* st, int64_t* refined, int64_t* solved);
/* Equilibration (default on; UMFPACK scales rows by default too, Control[UMFPACK_SCALE]): every
 * system is factorised as S A S with s_i = 2^-floor(log2|a_ii| / 2) -- powers of two, so the
 * scaled entries are exact -- and right-hand sides / solutions are converted on the way in and
 * out.  Results are those of the unscaled system up to rounding; what changes is the backward
 * error of the block factorisation with tile-local pivoting, which for the plate matrices (1e7
 * between the in-plane and bending blocks) drops to that of a CPU sparse LU in both the plain
 * and the transposed solve. */
int jpb_set_equilibration(jpb_state* st, int on);
/* correction-driven refinement on top of the above: up to `extra_steps` further steps
 * (default 0 = off), each executed only for the systems whose last correction was still larger
 * than tol * ||x|| (max-norms, default 1e-11); the other systems' blocks exit at once.  Meant
 * for matrices on which the static-pivot factorisation has a poor backward error, so that one
 * step does not reach the working-precision floor.  For the plate sweeps one step does
 * (measured: further steps move u by less than cond(A) * eps), hence off.
 * umfpack_zi_solve refines until the backward error stops improving, at most 2 steps. */
int jpb_set_adaptive_refinement(jpb_state* st, int extra_steps, double tol);
/* number of concurrent sub-waves (CUDA streams) a fused evaluation is cut into (default 1;
 * 1 = strictly sequential phases, which is what the per-phase timers of jpb_last_timings
 * need to be meaningful).  No counterpart in the reference (OpenMP over items,
 * InnerState.h:277, is its analogue). */
int jpb_set_streams(jpb_state* st, int n);
/* Residual acceptance.  Pivoting is static with partial pivoting inside the 32-wide pivot
 * tiles only, which is safe for the plate matrices of jax_plate but
 * not for arbitrary ones, so every system is also judged by its normwise backward error
 * omega = ||b - A x||_2 / || |A| |x| + |b| ||_2 (moduli in the 1-norm of (re, im)):
 *   jpb_solve      : omega <= tol_final for the solution returned (one more pass over A per
 *                    call; default 1e-10; a healthy solve gives ~1e-16);
 *   jpb_eval_*     : omega of the UNREFINED solution <= tol_unrefined (default 1e-6; the
 *                    refinement step that follows contracts it by ~cond(A) * omega), tested
 *                    on the residual the refinement computes anyway -- no extra pass.
 * A system that fails gets status bit 2 and the call returns JPB_ESINGULAR.  (UMFPACK's
 * threshold pivoting has no such restriction; the reference raises only for status != 0,
 * umfpack_interface.h:11-18.) */
int jpb_set_residual_tolerance(jpb_state* st, double tol_final, double tol_unrefined);

/* ---- pattern registration (InnerState::add_mat) ------------------------------------
 * indices/indptr: CSC of the N x N pattern; indices_T/indptr_T/perm: the reference's CSR
 * view and CSR->CSC permutation (Sparse.py:33-34).  Only the CSC arrays are used for the
 * analysis; the CSR view is rebuilt exactly (so transposed products are right for
 * structurally unsymmetric patterns too) and `perm`, when given, is checked against it.
 * index_bytes = 4 or 8.  *solver_num receives the 0-based handle (insertion index). */
int jpb_add_mat(jpb_state* st, int64_t N, int64_t nnz, const void* indices, const void* indptr,
                const void* indices_T, const void* indptr_T, const void* perm, int index_bytes,
                int* solver_num);

/* symbolic statistics: info[0]=N, [1]=nnz, [2]=nnz(L+U), [3]=real flops of one numeric
 * factorisation, [4]=front slots (complex elements per system), [5]=#fronts, [6]=#levels,
 * [7]=largest front order, [8]=largest pivot block, [9]=vector workspace slots */
int jpb_symbolic_info(const jpb_state* st, int solver_num, double* info10);
/* fill-reducing permutation found by the analysis: perm[new] = old, length N */
int jpb_get_ordering(const jpb_state* st, int solver_num, int32_t* perm);

/* ---- InnerState::solve / matvec ----------------------------------------------------
 * mode table (Sparse.py:238-277, InnerState.h:187-232):
 *   0: data[nnz]      b[N]         -> out[N]
 *   1: data[B,nnz]    b[N]         -> out[B,N]
 *   2: data[nnz]      b[B,N]       -> out[B,N]      (one factorisation, B right-hand sides)
 *   3: data[B,nnz]    b[B,N]       -> out[B,N]
 *   4: data[B1,nnz]   b[B0,B1,N]   -> out[B0,B1,N]  (B1 factorisations, B0 rhs each)
 * nbatch0 = B (modes 1-3) or B0 (mode 4); nbatch1 = B1 (mode 4 only).
 * transpose != 0 solves A^T x = b (plain transpose, UMFPACK_Aat, InnerState.h:183-185).
 * data / b / out (mat / vec / out) may be HOST or DEVICE pointers (unified addressing decides per
 * copy), which is what lets the XLA-FFI handlers of csrc/xla_ffi.cc pass XLA's device buffers
 * straight through; `status` is always host memory.
 * status (optional, one int per system): bit 0 (1) zero / non-finite pivot block, bit 1 (2)
 * backward error above the tolerance of jpb_set_residual_tolerance. */
int jpb_solve(jpb_state* st, int solver_num, int mode, int transpose, const double* data,
              const double* b, int64_t nbatch0, int64_t nbatch1, double* out, int* status);
int jpb_matvec(jpb_state* st, int solver_num, int mode, int transpose, const double* mat,
               const double* vec, int64_t nbatch0, int64_t nbatch1, double* out);

/* data[b,e] = sum_k coef[b,k] * mats[k,e]   (Problem.py:440-445), coef complex, mats real */
int jpb_assemble(jpb_state* st, int64_t nnz, int K, int64_t B, const double* mats,
                 const double* coef, double* data);

/* ---- fused per-frequency evaluation -------------------------------------------------
 * mats [K,nnz] real in the canonical CSC order, rhs_vec [N] real, Q [R,N] real read-out
 * rows (row means of the interpolation matrices, Problem.py:454-462).  Kept on the device. */
int jpb_register_components(jpb_state* st, int solver_num, int K, const double* mats,
                            const double* rhs_vec, int R, const double* Q);

/* For b < B: A_b = sum_k coef[b,k] M_k ; u_b = A_b^-1 (rhs_scale[b] * rhs_vec) ;
 * y[b,r] = Q_r . u_b.  u (optional) [B,N] complex, y [B,R] complex. */
int jpb_eval_forward(jpb_state* st, int solver_num, int64_t B, const double* coef,
                     const double* rhs_scale, double* u, double* y, int* status);

/* Forward + adjoint + gradient for the unsymm read-out  fr = sqrt(sum_r w[r] |y_r|^2)
 * (Problem.py:464-477) and loss `loss_type` against ref[B] with mean over F_total
 * frequencies (F_total >= B allows sharding one sweep over several calls / GPUs):
 *   fr[B] real; loss_sum = sum_b loss term (caller divides by F_total);
 *   ct_coef[B,K] complex = -lambda_b^T M_k u_b ; ct_s[B] complex = lambda_b^T rhs_vec
 * with lambda_b = A_b^-T ubar_b, ubar_b the JAX-convention cotangent of u_b
 * (SURVEY App. D).  One numeric factorisation per system serves both solves.
 * All arrays are HOST memory (any alignment, pageable is fine): inputs are packed into a
 * page-locked mirror owned by the state and go to the device in one asynchronous copy,
 * results come back the same way; the call returns after the stream has drained. */
int jpb_eval_loss_grad(jpb_state* st, int solver_num, int64_t B, const double* coef,
                       const double* rhs_scale, const double* ref, const double* w,
                       int loss_type, int64_t F_total, double* fr, double* loss_sum,
                       double* ct_coef, double* ct_s, int* status);

/* same as jpb_eval_loss_grad with every array already in device memory (zero-copy for
 * torch tensors); runs on `stream` (a cudaStream_t) without synchronising the host. */
int jpb_eval_loss_grad_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                           const double* rhs_scale, const double* ref, const double* w,
                           int loss_type, int64_t F_total, double* fr, double* loss_sum,
                           double* ct_coef, double* ct_s, int* status, void* stream);

/* Batched parameter candidates (BASELINE configs[4]; jax.vmap over candidates on top of the
 * vmap over omega, Problem.py:516): the B systems are the flattened (candidate, frequency)
 * pairs this call owns, cand[b] in [0, ncand) names the candidate of system b, and
 * loss_sum[c] = sum of the loss terms of candidate c (caller divides by F_total, the sweep
 * length per candidate).  Everything else as jpb_eval_loss_grad; ct_coef / ct_s stay per
 * system, so the caller's chain rule yields one gradient per candidate. */
int jpb_eval_candidates(jpb_state* st, int solver_num, int64_t B, const double* coef,
                        const double* rhs_scale, const double* ref, const double* w,
                        int loss_type, int64_t F_total, const int32_t* cand, int ncand,
                        double* fr, double* loss_sum, double* ct_coef, double* ct_s, int* status);
int jpb_eval_candidates_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                            const double* rhs_scale, const double* ref, const double* w,
                            int loss_type, int64_t F_total, const int32_t* cand, int ncand,
                            double* fr, double* loss_sum, double* ct_coef, double* ct_s,
                            int* status, void* stream);

/* Gauss-Newton / Levenberg-Marquardt Jacobian rows (BASELINE configs[2]; the model the
 * reference sketches in Optimizers.py:139-144 and would obtain through mode-4 multi-RHS
 * sensitivity solves, Sparse.py:267-270 / InnerState.h:289-304).  Same pipeline as
 * jpb_eval_loss_grad with the cotangent of fr itself, ybar = w * conj(y) / fr:
 *   ct_coef[b,k] = d fr_b / d coef[b,k],  ct_s[b] = d fr_b / d rhs_scale[b]   (JAX convention)
 * so that  J[b,p] = d fr_b / d p = Re(sum_k ct_coef[b,k] dcoef[b,k]/dp + ct_s[b] ds[b]/dp).
 * One factorisation and ONE adjoint solve per frequency give the whole row, where forward
 * sensitivities need P right-hand sides. */
int jpb_eval_jacobian(jpb_state* st, int solver_num, int64_t B, const double* coef,
                      const double* rhs_scale, const double* w, double* fr, double* ct_coef,
                      double* ct_s, int* status);
int jpb_eval_jacobian_dev(jpb_state* st, int solver_num, int64_t B, const double* coef,
                          const double* rhs_scale, const double* w, double* fr, double* ct_coef,
                          double* ct_s, int* status, void* stream);

/* ---- instrumentation ----------------------------------------------------------------
 * phase timers (ms, CUDA events on the state's stream) of the last jpb_eval_* / jpb_solve
 * call: [0] factor, [1] forward solve(s), [2] adjoint solve(s), [3] residual/refinement
 * SpMV, [4] gradient, [5] read-out/loss, [6] H2D, [7] D2H; [8] kernel launches issued;
 * [9] systems per wave */
int jpb_last_timings(const jpb_state* st, double* t10);
/* FP64 tensor-core (DMMA m8n8k4) peak of this device in TFLOP/s, register-resident loop */
int jpb_dmma_peak(jpb_state* st, double* tflops);
/* self test of the dense-front kernels against scalar device loops; returns max rel err */
int jpb_selftest(jpb_state* st, double* max_err);

/* ---- allocation-free residual check ---------------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * For nsys systems on the pattern `solver_num` (registered earlier with jpb_add_mat, which is
 * where all allocation happens): r[b] = rhs[b] - op(A_b) x[b] and the normwise backward error
 * omega[b] = ||r[b]||_2 / || |A_b||x[b]| + |rhs[b]| ||_2 (moduli as |re| + |im|), the acceptance
 * test of the solves.  It is the residual kernel of the refinement step (spmv_kernel<NORMS>)
 * with its per-system norm accumulators in a file-scope `static __device__` array of
 * 3 * JPB_NOALLOC_MAX_SYSTEMS doubles (96 KB on every GPU that loads the library, until exit)
 * instead of a grow-only workspace: the call makes no cudaMalloc / cudaFree / allocating
 * library call.  All five arrays are DEVICE pointers on the state's device, complex128 as
 * interleaved doubles: mat [nsys, nnz] in the caller's CSC value order, x / rhs / r [nsys, N],
 * omega [nsys] (real).  transpose != 0: op(A) = A^T (plain transpose, as jpb_solve).
 * stream: a cudaStream_t to launch on (the call then returns without synchronising) or NULL =
 * the state's stream, synchronised before return.  nsys > JPB_NOALLOC_MAX_SYSTEMS is rejected
 * with JPB_EINVAL (split the batch).  Every call on a device shares the one array, which
 * keeps the previous call's contents (the call clears what it uses first): calls that overlap
 * on the GPU (different streams, or graphs replayed concurrently) race on it -- serialise them. */
#define JPB_NOALLOC_MAX_SYSTEMS 4096
int jpb_residual_noalloc(jpb_state* st, int solver_num, int transpose, int64_t nsys,
                         const double* mat, const double* x, const double* rhs, double* r,
                         double* omega, void* stream);

/* ---- host-only view of the symbolic analysis (no CUDA device needed) ------------------
 * Used by the CPU test-suite to replay the multifrontal schedule in numpy and by tools
 * that want the ordering / fill statistics.  sizes: [0]=nfronts, [1]=len(rowidx),
 * [2]=len(relmap), [3]=nnz, [4]=nlevels, [5]=front_slots, [6]=nnz_lu, [7]=N. */
typedef struct jpb_host_symbolic jpb_host_symbolic;
int jpb_host_analyze(int64_t N, int64_t nnz, const void* indices, const void* indptr,
                     int index_bytes, jpb_host_symbolic** out, char* errbuf, int errlen);
int jpb_host_sizes(const jpb_host_symbolic* h, int64_t* sizes8, double* flops);
int jpb_host_export(const jpb_host_symbolic* h, int32_t* perm, int32_t* f_s, int32_t* f_k,
                    int32_t* f_ld, int64_t* f_off, int32_t* f_rowptr, int32_t* rowidx,
                    int32_t* f_parent, int32_t* f_level, int64_t* f_reloff, int32_t* relmap,
                    int64_t* f_entptr, int32_t* ent_src, int32_t* ent_slot);
void jpb_host_free(jpb_host_symbolic* h);

#ifdef __cplusplus
}
#endif
#endif
