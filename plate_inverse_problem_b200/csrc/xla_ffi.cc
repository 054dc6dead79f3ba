// XLA FFI custom-call handlers over the C-ABI of include/jax_plate_b200.h: the binding
// north_star describes ("host code stays Python/JAX and calls CUDA through a thin C-ABI
// registered as an XLA FFI custom call with a custom_vjp").  They replace the Python host
// callbacks of the reference's two primitives,
//     cpu_spsolve    /root/reference/source/jax_plate/Sparse.py:187-197, 224-229
//     custom_matvec  /root/reference/source/jax_plate/Sparse.py:150-160, 178-183
// and add the fused per-sweep calls behind Problem._solve / getLossFunction
// (Problem.py:432-477, 948-975).  plate_inverse_problem_b200/jax_binding.py registers them.
//
// NOT BUILT IN THIS IMAGE: jax / jaxlib -- and with them xla/ffi/api/ffi.h -- are absent
// (SURVEY 0.1).  __graft_entry__.build() compiles this file into libjaxplate_b200_ffi.so only
// when `import jax` succeeds:
//     nvcc -shared -Xcompiler -fPIC -std=c++17 -I$(python -c "import jax; print(jax.ffi.include_dir())")
//          -Iinclude csrc/xla_ffi.cc -L. -ljaxplate_b200 -o libjaxplate_b200_ffi.so
// The state handle travels as an int64 attribute (the address jpb_create returned); XLA's
// buffers are device memory on the CUDA platform, which jpb_solve / jpb_matvec accept as is.
#include <cstdint>

#include <cuda_runtime.h>

#include "xla/ffi/api/ffi.h"

#include "../../include/jax_plate_b200.h"

namespace ffi = xla::ffi;

namespace {

using C128 = ffi::Buffer<ffi::C128>;
using F64 = ffi::Buffer<ffi::F64>;
using RC128 = ffi::ResultBuffer<ffi::C128>;
using RF64 = ffi::ResultBuffer<ffi::F64>;

jpb_state* handle(int64_t h) { return reinterpret_cast<jpb_state*>(static_cast<intptr_t>(h)); }

ffi::Error done(jpb_state* st, int rc) {
    if (rc == JPB_OK) return ffi::Error::Success();
    return ffi::Error(rc == JPB_EINVAL ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal,
                      jpb_last_error(st));
}

// batch sizes of the `_mode` table (Sparse.py:238-277) from the operand shapes
void batch_sizes(int64_t mode, const C128& data, const C128& b, int64_t* nb0, int64_t* nb1) {
    auto dd = data.dimensions();
    auto bd = b.dimensions();
    *nb0 = *nb1 = 0;
    if (mode == 1) *nb0 = dd[0];
    else if (mode == 2 || mode == 3) *nb0 = bd[0];
    else if (mode == 4) { *nb0 = bd[0]; *nb1 = bd[1]; }
}

// the library runs on its own stream and returns after it has drained; order it after the
// work XLA has queued on the call's stream
ffi::Error SolveImpl(cudaStream_t stream, int64_t state, int64_t solver_num, bool transpose, int64_t mode,
                     C128 data, C128 b, RC128 out) {
    jpb_state* st = handle(state);
    if (cudaStreamSynchronize(stream) != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, "stream sync failed");
    int64_t nb0, nb1;
    batch_sizes(mode, data, b, &nb0, &nb1);
    return done(st, jpb_solve(st, (int)solver_num, (int)mode, transpose ? 1 : 0,
                              reinterpret_cast<const double*>(data.typed_data()),
                              reinterpret_cast<const double*>(b.typed_data()), nb0, nb1,
                              reinterpret_cast<double*>(out->typed_data()), nullptr));
}

ffi::Error MatvecImpl(cudaStream_t stream, int64_t state, int64_t solver_num, bool transpose, int64_t mode,
                      C128 mat, C128 vec, RC128 out) {
    jpb_state* st = handle(state);
    if (cudaStreamSynchronize(stream) != cudaSuccess) return ffi::Error(ffi::ErrorCode::kInternal, "stream sync failed");
    int64_t nb0, nb1;
    batch_sizes(mode, mat, vec, &nb0, &nb1);
    return done(st, jpb_matvec(st, (int)solver_num, (int)mode, transpose ? 1 : 0,
                               reinterpret_cast<const double*>(mat.typed_data()),
                               reinterpret_cast<const double*>(vec.typed_data()), nb0, nb1,
                               reinterpret_cast<double*>(out->typed_data())));
}

// fused sweep: value and the cotangents of (coef, rhs_scale) from ONE factorisation per system;
// runs on XLA's stream without a host synchronisation
ffi::Error EvalLossGradImpl(cudaStream_t stream, int64_t state, int64_t solver_num, int64_t loss_type,
                            int64_t f_total, C128 coef, C128 scale, F64 ref, F64 w, RF64 fr, RF64 loss_sum,
                            RC128 ct_coef, RC128 ct_s) {
    jpb_state* st = handle(state);
    const int64_t B = coef.dimensions()[0];
    return done(st, jpb_eval_loss_grad_dev(
                        st, (int)solver_num, B, reinterpret_cast<const double*>(coef.typed_data()),
                        reinterpret_cast<const double*>(scale.typed_data()), ref.typed_data(), w.typed_data(),
                        (int)loss_type, f_total, fr->typed_data(), loss_sum->typed_data(),
                        reinterpret_cast<double*>(ct_coef->typed_data()),
                        reinterpret_cast<double*>(ct_s->typed_data()), nullptr, stream));
}

// Gauss-Newton rows: fr and d fr / d (coef, rhs_scale)
ffi::Error EvalJacobianImpl(cudaStream_t stream, int64_t state, int64_t solver_num, C128 coef, C128 scale,
                            F64 w, RF64 fr, RC128 ct_coef, RC128 ct_s) {
    jpb_state* st = handle(state);
    const int64_t B = coef.dimensions()[0];
    return done(st, jpb_eval_jacobian_dev(
                        st, (int)solver_num, B, reinterpret_cast<const double*>(coef.typed_data()),
                        reinterpret_cast<const double*>(scale.typed_data()), w.typed_data(), fr->typed_data(),
                        reinterpret_cast<double*>(ct_coef->typed_data()),
                        reinterpret_cast<double*>(ct_s->typed_data()), nullptr, stream));
}

}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(
    JpbSolve, SolveImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Attr<int64_t>("state").Attr<int64_t>("solver_num").Attr<bool>("transpose").Attr<int64_t>("mode")
        .Arg<C128>().Arg<C128>().Ret<C128>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(
    JpbMatvec, MatvecImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Attr<int64_t>("state").Attr<int64_t>("solver_num").Attr<bool>("transpose").Attr<int64_t>("mode")
        .Arg<C128>().Arg<C128>().Ret<C128>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(
    JpbEvalLossGrad, EvalLossGradImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Attr<int64_t>("state").Attr<int64_t>("solver_num").Attr<int64_t>("loss_type").Attr<int64_t>("f_total")
        .Arg<C128>().Arg<C128>().Arg<F64>().Arg<F64>()
        .Ret<F64>().Ret<F64>().Ret<C128>().Ret<C128>());

XLA_FFI_DEFINE_HANDLER_SYMBOL(
    JpbEvalJacobian, EvalJacobianImpl,
    ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>()
        .Attr<int64_t>("state").Attr<int64_t>("solver_num")
        .Arg<C128>().Arg<C128>().Arg<F64>()
        .Ret<F64>().Ret<C128>().Ret<C128>());
