"""B200-native implementation of jax_plate's frequency-domain forward / adjoint / gradient
hot path.  The compute lives in libjaxplate_b200.so (hand-written sm_100a
CUDA behind the C-ABI of include/jax_plate_b200.h); this package is the thin host layer
that mirrors the reference's `jax_plate_lib` / `jax_plate.Sparse` / `Problem` interface."""

__all__ = ["jax_plate_lib", "Sparse", "Problem", "Material", "synthetic"]
