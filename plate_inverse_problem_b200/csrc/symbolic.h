// Symbolic analysis for the batched multifrontal LU (host, once per sparsity pattern).
//
// Replaces what `umfpack_zi_symbolic` does for the reference at
// /root/reference/source/jax_plate_lib/include/InnerState.h:159-160 (pattern-only analysis
// cached per `solver_num`): fill-reducing ordering, elimination tree, supernodes, and --
// new here, because all systems of a frequency sweep share it -- the complete static
// schedule the GPU kernels run (front index lists, extend-add maps, level schedule, and
// the map from every CSC entry of A to its slot in front storage).
#pragma once
#include <cstdint>
#include <vector>

namespace jpb {

struct SymbolicOptions {
    int leaf_size = 96;        // nested dissection stops below this many vertices
    double balance = 0.30;     // each side of a separator keeps >= balance * n vertices
    int relax_small = 8;       // amalgamate unconditionally up to this many pivots
    int relax_mid = 32;        // ... up to here if zero fraction < relax_mid_frac
    double relax_mid_frac = 0.35;
    int relax_big = 96;
    double relax_big_frac = 0.12;
    double relax_any_frac = 0.04;
};

struct Symbolic {
    int N = 0;
    int64_t nnz = 0;
    // perm[new] = old ; pinv[old] = new   (symmetric permutation P A P^T)
    std::vector<int> perm, pinv;

    int nfronts = 0;
    std::vector<int> f_c0;       // first pivot column (permuted numbering)
    std::vector<int> f_k;        // number of pivots
    std::vector<int> f_s;        // front order (pivots + contribution rows)
    std::vector<int> f_ld;       // leading dimension of the row-major front (s rounded up to even)
    std::vector<int64_t> f_off;  // first slot of the s x ld row-major front (elements)
    std::vector<int> f_voff;     // first slot of the front-local vector (length s)
    std::vector<int> f_rowptr;   // [nfronts+1] into rowidx
    std::vector<int> rowidx;     // permuted global indices, first k of a front = its pivots
    std::vector<int> rowidx_orig; // perm[rowidx]: the same rows in the caller's numbering
    std::vector<int> f_parent;   // parent front or -1
    std::vector<int> f_level;    // 0 = leaves
    std::vector<int> f_childptr; // [nfronts+1] into children
    std::vector<int> children;   // child front ids, ascending
    // for child c: relmap[f_reloff[c] + (i - k_c)] = parent-local row of child row i >= k_c
    std::vector<int64_t> f_reloff;
    std::vector<int> relmap;

    int nlevels = 0;
    std::vector<int> levelptr;    // [nlevels+1] into levelfronts
    std::vector<int> levelfronts; // fronts of a level, largest first

    // entries of A grouped by the front that owns them: for q in [f_entptr[f], f_entptr[f+1])
    // entry ent_src[q] (position in the canonical CSC order handed to add_mat) lands on
    // front-local slot ent_slot[q] = i_local * ld + j_local.
    std::vector<int64_t> f_entptr;
    std::vector<int> ent_src;
    std::vector<int> ent_slot;

    int64_t front_slots = 0;   // sum s*ld
    int64_t vec_slots = 0;     // sum s
    int64_t nnz_lu = 0;        // entries of L+U (diagonal counted once)
    double flops = 0;          // real flops of one numeric factorisation (complex fma = 8)
    int max_s = 0, max_k = 0;
};

// indptr/indices: CSC of the N x N pattern (any integer width converted by the caller).
// Throws std::runtime_error on malformed input.
void analyze(int N, const int64_t* indptr, const int64_t* indices, const SymbolicOptions& opt,
             Symbolic& out);

// exposed for tests
void nested_dissection_order(int N, const std::vector<int64_t>& xadj,
                             const std::vector<int>& adj, const SymbolicOptions& opt,
                             std::vector<int>& perm);

}  // namespace jpb
