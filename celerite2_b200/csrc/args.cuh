// Kernel argument blocks and the per-width launcher declarations.
// The launchers are defined in inst_sweep.cu / inst_factor.cu, which are compiled once per width JT
// (-DC2B_JT=...), so that the widths build in parallel.
#pragma once
#include "common.cuh"

namespace c2b {

// Scratch layout of a scanned op: nodes (leaves first) + states.
struct ScanBuffers {
  double *nodes  = nullptr;
  double *states = nullptr;
  int64_t node_stride = 0;
};

struct SweepArgs {
  int64_t N, batch, C, L;
  int J, K;
  const double *t, *c, *A, *B, *Y;  // A = "output" vector (dotted with the state), B = "input" vector
  double *Z, *F;
  int64_t t_bs, c_bs, y_bs;  // batch strides (0 when shared)
  // scan plumbing
  double *leaves;        // [batch][C][esize]   (summary kernels write here)
  const double *carry;   // [batch][total][ssize]: level-0 part holds carry per chunk (apply kernels read)
  int64_t node_stride;   // nodes per series (Levels::total)
};

struct SweepRevArgs {
  int64_t N, batch, C, L;
  int J, K;
  const double *t, *c, *A, *B, *Y, *Z, *F, *bZ;
  double *bt, *bA, *bB, *bY;
  int64_t t_bs, c_bs, y_bs;
  double *leaves;
  const double *carry;
  int64_t node_stride;
  double *bc_part;   // [batch][C][J]  per-chunk partial sums of bc
  double *bt_edge;   // [batch][C]     the one bt term each chunk owes to the row just outside it
};

struct FactorArgs {
  int64_t N, batch, C, L;
  int J;
  const double *t, *c, *a, *U, *V;
  double *d, *W, *S;
  int64_t *status;
  int64_t t_bs, c_bs;
  double *leaves;
  const double *carry;
  int64_t node_stride;
};

struct FactorRevArgs {
  int64_t N, batch, C, L;
  int J;
  const double *t, *c, *U, *d, *W, *S, *bd, *bW;
  double *bt, *ba, *bU, *bV;
  int64_t t_bs, c_bs;
  double *leaves;
  const double *carry;
  int64_t node_stride;
  double *bc_part, *bt_edge;
  int acc;  // C2B_ACCUMULATE: bt, bc, bU are added to
};

template <int JT, bool SOLVE, bool LOWER>
int launch_sweep(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream);
template <int JT, bool SOLVE, bool LOWER>
int launch_sweep_rev(SweepRevArgs &a, const ScanBuffers &sb, double *bc, cudaStream_t stream);
template <int JT>
int launch_factor(FactorArgs &g, const ScanBuffers &sb, cudaStream_t stream);
template <int JT>
int launch_factor_rev(FactorRevArgs &g, const ScanBuffers &sb, double *bc, cudaStream_t stream);

}  // namespace c2b
