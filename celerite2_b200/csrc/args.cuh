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

// One op of a series that is split into contiguous segments (one per GPU): see c2b_segment_t in the public header.
//   phase 0: the whole op (default).  phase 1: summary pass + up-sweep, the segment's summary element goes to
//   summary_out (the tree stays in the caller's scratch).  phase 2: state_in seeds the top of the tree, down-sweep,
//   apply pass; reverse ops hand the bt term owed to the row just outside the segment to edge_out.
struct Seg {
  int phase;
  const double *state_in;
  double *summary_out;
  double *edge_out;
};

struct SweepArgs {
  int64_t N, batch, C, L;
  Seg seg;
  int64_t tmin, tmax;  // valid time-stamp indices: [0, N-1], or [-1, ...] / [..., N] for a segment with a neighbour
  int J, K;
  const double *t, *c, *A, *B, *Y;  // A = "output" vector (dotted with the state), B = "input" vector
  double *Z, *F;
  const double *zdiv;    // solves only, nullable: the stored result is z / zdiv[n] (the "/ d" between the two sweeps of
                         // apply_inverse, python/celerite2/numpy.py:94-98, folded into the store); the recursion uses z
  int64_t t_bs, c_bs, y_bs;  // batch strides (0 when shared)
  // scan plumbing
  double *leaves;        // [batch][C][esize]   (summary kernels write here)
  const double *carry;   // [batch][total][ssize]: level-0 part holds carry per chunk (apply kernels read)
  int64_t node_stride;   // nodes per series (Levels::total)
};

struct SweepRevArgs {
  int64_t N, batch, C, L;
  Seg seg;
  int64_t tmin, tmax;
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
  Seg seg;
  int64_t tmin, tmax;
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
  Seg seg;
  int64_t tmin, tmax;
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
// running sums H_m (state after every row) of the rectangular products, K = 1: a.t = t2, a.B = V, a.Y = Y, a.F = H (inst_sweep.cu);
// returns C2B_ERR_INVALID_ARGUMENT without touching anything when the width has no lane-group kernels
template <int JT, bool UPPER>
int launch_gm_states(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream);
// state entering segment `which` of nseg from the segments' summaries; kind: 0 factor, 1 factor_rev (inst_factor.cu),
// 2 solve_lower, 3 solve_lower_rev, 4 solve_upper, 5 solve_upper_rev (inst_sweep.cu, K = 1)
template <int JT>
int launch_seg_carry_factor(int kind, int nseg, int which, const double *summaries, double *out, cudaStream_t stream);
template <int JT>
int launch_seg_carry_sweep(int kind, int nseg, int which, const double *summaries, double *out, cudaStream_t stream);
// CTAs per SM that the four K = 1 lane-group kernels of a solve (summary, apply, and their reverse-mode pair) can all keep
// resident, from the occupancy calculator (registers and shared memory as compiled); 0 for a run-time width (inst_sweep.cu)
template <int JT>
int sweep_resident_ctas();

}  // namespace c2b
