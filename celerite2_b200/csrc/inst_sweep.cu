// Launchers of the sweep kernels for ONE compile-time width (C2B_JT); see args.cuh.
#include "host_common.cuh"
#include <stdlib.h>

#include "sweep.cuh"
#include "gsweep.cuh"
#include "csweep_rev.cuh"

#ifndef C2B_JT
#error "compile with -DC2B_JT=<width> (0 = run-time width)"
#endif

namespace c2b {

// K == 1 solves with a compile-time width up to 8: lane-group kernels (gsweep.cuh).  seg.phase selects the whole op
// or one of the two halves of a segment of a longer series (args.cuh).
template <int JT, bool SOLVE, bool LOWER>
static int launch_sweep_group(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream) {
  if constexpr (JT > 0 && JT <= 16 && !SOLVE) {
    // matmuls: diagonal-affine chunk maps
    constexpr int GP    = kGroupThreadsS / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((a.batch * a.C + GP - 1) / GP);
    if (a.C > 1) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
      C2B_LAUNCH("gsweep_msummary_kernel", stream, launch_k(gsweep_msummary_kernel<JT, LOWER>, grid, kGroupThreadsS, 0, stream, a));
      int rc = run_scan<DiagV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
      if (rc) return rc;
      a.carry = sb.states;
    }
    C2B_LAUNCH("gsweep_apply_kernel", stream, launch_k(gsweep_apply_kernel<JT, SOLVE, LOWER>, grid, kGroupThreadsS, 0, stream, a));
  }
  if constexpr (JT > 0 && JT <= 16 && SOLVE) {
    const int phase = a.seg.phase;
    int rc;
    if (a.C > 1 || phase != 0) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
    }
    if ((phase == 0 && a.C > 1) || phase == 1) {
      constexpr int GPS   = kGroupThreadsS / GroupOf<JT + 1>::G;
      const unsigned grid = (unsigned)((a.batch * a.C + GPS - 1) / GPS);
      C2B_LAUNCH("gsweep_summary_kernel", stream, launch_k(gsweep_summary_kernel<JT, LOWER>, grid, kGroupThreadsS, 0, stream, a));
    }
    if (phase == 1) {
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream, 1))) return rc;
      return seg_export_root<AffV<JT>>(a.J, 1, a.C, sb.nodes, a.seg.summary_out, stream);
    }
    if (phase == 2) {
      if ((rc = seg_import_state<AffV<JT>>(a.J, 1, a.C, sb.states, a.seg.state_in, stream))) return rc;
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, a.seg.state_in, stream, 2))) return rc;
      a.carry = sb.states;
    } else if (a.C > 1) {
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream))) return rc;
      a.carry = sb.states;
    }
    constexpr int GP    = kGroupThreadsS / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((a.batch * a.C + GP - 1) / GP);
    C2B_LAUNCH("gsweep_apply_kernel", stream, launch_k(gsweep_apply_kernel<JT, SOLVE, LOWER>, grid, kGroupThreadsS, 0, stream, a));
  }
  return C2B_OK;
}

template <int JT, bool SOLVE, bool LOWER>
static int launch_sweep_rev_group(SweepRevArgs &a, const ScanBuffers &sb, double *bc, cudaStream_t stream) {
  if constexpr (JT > 0 && JT <= 16 && !SOLVE) {
    constexpr int GP    = kGroupThreadsS / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((a.batch * a.C + GP - 1) / GP);
    if (a.C > 1) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
      C2B_LAUNCH("gsweep_rev_msummary_kernel", stream, launch_k(gsweep_rev_msummary_kernel<JT, LOWER>, grid, kGroupThreadsS, 0, stream, a));
      int rc = run_scan<DiagV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
      if (rc) return rc;
      a.carry = sb.states;
    }
    C2B_LAUNCH("gsweep_rev_apply_kernel", stream, launch_k(gsweep_rev_apply_kernel<JT, SOLVE, LOWER>, grid, kGroupThreadsS, 0, stream, a));
    C2B_LAUNCH("rev_finalize_kernel", stream,
               launch_k(rev_finalize_kernel, dim3((unsigned)((a.C + 1023) / 1024), (unsigned)a.batch), 1024, 0, stream, a.N, a.C, a.L, a.J, LOWER, a.bc_part, a.bt_edge, a.bt, bc, false));
  }
  if constexpr (JT > 0 && JT <= 16 && SOLVE) {
    const int phase = a.seg.phase;
    int rc;
    if (a.C > 1 || phase != 0) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
    }
    if ((phase == 0 && a.C > 1) || phase == 1) {
      constexpr int GPS   = kGroupThreadsS / GroupOf<JT + 1>::G;
      const unsigned grid = (unsigned)((a.batch * a.C + GPS - 1) / GPS);
      C2B_LAUNCH("gsweep_rev_summary_kernel", stream, launch_k(gsweep_rev_summary_kernel<JT, LOWER>, grid, kGroupThreadsS, 0, stream, a));
    }
    if (phase == 1) {
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream, 1))) return rc;
      return seg_export_root<AffV<JT>>(a.J, 1, a.C, sb.nodes, a.seg.summary_out, stream);
    }
    if (phase == 2) {
      if ((rc = seg_import_state<AffV<JT>>(a.J, 1, a.C, sb.states, a.seg.state_in, stream))) return rc;
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, a.seg.state_in, stream, 2))) return rc;
      a.carry = sb.states;
    } else if (a.C > 1) {
      if ((rc = run_scan<AffV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream))) return rc;
      a.carry = sb.states;
    }
    constexpr int GP    = kGroupThreadsS / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((a.batch * a.C + GP - 1) / GP);
    C2B_LAUNCH("gsweep_rev_apply_kernel", stream, launch_k(gsweep_rev_apply_kernel<JT, SOLVE, LOWER>, grid, kGroupThreadsS, 0, stream, a));
    C2B_LAUNCH("rev_finalize_kernel", stream,
               launch_k(rev_finalize_kernel, dim3((unsigned)((a.C + 1023) / 1024), (unsigned)a.batch), 1024, 0, stream, a.N, a.C, a.L, a.J, LOWER, a.bc_part, a.bt_edge, a.bt, bc, false));
    if (phase == 2 && a.seg.edge_out)  // the chunk at the far end of the sweep owes its term to the neighbouring segment
      C2B_CUDA(cudaMemcpyAsync(a.seg.edge_out, a.bt_edge + (LOWER ? 0 : a.C - 1), sizeof(double), cudaMemcpyDeviceToDevice, stream));
  }
  return C2B_OK;
}

// K > 1 sweeps (solves and products) with a compile-time width: one CTA per chunk, one thread per column (csweep_tile_kernel)
template <int JT, bool SOLVE, bool LOWER>
static int launch_sweep_columns(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream) {
  if constexpr (JT > 0) {
    auto threads_for = [](int ncol) {
      int nt = (ncol + 31) / 32 * 32;
      return nt > kColThreads ? kColThreads : nt;
    };
    const unsigned grid = (unsigned)(a.batch * a.C);
    if (a.C > 1) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
      const int nt  = threads_for(SOLVE ? JT + a.K : a.K);
      const size_t smem = csweep_smem_doubles<JT>(nt < a.K ? nt : a.K) * sizeof(double);
      C2B_LAUNCH("csweep_summary_kernel", stream, (launch_k(csweep_tile_kernel<JT, LOWER, true, SOLVE>, grid, nt, smem, stream, a)));
      int rc;
      if (SOLVE)
        rc = run_scan<AffV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
      else
        rc = run_scan<DiagV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
      if (rc) return rc;
      a.carry = sb.states;
    }
    const int nt      = threads_for(a.K);
    const size_t smem = csweep_smem_doubles<JT>(nt < a.K ? nt : a.K) * sizeof(double);
    C2B_LAUNCH("csweep_apply_kernel", stream, (launch_k(csweep_tile_kernel<JT, LOWER, false, SOLVE>, grid, nt, smem, stream, a)));
  }
  return C2B_OK;
}

template <int JT, bool SOLVE, bool LOWER>
int launch_sweep(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream) {
  if (a.seg.phase != 0 && !(JT > 0 && JT <= 16 && SOLVE && a.K == 1)) {
    set_error("segments of a longer series: solve_lower / solve_upper with one right-hand side and J in 1..16 only");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (JT > 0 && JT <= 16 && a.K == 1 && (a.seg.phase != 0 || !no_group_kernels()))
    return launch_sweep_group<JT, SOLVE, LOWER>(a, sb, stream);
  if (JT > 0 && a.K > 1 && !no_group_kernels()) return launch_sweep_columns<JT, SOLVE, LOWER>(a, sb, stream);
  unsigned grid, block;
  size_t smem;
  int rcs;
  if (a.C > 1) {
    a.leaves      = sb.nodes;
    a.node_stride = sb.node_stride;
    const int ncol = SOLVE ? a.J + a.K : a.K;
    if (SOLVE && a.K == 1) {
      rcs = staged_shape(sweep_summary1_kernel<JT, LOWER>, a.batch * a.C, JT > 0, kSweepSlots, SweepRow<JT>::ROWD, &grid, &block, &smem);
      if (rcs) return rcs;
      C2B_LAUNCH("sweep_summary1_kernel", stream, launch_k(sweep_summary1_kernel<JT, LOWER>, grid, block, smem, stream, a));
    } else {
      rcs = staged_shape(sweep_summary_kernel<JT, SOLVE, LOWER>, a.batch * a.C * ncol, JT > 0, kSweepSlots, SweepRow<JT>::ROWD, &grid,
                         &block, &smem);
      if (rcs) return rcs;
      C2B_LAUNCH("sweep_summary_kernel", stream, launch_k(sweep_summary_kernel<JT, SOLVE, LOWER>, grid, block, smem, stream, a));
    }
    int rc;
    if (SOLVE)
      rc = run_scan<AffV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
    else
      rc = run_scan<DiagV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    a.carry = sb.states;
  }
  rcs = staged_shape(sweep_apply_kernel<JT, SOLVE, LOWER>, a.batch * a.C * a.K, JT > 0, kSweepSlots, SweepRow<JT>::ROWD, &grid, &block, &smem);
  if (rcs) return rcs;
  C2B_LAUNCH("sweep_apply_kernel", stream, launch_k(sweep_apply_kernel<JT, SOLVE, LOWER>, grid, block, smem, stream, a));
  return C2B_OK;
}

template <int JT, bool SOLVE, bool LOWER>
int launch_sweep_rev(SweepRevArgs &a, const ScanBuffers &sb, double *bc, cudaStream_t stream) {
  if (a.seg.phase != 0 && !(JT > 0 && JT <= 16 && SOLVE && a.K == 1)) {
    set_error("segments of a longer series: solve_lower_rev / solve_upper_rev with one right-hand side and J in 1..16 only");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (JT > 0 && JT <= 16 && a.K == 1 && (a.seg.phase != 0 || !no_group_kernels()))
    return launch_sweep_rev_group<JT, SOLVE, LOWER>(a, sb, bc, stream);
  unsigned grid, block;
  size_t smem;
  int rcs;
  if (a.C > 1) {
    a.leaves      = sb.nodes;
    a.node_stride = sb.node_stride;
    const int ncol = SOLVE ? a.J + a.K : a.K;
    bool tiled_summary = false;
    if constexpr (JT > 0) {
      if (a.K > 1 && !no_group_kernels()) {
        int nt = (ncol + 31) / 32 * 32;
        if (nt > kColThreads) nt = kColThreads;
        const size_t sm = csweep_rev_summary_smem_doubles<JT>(nt < a.K ? nt : a.K) * sizeof(double);
        if (sm <= 48 * 1024) {
          C2B_LAUNCH("csweep_rev_summary_kernel", stream,
                     (launch_k(csweep_rev_summary_kernel<JT, SOLVE, LOWER>, (unsigned)(a.batch * a.C), nt, sm, stream, a)));
          tiled_summary = true;
        }
      }
    }
    if (tiled_summary) {
    } else if (SOLVE && a.K == 1) {
      rcs = staged_shape(sweep_rev_summary1_kernel<JT, LOWER>, a.batch * a.C, JT > 0, kSweepRevSlots, SweepRevRow<JT>::ROWD, &grid, &block,
                         &smem);
      if (rcs) return rcs;
      C2B_LAUNCH("sweep_rev_summary1_kernel", stream, launch_k(sweep_rev_summary1_kernel<JT, LOWER>, grid, block, smem, stream, a));
    } else {
      rcs = staged_shape(sweep_rev_summary_kernel<JT, SOLVE, LOWER>, a.batch * a.C * ncol, JT > 0, kSweepRevSlots, SweepRevRow<JT>::ROWD,
                         &grid, &block, &smem);
      if (rcs) return rcs;
      C2B_LAUNCH("sweep_rev_summary_kernel", stream, launch_k(sweep_rev_summary_kernel<JT, SOLVE, LOWER>, grid, block, smem, stream, a));
    }
    int rc;
    if (SOLVE)
      rc = run_scan<AffV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
    else
      rc = run_scan<DiagV<JT>>(a.J, a.K, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    a.carry = sb.states;
  }
  bool tiled = false;
  if constexpr (JT > 0) {
    // K > 1, compiled width: CTA per chunk, thread per column, panel sums for the cotangents shared by the columns
    if (a.K > 1 && !no_group_kernels()) {
      int nt = (a.K + 31) / 32 * 32;
      if (nt > kColThreads) nt = kColThreads;
      const size_t sm = csweep_rev_smem_doubles<JT>(nt < a.K ? nt : a.K, nt) * sizeof(double);
      if (sm <= 200 * 1024) {
        if (sm > 48 * 1024)
          C2B_CUDA(cudaFuncSetAttribute(csweep_rev_tile_kernel<JT, SOLVE, LOWER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        C2B_LAUNCH("csweep_rev_apply_kernel", stream,
                   (launch_k(csweep_rev_tile_kernel<JT, SOLVE, LOWER>, (unsigned)(a.batch * a.C), nt, sm, stream, a)));
        tiled = true;
      }
    }
  }
  if (!tiled) {
    rcs = staged_shape(sweep_rev_apply_kernel<JT, SOLVE, LOWER>, a.batch * a.C, JT > 0, kSweepRevSlots, SweepRevRow<JT>::ROWD, &grid, &block,
                       &smem);
    if (rcs) return rcs;
    C2B_LAUNCH("sweep_rev_apply_kernel", stream, launch_k(sweep_rev_apply_kernel<JT, SOLVE, LOWER>, grid, block, smem, stream, a));
  }
  C2B_LAUNCH("rev_finalize_kernel", stream, launch_k(rev_finalize_kernel, dim3((unsigned)((a.C + 1023) / 1024), (unsigned)a.batch), 1024, 0, stream, a.N, a.C, a.L, a.J, LOWER, a.bc_part, a.bt_edge, a.bt, bc, false));
  return C2B_OK;
}

template <int JT, bool UPPER>
int launch_gm_states(SweepArgs &a, const ScanBuffers &sb, cudaStream_t stream) {
  if constexpr (JT > 0 && JT <= 16) {
    constexpr bool LOWER = !UPPER;
    constexpr int GP     = kGroupThreadsS / GroupOf<JT>::G;
    const unsigned grid  = (unsigned)((a.batch * a.C + GP - 1) / GP);
    if (a.C > 1) {
      a.leaves      = sb.nodes;
      a.node_stride = sb.node_stride;
      C2B_LAUNCH("gsweep_msummary_kernel", stream, (launch_k(gsweep_msummary_kernel<JT, LOWER, false>, grid, kGroupThreadsS, 0, stream, a)));
      int rc = run_scan<DiagV<JT>>(a.J, 1, a.batch, a.C, sb.nodes, sb.states, nullptr, stream);
      if (rc) return rc;
      a.carry = sb.states;
    }
    C2B_LAUNCH("gsweep_states_kernel", stream, (launch_k(gsweep_msummary_kernel<JT, LOWER, true>, grid, kGroupThreadsS, 0, stream, a)));
    return C2B_OK;
  } else {
    return C2B_ERR_INVALID_ARGUMENT;
  }
}
template int launch_gm_states<C2B_JT, false>(SweepArgs &, const ScanBuffers &, cudaStream_t);
template int launch_gm_states<C2B_JT, true>(SweepArgs &, const ScanBuffers &, cudaStream_t);

template <int JT>
int launch_seg_carry_sweep(int kind, int nseg, int which, const double *summaries, double *out, cudaStream_t stream) {
  // solve_lower (2) and solve_upper_rev (5) walk upwards, solve_lower_rev (3) and solve_upper (4) downwards
  return run_seg_carry<AffV<JT>>(JT, 1, nseg, which, kind == 3 || kind == 4, summaries, out, stream);
}

template <int JT>
int sweep_resident_ctas() {
  if constexpr (JT > 0 && JT <= 16) {
    static int cached[64] = {0};
    int dev               = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
    if (cached[dev] == 0) {
      int a = 0, b = 0, c = 0, d = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, gsweep_summary_kernel<JT, true>, kGroupThreadsS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, gsweep_apply_kernel<JT, true, true>, kGroupThreadsS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c, gsweep_rev_summary_kernel<JT, true>, kGroupThreadsS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d, gsweep_rev_apply_kernel<JT, true, true>, kGroupThreadsS, 0);
      int m = a < b ? a : b;
      m     = m < c ? m : c;
      m     = m < d ? m : d;
      cached[dev] = m > 0 ? m : -1;
    }
    return cached[dev] > 0 ? cached[dev] : 0;
  } else {
    return 0;
  }
}
template int sweep_resident_ctas<C2B_JT>();
template int launch_seg_carry_sweep<C2B_JT>(int, int, int, const double *, double *, cudaStream_t);

#define INST(S, L)                                                                                                                        \
  template int launch_sweep<C2B_JT, S, L>(SweepArgs &, const ScanBuffers &, cudaStream_t);                                                \
  template int launch_sweep_rev<C2B_JT, S, L>(SweepRevArgs &, const ScanBuffers &, double *, cudaStream_t);
INST(true, true)
INST(true, false)
INST(false, true)
INST(false, false)

}  // namespace c2b
