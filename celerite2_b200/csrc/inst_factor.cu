// Launchers of the factor / factor_rev kernels for ONE compile-time width (C2B_JT); see args.cuh.
#include <stdlib.h>

#include "host_common.cuh"
#include "factor.cuh"
#include "gfactor.cuh"
#include "gfactor_rev.cuh"

#ifndef C2B_JT
#error "compile with -DC2B_JT=<width> (0 = run-time width)"
#endif

namespace c2b {

// compile-time widths up to 16: lane-group kernels (gfactor.cuh).  seg.phase: see args.cuh.
template <int JT>
static int launch_factor_group(FactorArgs &g, const ScanBuffers &sb, cudaStream_t stream) {
  if constexpr (JT > 0 && JT <= 16) {
    constexpr int GP    = kGroupThreads / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((g.batch * g.C + GP - 1) / GP);
    const int phase     = g.seg.phase;
    int rc;
    if (g.C > 1 || phase != 0) {
      g.leaves      = sb.nodes;
      g.node_stride = sb.node_stride;
    }
    if ((phase == 0 && g.C > 1) || phase == 1)
      C2B_LAUNCH("gfactor_summary_kernel", stream, launch_k(gfactor_summary_kernel<JT>, grid, kGroupThreads, 0, stream, g));
    if (phase == 1) {
      if ((rc = run_scan<Lft<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream, 1))) return rc;
      return seg_export_root<Lft<JT>>(g.J, 1, g.C, sb.nodes, g.seg.summary_out, stream);
    }
    if (phase == 2) {
      if ((rc = seg_import_state<Lft<JT>>(g.J, 1, g.C, sb.states, g.seg.state_in, stream))) return rc;
      if ((rc = run_scan<Lft<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, g.seg.state_in, stream, 2))) return rc;
      g.carry = sb.states;
    } else if (g.C > 1) {
      if ((rc = run_scan<Lft<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream))) return rc;
      g.carry = sb.states;
    }
    // workspace rows straight from registers (16-byte stores of each lane's column): r02b / r02e, -7 % on the factor
    // kernels against staging them in shared memory first; C2B_S_STAGED=1 keeps the staged form for A/B runs
    static const bool direct_s = getenv("C2B_S_STAGED") == nullptr;
    // Chunked series: rows collected in shared memory and written by the TMA engine (bulk stores, gfactor.cuh); r02k:
    // factor apply 0.121 -> 0.107 ms (c2), 1.55 -> 1.45 ms (c5).  Whole series per lane group (c3) are faster with the
    // direct stores (5.70 vs 6.23 ms).  C2B_S_BULK=0 switches the bulk copies of factor and factor_rev off.
    static const bool bulk_s = getenv("C2B_S_BULK") == nullptr || atoi(getenv("C2B_S_BULK")) != 0;
    if constexpr (JT % 2 == 0) {
      if (bulk_s && g.C > 1 && g.S && (reinterpret_cast<uintptr_t>(g.S) & 15) == 0) {
        C2B_LAUNCH("gfactor_apply_kernel", stream, (launch_k(gfactor_apply_kernel<JT, false, true>, grid, kGroupThreads, 0, stream, g)));
        return C2B_OK;
      }
    }
    if (direct_s && g.S)
      C2B_LAUNCH("gfactor_apply_kernel", stream, (launch_k(gfactor_apply_kernel<JT, true>, grid, kGroupThreads, 0, stream, g)));
    else
      C2B_LAUNCH("gfactor_apply_kernel", stream, (launch_k(gfactor_apply_kernel<JT, false>, grid, kGroupThreads, 0, stream, g)));
  }
  return C2B_OK;
}

template <int JT>
int launch_factor(FactorArgs &g, const ScanBuffers &sb, cudaStream_t stream) {
  if (g.seg.phase != 0 && !(JT > 0 && JT <= 16)) {
    set_error("segments of a longer series need a compiled width (J in 1..16)");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (JT > 0 && JT <= 16 && (g.seg.phase != 0 || !no_group_kernels())) return launch_factor_group<JT>(g, sb, stream);
  unsigned grid, block;
  size_t smem;
  int rcs;
  if (g.C > 1) {
    g.leaves      = sb.nodes;
    g.node_stride = sb.node_stride;
    rcs = staged_shape(factor_summary_kernel<JT>, g.batch * g.C, JT > 0, kFactorSlots, FactorRow<JT>::ROWD, &grid, &block, &smem);
    if (rcs) return rcs;
    C2B_LAUNCH("factor_summary_kernel", stream, launch_k(factor_summary_kernel<JT>, grid, block, smem, stream, g));
    int rc = run_scan<Lft<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    g.carry = sb.states;
  }
  rcs = staged_shape(factor_apply_kernel<JT>, g.batch * g.C, JT > 0, kFactorSlots, FactorRow<JT>::ROWD, &grid, &block, &smem);
  if (rcs) return rcs;
  C2B_LAUNCH("factor_apply_kernel", stream, launch_k(factor_apply_kernel<JT>, grid, block, smem, stream, g));
  return C2B_OK;
}

template <int JT>
static int launch_factor_rev_group(FactorRevArgs &g, const ScanBuffers &sb, double *bc, cudaStream_t stream) {
  if constexpr (JT > 0 && JT <= 16) {
    constexpr int GP    = kGroupThreadsR / GroupOf<JT>::G;
    const unsigned grid = (unsigned)((g.batch * g.C + GP - 1) / GP);
    const int phase     = g.seg.phase;
    int rc;
    if (g.C > 1 || phase != 0) {
      g.leaves      = sb.nodes;
      g.node_stride = sb.node_stride;
    }
    if ((phase == 0 && g.C > 1) || phase == 1)
      C2B_LAUNCH("gfactor_rev_summary_kernel", stream, launch_k(gfactor_rev_kernel<JT, false>, grid, kGroupThreadsR, 0, stream, g));
    if (phase == 1) {
      if ((rc = run_scan<Cong<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream, 1))) return rc;
      return seg_export_root<Cong<JT>>(g.J, 1, g.C, sb.nodes, g.seg.summary_out, stream);
    }
    if (phase == 2) {
      if ((rc = seg_import_state<Cong<JT>>(g.J, 1, g.C, sb.states, g.seg.state_in, stream))) return rc;
      if ((rc = run_scan<Cong<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, g.seg.state_in, stream, 2))) return rc;
      g.carry = sb.states;
    } else if (g.C > 1) {
      if ((rc = run_scan<Cong<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream))) return rc;
      g.carry = sb.states;
    }
    // workspace rows of a tile by one bulk copy (gfactor_rev.cuh); r02k: 0.156 -> 0.154 ms (c2), 2.49 -> 2.38 ms (c5)
    static const bool bulk_s = getenv("C2B_S_BULK") == nullptr || atoi(getenv("C2B_S_BULK")) != 0;
    bool launched            = false;
    if constexpr (JT % 2 == 0) {
      if (bulk_s && (reinterpret_cast<uintptr_t>(g.S) & 15) == 0) {
        C2B_LAUNCH("gfactor_rev_apply_kernel", stream, (launch_k(gfactor_rev_kernel<JT, true, true>, grid, kGroupThreadsR, 0, stream, g)));
        launched = true;
      }
    }
    if (!launched) C2B_LAUNCH("gfactor_rev_apply_kernel", stream, launch_k(gfactor_rev_kernel<JT, true>, grid, kGroupThreadsR, 0, stream, g));
    C2B_LAUNCH("rev_finalize_kernel", stream,
               launch_k(rev_finalize_kernel, dim3((unsigned)((g.C + 1023) / 1024), (unsigned)g.batch), 1024, 0, stream, g.N, g.C, g.L, g.J, true, g.bc_part, g.bt_edge, g.bt, bc, g.acc != 0));
    if (phase == 2 && g.seg.edge_out)
      C2B_CUDA(cudaMemcpyAsync(g.seg.edge_out, g.bt_edge, sizeof(double), cudaMemcpyDeviceToDevice, stream));
  }
  return C2B_OK;
}

template <int JT>
int launch_factor_rev(FactorRevArgs &g, const ScanBuffers &sb, double *bc, cudaStream_t stream) {
  if (g.seg.phase != 0 && !(JT > 0 && JT <= 16)) {
    set_error("segments of a longer series need a compiled width (J in 1..16)");
    return C2B_ERR_INVALID_ARGUMENT;
  }
  if (JT > 0 && JT <= 16 && (g.seg.phase != 0 || !no_group_kernels())) return launch_factor_rev_group<JT>(g, sb, bc, stream);
  unsigned grid, block;
  size_t smem;
  int rcs;
  if (g.C > 1) {
    g.leaves      = sb.nodes;
    g.node_stride = sb.node_stride;
    rcs = staged_shape(factor_rev_summary1_kernel<JT>, g.batch * g.C, JT > 0, kFactorRevSlots, FactorRevRow<JT>::ROWD, &grid, &block, &smem);
    if (rcs) return rcs;
    C2B_LAUNCH("factor_rev_summary1_kernel", stream, launch_k(factor_rev_summary1_kernel<JT>, grid, block, smem, stream, g));
    int rc = run_scan<Cong<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    g.carry = sb.states;
  }
  rcs = staged_shape(factor_rev_apply_kernel<JT>, g.batch * g.C, JT > 0, kFactorRevSlots, FactorRevRow<JT>::ROWD, &grid, &block, &smem);
  if (rcs) return rcs;
  C2B_LAUNCH("factor_rev_apply_kernel", stream, launch_k(factor_rev_apply_kernel<JT>, grid, block, smem, stream, g));
  C2B_LAUNCH("rev_finalize_kernel", stream, launch_k(rev_finalize_kernel, dim3((unsigned)((g.C + 1023) / 1024), (unsigned)g.batch), 1024, 0, stream, g.N, g.C, g.L, g.J, true, g.bc_part, g.bt_edge, g.bt, bc, g.acc != 0));
  return C2B_OK;
}

template <int JT>
int launch_seg_carry_factor(int kind, int nseg, int which, const double *summaries, double *out, cudaStream_t stream) {
  // factor walks upwards (earlier segments = lower rows), factor_rev downwards
  if (kind == 0) return run_seg_carry<Lft<JT>>(JT, 1, nseg, which, false, summaries, out, stream);
  return run_seg_carry<Cong<JT>>(JT, 1, nseg, which, true, summaries, out, stream);
}

template int launch_seg_carry_factor<C2B_JT>(int, int, int, const double *, double *, cudaStream_t);
template int launch_factor<C2B_JT>(FactorArgs &, const ScanBuffers &, cudaStream_t);
template int launch_factor_rev<C2B_JT>(FactorRevArgs &, const ScanBuffers &, double *, cudaStream_t);

}  // namespace c2b
