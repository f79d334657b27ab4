// Launchers of the factor / factor_rev kernels for ONE compile-time width (C2B_JT); see args.cuh.
#include "host_common.cuh"
#include "factor.cuh"

#ifndef C2B_JT
#error "compile with -DC2B_JT=<width> (0 = run-time width)"
#endif

namespace c2b {

template <int JT>
int launch_factor(FactorArgs &g, const ScanBuffers &sb, cudaStream_t stream) {
  unsigned grid, block;
  launch_shape(g.batch * g.C, &grid, &block);
  if (g.C > 1) {
    g.leaves      = sb.nodes;
    g.node_stride = sb.node_stride;
    C2B_LAUNCH("factor_summary_kernel", stream, factor_summary_kernel<JT><<<grid, block, 0, stream>>>(g));
    int rc = run_scan<Lft<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    g.carry = sb.states;
  }
  C2B_LAUNCH("factor_apply_kernel", stream, factor_apply_kernel<JT><<<grid, block, 0, stream>>>(g));
  return C2B_OK;
}

template <int JT>
int launch_factor_rev(FactorRevArgs &g, const ScanBuffers &sb, double *bc, cudaStream_t stream) {
  unsigned grid, block;
  if (g.C > 1) {
    g.leaves      = sb.nodes;
    g.node_stride = sb.node_stride;
    launch_shape(g.batch * g.C * (g.J + 1), &grid, &block);
    C2B_LAUNCH("factor_rev_summary_kernel", stream, factor_rev_summary_kernel<JT><<<grid, block, 0, stream>>>(g));
    int rc = run_scan<Cong<JT>>(g.J, 1, g.batch, g.C, sb.nodes, sb.states, nullptr, stream);
    if (rc) return rc;
    g.carry = sb.states;
  }
  launch_shape(g.batch * g.C, &grid, &block);
  C2B_LAUNCH("factor_rev_apply_kernel", stream, factor_rev_apply_kernel<JT><<<grid, block, 0, stream>>>(g));
  C2B_LAUNCH("rev_finalize_kernel", stream, rev_finalize_kernel<<<(unsigned)g.batch, 256, 0, stream>>>(g.N, g.C, g.L, g.J, true, g.bc_part, g.bt_edge, g.bt, bc));
  return C2B_OK;
}

template int launch_factor<C2B_JT>(FactorArgs &, const ScanBuffers &, cudaStream_t);
template int launch_factor_rev<C2B_JT>(FactorRevArgs &, const ScanBuffers &, double *, cudaStream_t);

}  // namespace c2b
