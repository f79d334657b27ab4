// Lane-private TMA tile pipeline for the thread-per-chunk recursions.
//
// One thread walks one chunk row by row (the recursion is sequential along n), so a warp reads and writes 32
// streams that lie a full chunk length apart.  Doing that with per-lane loads and stores makes every memory
// instruction touch 32 different cache lines; the r01b profile shows those kernels bound by L1/LSU wavefronts
// (116 cycles per row per SM against a 20-cycle HBM budget), not by DRAM bandwidth or FP64 issue.
//
// Here every lane moves ITS chunk in tiles of T rows with the bulk-copy engine instead:
//   load : one cp.async.bulk (TMA, UBLKCP in SASS) per array and tile, T x w x 8 contiguous bytes from global
//          into the lane's slot of shared memory, completion signalled on the lane's own mbarrier; tiles are
//          double-buffered, so tile i+1 streams in while tile i is computed;
//   store: the lane writes its results into its output slot and hands the finished tile to one bulk store per
//          array (cp.async.bulk.global.shared::cta), tracked with bulk groups.
// No LSU wavefronts, no address arithmetic per element, full-line DRAM bursts.  Bulk copies need 16-byte aligned
// addresses and sizes; a piece that is not (odd row counts at a series end, odd J with odd offsets) is moved with
// plain loads/stores by the lane -- correct, just slow, and rare (chunk lengths and T are even).
// Rows sit in MEMORY order inside a tile; descending sweeps read them backwards.
#pragma once
#include "common.cuh"

namespace c2b {

C2B_DEV unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

C2B_DEV void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
C2B_DEV void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
C2B_DEV void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
C2B_DEV void bulk_load(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
C2B_DEV void bulk_store(void *gmem_dst, const void *smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
C2B_DEV void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
C2B_DEV void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
C2B_DEV void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
C2B_DEV void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Geometry of one lane's chunk.
struct ChunkGeom {
  int64_t n0;       // first row
  int cnt;          // rows (0 = this lane has no chunk)
  int64_t N;        // rows of the series
  bool descending;  // processing order
};

// T rows per tile; SIN / SOUT = doubles per lane slot for one input stage / the output tile (multiples of 2).
template <int T, int SIN, int SOUT>
struct TilePipe {
  static_assert(T % 2 == 0 && SIN % 2 == 0 && SOUT % 2 == 0, "16-byte granularity");
  static constexpr int kThreads = 32;  // one warp per CTA
  // shared memory per CTA, in bytes
  static constexpr size_t kSmemBytes = (size_t)kThreads * (2 * SIN + SOUT) * 8 + (size_t)kThreads * 2 * 8;

  ChunkGeom g;
  double *sin0, *sin1, *sout;
  unsigned long long *bar0, *bar1;
  unsigned phase0, phase1;
  unsigned pending_tx;  // bytes announced for the tile being fetched
  int ntiles;

  C2B_DEV void init(double *smem, const ChunkGeom &geom) {
    g              = geom;
    const int lane = threadIdx.x;
    // slot layout [lane][...] with a stride that is 2 mod 4 doubles where possible: 16-byte aligned for the bulk
    // engine, at most 2-way bank conflicts for the lane-private 8-byte accesses of the compute phase
    sin0   = smem + (size_t)lane * SIN;
    sin1   = smem + (size_t)kThreads * SIN + (size_t)lane * SIN;
    sout   = smem + (size_t)2 * kThreads * SIN + (size_t)lane * SOUT;
    bar0   = (unsigned long long *)(smem + (size_t)kThreads * (2 * SIN + SOUT)) + 2 * lane;
    bar1   = bar0 + 1;
    phase0 = phase1 = 0;
    ntiles          = (g.cnt + T - 1) / T;
    mbar_init(bar0, 1);
    mbar_init(bar1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_async_smem();
    __syncwarp();
  }

  // rows of tile `tile` that exist in the chunk, and the memory row the tile starts at after clipping
  C2B_DEV int rows_in_tile(int tile) const {
    const int rem = g.cnt - tile * T;
    return rem <= 0 ? 0 : (rem < T ? rem : T);
  }
  // unclipped first memory row of the tile (may be below n0 for the last, partial tile of a descending sweep)
  C2B_DEV int64_t tile_lo(int tile) const { return g.descending ? g.n0 + g.cnt - (int64_t)(tile + 1) * T : g.n0 + (int64_t)tile * T; }
  C2B_DEV int64_t row_of(int tile, int rp) const {
    const int64_t pr = (int64_t)tile * T + rp;
    return g.descending ? g.n0 + g.cnt - 1 - pr : g.n0 + pr;
  }

  // The slot about to be refilled was read with ordinary loads; order those before the bulk engine's writes.
  C2B_DEV void begin_fetch() {
    fence_async_smem();
    pending_tx = 0;
  }
  // One input segment: W doubles per row, HLO / HHI halo rows below / above the tile (0 or 1).  `series` points at
  // row 0 of this lane's series.  The slot keeps T + HLO + HHI rows at offset `off`; slot row 0 <-> memory row
  // tile_lo - HLO.
  template <int W, int HLO, int HHI>
  C2B_DEV void fetch(int stage, int tile, const double *series, int off) {
    if (g.cnt == 0 || tile >= ntiles) return;
    int64_t lo = tile_lo(tile) - HLO, hi = tile_lo(tile) + T + HHI;
    const int64_t cl = g.n0 - HLO < 0 ? 0 : g.n0 - HLO;
    const int64_t ch = g.n0 + g.cnt + HHI > g.N ? g.N : g.n0 + g.cnt + HHI;
    const int64_t l2 = lo < cl ? cl : lo, h2 = hi > ch ? ch : hi;
    if (h2 <= l2) return;
    double *dst          = (stage ? sin1 : sin0) + off + (l2 - lo) * W;
    const double *src    = series + l2 * W;
    const unsigned bytes = (unsigned)((h2 - l2) * W * 8);
    if ((((uintptr_t)src | (uintptr_t)dst | bytes) & 15u) == 0) {
      pending_tx += bytes;
      bulk_load(dst, src, bytes, stage ? bar1 : bar0);
    } else {
      const int n = (int)((h2 - l2) * W);
      for (int e = 0; e < n; ++e) dst[e] = src[e];
    }
  }
  C2B_DEV void end_fetch(int stage) {
    // one arrival per tile and lane; announces the bytes the bulk copies of this tile will deliver
    mbar_expect_tx(stage ? bar1 : bar0, pending_tx);
  }
  C2B_DEV void wait_tile(int stage) {
    if (stage) {
      mbar_wait(bar1, phase1);
      phase1 ^= 1;
    } else {
      mbar_wait(bar0, phase0);
      phase0 ^= 1;
    }
  }

  // lane-private accessors; rp = row inside the tile in processing order
  template <int W, int HLO>
  C2B_DEV double in(int stage, int off, int rp, int x, int shift = 0) const {
    const int r = (g.descending ? T - 1 - rp : rp) + HLO + shift;
    return (stage ? sin1 : sin0)[off + r * W + x];
  }
  template <int W>
  C2B_DEV void out(int off, int rp, int x, double v) {
    const int r         = g.descending ? T - 1 - rp : rp;
    sout[off + r * W + x] = v;
  }

  // Before the first out() of a tile: the previous tile's bulk stores must have finished reading the slot.
  C2B_DEV void begin_tile_outputs() { bulk_wait_read(); }
  // after the last out() of a tile and before its first store(): order the lane's writes before the bulk engine's reads
  C2B_DEV void begin_stores() { fence_async_smem(); }
  // One output segment of the finished tile; rows_done = rows of the tile (processing order) that are valid.
  template <int W>
  C2B_DEV void store(int tile, double *series, int off, int rows_done) {
    if (rows_done <= 0) return;
    // valid memory rows: processing rows [0, rows_done) of this tile
    int64_t l2, h2;
    if (g.descending) {
      h2 = tile_lo(tile) + T;
      l2 = h2 - rows_done;
    } else {
      l2 = tile_lo(tile);
      h2 = l2 + rows_done;
    }
    const double *src    = sout + off + (l2 - tile_lo(tile)) * W;
    double *dst          = series + l2 * W;
    const unsigned bytes = (unsigned)((h2 - l2) * W * 8);
    if ((((uintptr_t)src | (uintptr_t)dst | bytes) & 15u) == 0) {
      bulk_store(dst, src, bytes);
    } else {
      const int n = (int)((h2 - l2) * W);
      for (int e = 0; e < n; ++e) dst[e] = src[e];
    }
  }
  // after the last store() of a tile
  C2B_DEV void end_tile_outputs() { bulk_commit(); }
  C2B_DEV void finish() { bulk_wait_all(); }
};

}  // namespace c2b
