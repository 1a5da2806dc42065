// bnl_pairs.cuh -- all two-way tables of a column set in one tensor-core pass.
//
// When a batch asks for the table of (nearly) every PAIR of a set of factor columns
// -- the first score_cache_fill of hc() (every i -> j of the empty graph,
// src/learning/score/hc.cache.lookup.c:52-72), level 0 of pc.stable (every x, y
// marginally, R/pcalgo.R:8-28) or mi_matrix_discrete (src/learning/local/
// mi.matrix.c:40-62) -- counting unit by unit repeats the row pass once per pair
// and sits on the shared-memory atomic rate. The same numbers are ONE contraction:
// with H the N x L 0/1 indicator matrix of the selected columns (one column of H
// per (factor, level)), G = H' H holds n[x_i = a, x_j = b] for every pair at once,
// and the marginal counts on its diagonal. Rows with an NA in a column have an
// all-zero indicator block there, so they drop out of exactly the pairs that
// involve that column, as fill_2d_table skips them
// (src/core/contingency.tables.c:93-130).
//
// G is formed on the 5th-generation tensor cores (tcgen05.mma kind::i8, u8 x u8 ->
// s32 accumulators in TMEM: exact, a cell is a row count < 2^31). H never exists
// in HBM: producer warps read the packed uint8 code columns (16 rows per 128-bit
// load), expand them to 0/1 operand bytes with byte-permutes and write them
// straight into the 128-byte-swizzled K-major shared-memory layout the MMA reads
// through its matrix descriptors.
#pragma once
#include "bnl_internal.h"

#define PAIR_BM 128                 // indicator rows of the A operand = TMEM lanes
#define PAIR_KROWS 128              // data rows per pipeline stage = 128 bytes of K per indicator row
#define PAIR_STAGES 4
// NPW producer warps expand codes and one more warp issues the MMAs (template
// parameter). A producer warp's own instruction stream is the critical path of a
// stage -- one warp issues about once in five cycles on its dependency chains -- so
// the wide tile runs 19 + 1 warps with 3 columns per thread rather than 11 + 1 with
// 5 (registers are allotted per 4 warps: the total is kept a multiple of 4).
#define PAIR_MAX_LEVELS 8           // one byte-permute selects among 8 table bytes

namespace bnlpair {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// try_wait parks the thread in hardware until the phase completes or the hint (ns)
// runs out, so a waiting warp takes no issue slots from the producers (a polling
// loop in the MMA warp cost a quarter of all instructions issued). Waits are
// bounded: a protocol slip must end in a trap, not in a hung device.
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
  for (int tries = 0; tries < 4096; tries++) {
    uint32_t done;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(smem_u32(b)), "r"(parity), "r"(1000000u) : "memory");
    if (done) return;
  }
  __trap();
}

// The MMA issuer must react within a fraction of a stage time (~0.3 us): it polls
// without a suspend hint (a parked thread wakes too late, the tensor pipe then idles
// until a burst of ready stages has piled up). One thread; ~3 instructions per poll.
__device__ __forceinline__ void mbar_wait_spin(uint64_t *b, uint32_t parity) {
  for (int tries = 0; tries < (1 << 28); tries++) {
    uint32_t done;
    asm volatile("{ .reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
    if (done) return;
  }
  __trap();
}

// shared-memory matrix descriptor, K-major operand in the 128-byte swizzle: rows of
// 128 bytes, 8-row groups 1024 bytes apart (stride byte offset), descriptor
// version 1 (sm_100), layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0;\n"
               "  tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p; }"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the barrier once every MMA issued so far has finished reading shared memory / writing TMEM
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// One-hot bytes of 4 rows for one level from their packed codes: `sel` holds the
// four codes as nibbles, the 8-byte table {lo, hi} has a 1 at the level's byte. A nibble of
// 0xF (NA, padding) selects "sign of byte 7", which is 0.
__device__ __forceinline__ uint32_t onehot4(uint32_t sel, uint32_t lo, uint32_t hi) {
  uint32_t r;      // prmt in PTX's generic mode: __byte_perm() drops bit 3 of a selector nibble
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(sel));
  return r;
}
// NA = true: some selected column holds NA bytes (0xFF) between valid codes; they are
// cut to 0xF first so that they cannot spill into the neighbouring code's nibble.
// Without NA the only 0xFF bytes are the padding past the last row, which sits at the
// HIGH end of a word and only overwrites nibbles that are 0xF anyway.
template <bool NA>
__device__ __forceinline__ uint32_t selector4(uint32_t w) {
  const uint32_t m = NA ? (w & 0x0F0F0F0Fu) : w;   // valid codes are < 8; NA 0xFF -> 0xF
  const uint32_t t = m | (m >> 4);                 // byte 0: codes 0, 1; byte 2: codes 2, 3
  return __byte_perm(t, 0u, 0x4420);
}
// predicated store: no branch around it
__device__ __forceinline__ void sts128_if(uint32_t on, uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("{ .reg .pred p; setp.ne.b32 p, %0, 0; @p st.shared.v4.b32 [%1], {%2, %3, %4, %5}; }"
               ::"r"(on), "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

}  // namespace bnlpair

// Tile (bm, bn) of G: indicator rows [bm * 128, +128) against indicator rows
// [bn * BN, +BN), over the data rows of one split. grid = tiles x splits, tile
// fastest, so the CTAs resident at one time work on the same slab of rows and a
// column's bytes come from HBM once. One CTA per SM.
//   warps 0..NPW-1  expand codes into the stage buffers; warps 0-7 then drain the accumulator
//   warp  NPW       waits for full stages and issues 4 MMAs (K = 32 each) per stage
// cols[] lists the selected columns in indicator order: {packed slot, first indicator, levels}.
//
// The stage loop is written without branches. A producer warp's own instruction
// stream is the critical path of a stage (12 warps, 3 per scheduler: one warp issues
// about once in five cycles), and in the first versions most of that stream was
// control flow -- an EMPTY stage loop (no loads, no expansion, no MMA) already took
// 4.1 of the kernel's 7.0 ms. So: every thread owns ITEMS column slots whether or not
// they are used (an unused slot loads column 0 and stores nothing), stores and loads
// are predicated instead of skipped, the trip count is padded to a multiple of the
// ring depth (a padded stage stores nothing and its MMAs are predicated off), and the
// levels past the fourth are compiled in only for the WIDE variant.
template <int BN, int ITEMS, int DEPTH, bool NA, bool WIDE, int NPW, int SUP, int NSTG>
__global__ void __launch_bounds__((NPW + 1) * 32, 1)
pair_onehot_kernel(const uint8_t *__restrict__ u8, size_t stride, const int4 *__restrict__ cols,
                   const int *__restrict__ ind2col, int L, const int2 *__restrict__ tiles, int ntiles,
                   int stages_per_split, int total_stages, int *__restrict__ G, int ldg, int atomic_out) {
  using namespace bnlpair;
  // NSTG operand buffers. NSTG > PAIR_STAGES marks the single-tile launch (L <= 128): its one tile is
  // diagonal, the A rows ARE the B rows, so a buffer holds the B rows only and twice as many fit.
  constexpr bool DIAG = NSTG > PAIR_STAGES;
  constexpr int A_BYTES = DIAG ? 0 : PAIR_BM * 128, B_BYTES = BN * 128, STAGE_BYTES = A_BYTES + B_BYTES;
  // SUP operand buffers share one full / empty barrier pair ("one handshake per SUP K-blocks")
  constexpr int NS = NSTG, NSS = NSTG / SUP;
  static_assert((NSS & (NSS - 1)) == 0 && NSS >= 2, "barrier ring indexed with a mask");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // the dynamic window is only guaranteed 16-byte alignment: round up to the swizzle atom
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full_bar[NS], empty_bar[NS], accum_bar;
  __shared__ uint32_t s_tmem;
  __shared__ int s_slot[256], s_row[256], s_nlev[256], s_off[256];
  __shared__ int s_range[4];

  const int tile = blockIdx.x % ntiles, split = blockIdx.x / ntiles;
  const int bm = tiles[tile].x, bn = tiles[tile].y;
  const int a0 = bm * PAIR_BM, b0 = bn * BN;
  // a diagonal tile: the A rows are a slice of the B rows, expanded once
  const bool a_in_b = a0 >= b0 && a0 + PAIR_BM <= b0 + BN;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NSS; s++) { mbar_init(&full_bar[s], NPW); mbar_init(&empty_bar[s], 1); }
    mbar_init(&accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // columns that reach into the B rows, then (unless shared) into the A rows
    s_range[0] = ind2col[b0];
    s_range[1] = ind2col[min(b0 + BN, L) - 1];
    s_range[2] = a_in_b ? 1 : ind2col[a0];
    s_range[3] = a_in_b ? 0 : ind2col[min(a0 + PAIR_BM, L) - 1];
  }
  if (warp == NPW) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "n"(BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  // rows that no column writes (beyond L, or levels outside the tile) must read as 0
  for (int i = threadIdx.x; i < NS * STAGE_BYTES / 16; i += (NPW + 1) * 32)
    reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  const int nB = s_range[1] - s_range[0] + 1, nA = s_range[3] - s_range[2] + 1, nitems = nB + nA;
  if (nitems > ITEMS * NPW * 4 || nitems > 256) __trap();      // the host's layout bounds the columns per tile (bnl_pairs.cu)
  for (int q = threadIdx.x; q < nitems; q += (NPW + 1) * 32) {     // the work list, one column per thread
    const bool inB = q < nB;
    const int4 ci = cols[inB ? s_range[0] + q : s_range[2] + q - nB];
    s_slot[q] = ci.x; s_row[q] = ci.y - (inB ? b0 : a0); s_nlev[q] = ci.z;
    s_off[q] = (inB ? A_BYTES : 0) | ((inB ? BN : PAIR_BM) << 20);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;

  const int st0 = split * stages_per_split;
  const int nst = min(st0 + stages_per_split, total_stages) - st0;     // >= 1 by the host's split
  const int nsup = (nst + SUP - 1) / SUP;                              // handshakes
  const int npad = (nsup + DEPTH - 1) / DEPTH * DEPTH;                 // trip count of both roles

  if (warp < NPW) {
    // ---------------- producers ----------------
    // Thread = (16-row chunk g, column lane cq): the 8 threads of a column read 128
    // contiguous bytes and write one whole 128-byte operand row per level (the swizzle
    // spreads their 16-byte pieces over all banks). The loads run DEPTH stages ahead in
    // a register ring -- one CTA per SM, so nothing else hides the memory latency -- and
    // the lines 16 stages ahead are pulled from DRAM into L2.
    const int g = threadIdx.x & 7, cq = threadIdx.x >> 3;
    // per column slot of this thread, fixed for the whole kernel: the column's start
    // (coff, 16-byte units), where its rows go inside a stage (sa), which levels fall
    // inside the tile (lm, one bit per level; 0: unused slot), and the swizzled 16-byte
    // piece of this thread's chunk in each of those rows (sw: one byte per level), so
    // that a level costs one byte-extract, one add, four byte-permutes and the store.
    uint32_t coff[ITEMS], sa[ITEMS], lm[ITEMS], sw_lo[ITEMS], sw_hi[ITEMS];
    const uint4 *base4 = reinterpret_cast<const uint4 *>(u8);
    const uint32_t stride16 = (uint32_t)(stride >> 4);
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
      const int q = cq + (NPW * 4) * i;
      const bool ok = q < nitems;
      const int row = ok ? s_row[q] : 0, nl = ok ? s_nlev[q] : 0, nr = ok ? (s_off[q] >> 20) : 0;
      coff[i] = (uint32_t)(ok ? s_slot[q] : 0) * stride16 + (uint32_t)g;
      sa[i] = (uint32_t)((ok ? (s_off[q] & 0xFFFFF) : 0) + row * 128);
      const int v0 = max(0, -row), v1 = min(nl, nr - row);
      lm[i] = v1 > v0 ? ((1u << v1) - (1u << v0)) : 0u;
      uint32_t lo = 0, hi = 0;
#pragma unroll
      for (int v = 0; v < 4; v++) {
        lo |= (uint32_t)((g ^ ((row + v) & 7)) << 4) << (8 * v);
        hi |= (uint32_t)((g ^ ((row + v + 4) & 7)) << 4) << (8 * v);
      }
      sw_lo[i] = lo; sw_hi[i] = hi;
    }
    // The CTAs resident at one time work on the same slab of rows (so that it stays in
    // L2); each starts at its own place in it and wraps around.
    const int rot = (int)(((long long)tile * nst) / ntiles);
    uint4 ring[DEPTH][SUP][ITEMS];
    auto load = [&](uint4(&r)[ITEMS], int it) {          // it < nst: a real stage
      int w = it + rot;
      w -= w >= nst ? nst : 0;
      const uint32_t so = (uint32_t)(st0 + w) * (PAIR_KROWS / 16);
#pragma unroll
      for (int i = 0; i < ITEMS; i++) r[i] = __ldg(base4 + (coff[i] + so));
      int wp = w + 16;
      wp -= wp >= nst ? nst : 0;
      const uint32_t sp = (uint32_t)(st0 + wp) * (PAIR_KROWS / 16);
      const uint32_t pf = g == 0 && it + 16 < nst;
#pragma unroll
      for (int i = 0; i < ITEMS; i++)
        asm volatile("{ .reg .pred p; setp.ne.b32 p, %1, 0; @p prefetch.global.L2 [%0]; }"
                     ::"l"(base4 + (coff[i] + sp)), "r"(pf & (lm[i] != 0u)));
    };
#pragma unroll
    for (int d = 0; d < DEPTH; d++)
#pragma unroll
      for (int h = 0; h < SUP; h++) load(ring[d][h], min(d * SUP + h, nst - 1));
    const uint32_t smem0 = smem_u32(smem);
    for (int su0 = 0; su0 < npad; su0 += DEPTH) {
#pragma unroll
      for (int d = 0; d < DEPTH; d++) {
        const int su = su0 + d;
        const int slot = su & (NSS - 1);
        if (su >= NSS) mbar_wait_spin(&empty_bar[slot], ((su / NSS) - 1) & 1);
#pragma unroll
        for (int h = 0; h < SUP; h++) {
          const int it = su * SUP + h;
          const uint32_t sb = smem0 + (uint32_t)(slot * SUP + h) * STAGE_BYTES;
          const uint32_t live = it < nst ? 0xFFu : 0u;    // a padded stage stores nothing
#pragma unroll
          for (int i = 0; i < ITEMS; i++) {
            const uint32_t s0 = selector4<NA>(ring[d][h][i].x), s1 = selector4<NA>(ring[d][h][i].y),
                           s2 = selector4<NA>(ring[d][h][i].z), s3 = selector4<NA>(ring[d][h][i].w);
            const uint32_t rb = sb + sa[i], m = lm[i] & live;
#pragma unroll
            for (int v = 0; v < 4; v++)
              sts128_if(m & (1u << v), rb + v * 128 + __byte_perm(sw_lo[i], 0u, 0x4440 + v), onehot4(s0, 1u << (8 * v), 0u),
                        onehot4(s1, 1u << (8 * v), 0u), onehot4(s2, 1u << (8 * v), 0u), onehot4(s3, 1u << (8 * v), 0u));
            if (WIDE) {                                    // factors with 5-8 levels
#pragma unroll
              for (int v = 0; v < 4; v++)
                sts128_if(m & (16u << v), rb + (v + 4) * 128 + __byte_perm(sw_hi[i], 0u, 0x4440 + v),
                          onehot4(s0, 0u, 1u << (8 * v)), onehot4(s1, 0u, 1u << (8 * v)), onehot4(s2, 0u, 1u << (8 * v)),
                          onehot4(s3, 0u, 1u << (8 * v)));
            }
          }
        }
        // generic-proxy writes -> visible to the tensor core's async-proxy reads
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        asm volatile("{ .reg .pred p; setp.eq.u32 p, %1, 0; @p mbarrier.arrive.shared::cta.b64 _, [%0]; }"
                     ::"r"(smem_u32(&full_bar[slot])), "r"((uint32_t)lane) : "memory");
#pragma unroll
        for (int h = 0; h < SUP; h++)                     // past the end: the last stage again, never used
          load(ring[d][h], min((su + DEPTH) * SUP + h, nst - 1));
      }
    }
  } else {
    // ---------------- MMA issuer ----------------
    // The whole warp walks the stage loop and one lane issues.
    // instruction descriptor: D = s32 (c_format 2), A and B unsigned 8-bit, both K-major, N >> 3, M >> 4
    const uint32_t idesc = (2u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(PAIR_BM >> 4) << 24);
    const uint32_t a_shift = a_in_b ? (uint32_t)(A_BYTES + (a0 - b0) * 128) : 0u;
    const uint32_t smem0 = smem_u32(smem);
    for (int su = 0; su < npad; su++) {
      const int slot = su & (NSS - 1);
      mbar_wait_spin(&full_bar[slot], (su / NSS) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (lane == 0) {
#pragma unroll
        for (int h = 0; h < SUP; h++) {
          const int it = su * SUP + h;
          const uint32_t sb = smem0 + (uint32_t)(slot * SUP + h) * STAGE_BYTES;
          const uint64_t ad = make_desc_sw128(sb + a_shift), bd = make_desc_sw128(sb + A_BYTES);
          if (it < nst) {
#pragma unroll
            for (int k = 0; k < PAIR_KROWS / 32; k++)    // +32 bytes of K per step: +2 in the (address >> 4) field
              umma_i8(tmem, ad + 2 * k, bd + 2 * k, idesc, (it | k) != 0);
          }
        }
        umma_commit(&empty_bar[slot]);
      }
      __syncwarp();
    }
    if (lane == 0) umma_commit(&accum_bar);
    __syncwarp();
  }

  // ---------------- accumulator -> G ----------------
  // A thread holds one indicator row of the accumulator (its TMEM lane), 32 columns at
  // a time; the 32 x 32 block is turned through shared memory (the stage buffers are
  // free once the last MMA has committed) so that a warp writes 128 contiguous bytes
  // of a row of G per instruction.
  if (warp < 8) {
    mbar_wait(&accum_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int lg = warp & 3;                          // a warp reads the 32 TMEM lanes of its quarter
    int *tile_s = reinterpret_cast<int *>(smem) + warp * (32 * 33);
    for (int c0 = (warp >> 2) * 32; c0 < BN; c0 += 64) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 32; j++) tile_s[lane * 33 + j] = (int)v[j];
      __syncwarp();
      int *dst = G + (size_t)(a0 + lg * 32) * ldg + b0 + c0 + lane;
      const int rows = min(32, L - (a0 + lg * 32));
      for (int rr = 0; rr < rows; rr++) {
        const int val = tile_s[rr * 33 + lane];
        if (atomic_out) { if (val) atomicAdd(dst + (size_t)rr * ldg, val); }
        else dst[(size_t)rr * ldg] = val;
      }
      __syncwarp();
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == NPW) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(BN));
  }
}

// Unit tables out of G, in the layout bnl_counts returns (n[z][x][y], first
// variable fastest). A two-column unit with codes (a, b) = (col[0], col[1]) reads
// G[first(col0) + a][first(col1) + b] (upper triangle: smaller indicator first); a
// one-column unit reads the diagonal. A column without NA has no indicator for its
// last level (slot_eff < levels): those cells are the margin minus the rest of the
// row / column -- the other column's one-way count G[b][b] when this column can
// hold no NA, the row count N when neither can.
__global__ void __launch_bounds__(256)
pair_extract_kernel(const BnlUnit *__restrict__ units, const int *__restrict__ ids, int n,
                    const int *__restrict__ slot_first, const int *__restrict__ slot_eff, int nobs,
                    const int *__restrict__ G, int ldg, int *__restrict__ tables) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const BnlUnit &u = units[ids ? ids[t] : t];
  int *tab = tables + u.table_off;
  const int f0 = slot_first[u.col[0]], e0 = slot_eff[u.col[0]];
  auto diag = [&](int i) { return G[(size_t)i * ldg + i]; };
  if (u.ncols == 1) {
    const int l0 = u.cells;
    int sum = 0;
    for (int a = 0; a < e0; a++) { const int v = diag(f0 + a); tab[a] = v; sum += v; }
    if (e0 < l0) tab[l0 - 1] = nobs - sum;
    return;
  }
  const int f1 = slot_first[u.col[1]], e1 = slot_eff[u.col[1]], l0 = u.mul[1], l1 = u.cells / l0;
  int total = 0;
  for (int b = 0; b < e1; b++) {
    int colsum = 0;
    for (int a = 0; a < e0; a++) {
      const int ia = f0 + a, ib = f1 + b;
      const int v = ia <= ib ? G[(size_t)ia * ldg + ib] : G[(size_t)ib * ldg + ia];
      tab[a + l0 * b] = v;
      colsum += v;
    }
    if (e0 < l0) { const int v = diag(f1 + b) - colsum; tab[l0 - 1 + l0 * b] = v; colsum += v; }
    total += colsum;
  }
  if (e1 < l1) {
    for (int a = 0; a < e0; a++) {
      int rowsum = 0;
      for (int b = 0; b < e1; b++) rowsum += tab[a + l0 * b];
      const int v = diag(f0 + a) - rowsum;
      tab[a + l0 * (l1 - 1)] = v;
      total += v;
    }
    if (e0 < l0) tab[l0 - 1 + l0 * (l1 - 1)] = nobs - total;
  }
}
