// Dense-input forward on the tensor cores: feat (B, C, A) fp32 -> q (B, A), 64-channel networks.
//
// pk_dense_forward_kernel (pk_decide.cuh) walks the dense block row by row on the FP32 pipe: ~5 800 warp instructions
// per battle, 0.39 of the HBM roofline.  The block is dense in memory but the contraction does not care: layer 0 is
// the GEMM  H0[(b, a)][j] = sum_k X[b][k][a] W0[j][k],  and this kernel runs it -- and the hidden layers -- with
// tcgen05.mma.kind::tf32 on operands split x = hi + lo (hi = tf32(x), lo = x - hi exact), accumulating
// hi*hi + hi*lo + lo*hi in fp32 in TMEM, which keeps Q inside the 1e-5 fp32 tolerance of the parity tests.
//
//   tile       nb = 128 / A consecutive battles = nb * A accumulator rows (126 at A = 9); row r = (battle r / A,
//              action column r % A), which is also the flat index of q -- the result store is one coalesced run.
//   pairs      a CTA works on TWO tiles at a time against each W0 k-block: two accumulators, one A stage per tile, the
//              W0 stream (as many bytes per tile as the features themselves) shared between them.
//   X stream   a k-block is 64 feature rows: per battle 256 A contiguous bytes at a 4-byte-aligned address.  A lane of
//              one of the two X producer warps copies the 16-byte-aligned cover of one battle's piece (2.3 KB) with a TMA
//              bulk load into a ring of four X stages; completion is counted in bytes on an mbarrier.  Nothing is read
//              twice.
//   A operand  lives in TENSOR MEMORY, not shared memory (tcgen05.mma with A from TMEM): accumulator row r is TMEM
//              lane r, and thread (lane quarter q, lane t) owns row 32 q + t.  Warps 0-7 feed tile 0 of the pair, warps
//              8-15 tile 1, so the two sets drift apart and one set's shared-memory reads overlap the other's TMEM
//              stores.  Within a set warp (q, half) reads X[row][32 half .. 32 half + 31] -- thirty-two LDS whose lanes
//              walk consecutive words -- splits them and stores the hi parts (and the lo parts when there are any:
//              one-hot rows are exact in tf32, and a group of lo columns that already holds zeros is not stored
//              again) with tcgen05.st.  No transposed copy in shared memory and no proxy fence.
//   W0         per k-block a 128-row B operand [tf32(W0) ; W0 - tf32(W0)] (32 KB, PkLayout::tc0, kept current by
//              pk_clamp_adam), streamed from L2 with one TMA bulk load per k-block and pair: MMA N = 128 gives X_hi W_hi
//              in accumulator columns 0-63 and X_hi W_lo in 64-127; the N = 64 lo MMA adds X_lo W_hi to columns 0-63.
//              The hidden matrices (PkLayout::tc) ride the same ring during the epilogue.
//   MMA warp   one thread issues; tcgen05.commit releases the stages.
//   epilogue   all sixteen warps, one tile at a time: thread (row, cg) reads channels 16 cg .. +15 of its row from the
//              accumulator, adds the two halves and the bias, applies ReLU, splits and writes the hidden layer's A
//              operand -- also in TMEM; the head is a 16-channel partial dot product per thread, summed in a fixed
//              order through shared memory.
// Tensor memory (512 columns): two layer-0 accumulators of 128, A stages 2 x 128 (the hidden layers' A operand and
// accumulator alias them during the epilogue).
//
// Measured on B200, 65 536 states: 380 - 396 us = 0.56 - 0.59 of the HBM roofline (FP32-pipe kernel:
// 569 us, 0.39); 4 096 states 31 - 33 us (54).  What was learned on the way, each from a timeline of the roles
// (tools/trace_dense.py) or a microbenchmark (tools/micro/bulk_load_bw.cu):
//   * A tiles in SHARED memory (transposed by LDS + STS): 136 KB of shared-memory traffic per 16 KB of input -- the
//     shared-memory pipe bound the kernel at 700 us.
//   * a TMA bulk load occupies the SM's TMA unit for ~75 cycles whatever its size (592 B .. 2.4 KB: 2.2 / 4.4 / 9.3 TB/s
//     chip-wide from L2); 32-row pieces (1.2 KB) capped the stream and every issuing thread waited its turn (840 us);
//     16-byte cp.async, a thousand requests per k-block, stalled the issuing warps (960 us).  64-row pieces: 414 us.
//     One warp issues its lanes' bulk loads ~115 cycles apart; two producer warps reach the unit's 75.
//   * a tcgen05.mma whose A operand is in TMEM issues every ~131 - 150 cycles whatever its N (64 or 128), whether
//     successive MMAs share an accumulator or alternate between two; two issuing warps are slower (~200).  Eight per
//     tile and k-block = ~1 200 cycles: with the X producer (~1 400 per k-block with its waits) this is what bounds
//     the kernel now.  Without the MMAs the kernel takes 320 us, without MMAs and TMEM stores 311 us.
//   * a CTA-pair form (clusters of two CTAs, each with its own tiles, the leader issuing tcgen05.mma.cta_group::2 with A
//     from both CTAs' TMEM so that one instruction covers 256 rows, each CTA staging half of every weight operand) was
//     built and passed the parity tests, but ran at 589 us against 419 on the same box: with one A stage per tile the
//     loop workers -> a_full -> MMA -> a_empty -> workers has no slack, and the remote arrivals and multicast commits
//     put two cluster-scope latencies into it, with both CTAs forced into lock step.  Not kept.  (The same pairing is
//     what lifted the wide GEMM from 83 to 93 % tensor-active: there the ring is five stages deep.)
//   * next: X through tensor maps (four boxes per k-block instead of fourteen bulk loads) and layer 0 as bf16 x 3-way
//     split with K = 16 per MMA (half the MMAs); neither is built.
// Reference semantics: MLPModel.forward mlp_model.py:21-30 on Agent._calc_q_vector[_batch] input (agent.py:23-30).
#pragma once

#include "pk_decide_tc.cuh"

#define PKD_TW 16                   // loading / splitting / epilogue warps
#define PKD_PRODUCER PKD_TW         // warp index of the TMA producer of the W0 stream (and the one-off weights)
#define PKD_XPRODUCER (PKD_TW + 1)  // first warp of the TMA producers of the X stream
#define PKD_XPW 2                   // X producer warps: lane i of warp j copies the piece of battle PKD_XPW i + j (a warp issues
                                    // its lanes' bulk loads one after the other, ~115 cycles each: two warps keep the TMA unit fed;
                                    // three measured 391 against 381 us at 65 536 states)
#define PKD_MMA (PKD_TW + 1 + PKD_XPW)  // warp index of the MMA issuer
#define PKD_WARPS (PKD_TW + 2 + PKD_XPW)
#define PKD_KBF 64                  // feature rows per k-block: a TMA bulk load costs ~75 cycles of the SM's TMA unit
                                    // whatever its size (tools/micro/bulk_load_bw.cu), so pieces must be >= 2 KB
#define PKD_XK 1                    // MMA k-blocks per X stage.  2 (one 4.6 KB bulk load per battle and 128 feature rows: half the
                                    // loads for the TMA unit, but only two X stages fit) passes parity and is SLOWER: 396 against
                                    // 381-385 us at 65 536 states -- the X side is not bound by the number of loads
#define PKD_XKF (PKD_KBF * PKD_XK)
#define PKD_A_STAGES 2
#define PKD_W_TILE 16384u           // [W0 hi ; W0 lo] of 32 input rows: 128 rows x 128 B (one PkLayout::tc0 tile)
#define PKD_W_STAGE 32768u          // one k-block: two tiles
#define PKD_W_STAGES 2
#define PKD_X_STAGES_MAX 4
#define PKD_TMEM_COLS 512u
#define PKD_COL_ACC0 0u             // layer-0 accumulators of the pair's two tiles, 128 columns each:
                                    // X_hi W_hi + X_lo W_hi | X_hi W_lo
#define PKD_COL_A 256u              // A stages (one per tile of the pair), 128 columns each: hi parts (64, column = k), lo parts (64)
#define PKD_COL_HA PKD_COL_A        // hidden A operand (per K-step of 8: hi (8), lo (8)): aliases A stage 0, idle during the epilogue
#define PKD_COL_HACC (PKD_COL_A + 128u)  // hidden-layer accumulator (H_hi W_hi + H_lo W_hi | H_hi W_lo): aliases A stage 1

struct PkdSmemPlan {
  unsigned w;         // W stages: W0 k-blocks, and the hidden matrices (PkLayout::tc, one stage each) during the epilogue
  unsigned x;         // X stages
  unsigned x_stage, x_stages, piece;  // bytes per stage, stages, bytes per battle slot
  unsigned small;     // b[0..n_layers-1], wout, bout
  unsigned part;      // head partials [4][128]
  unsigned flags;     // "some lo is non-zero" per (A stage, phase)
  unsigned bars;
  unsigned total;
};

// battles per 128-row tile (warp w copies battles w, w + 16: two passes at most)
__host__ __device__ inline int pkd_battles_per_tile(int A) { return 128 / A < 32 ? 128 / A : 32; }

__host__ __device__ inline PkdSmemPlan pkd_smem_plan(const PkLayout& lay, int A, unsigned max_smem) {
  PkdSmemPlan p;
  const unsigned nb = (unsigned)pkd_battles_per_tile(A);
  unsigned o = 0;
  p.w = o;
  o += PKD_W_STAGE * PKD_W_STAGES;
  p.small = o;
  o += ((unsigned)lay.n_layers * PKT_H + PKT_H + 4u) * 4u;
  p.part = o;
  o += 4u * 128u * 4u;
  p.flags = o;
  o += 32u;
  p.bars = o;
  o += 8u * (2u * PKD_X_STAGES_MAX + 2u * PKD_A_STAGES + 2u * PKD_W_STAGES + 4u) + 16u;
  o = (o + 127u) & ~127u;
  p.x = o;
  p.piece = 4u * PKD_XKF * (unsigned)A + 32u;  // the aligned cover of an X block's bytes at any 4-byte alignment
  p.x_stage = (nb * p.piece + 127u) & ~127u;
  const unsigned room = max_smem > o + 1024u ? max_smem - o - 1024u : 0u;
  p.x_stages = room / p.x_stage;
  if (p.x_stages > PKD_X_STAGES_MAX) p.x_stages = PKD_X_STAGES_MAX;
  p.total = o + p.x_stages * p.x_stage + 1024u;
  return p;
}

// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void pkd_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void pkd_tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void pkd_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float pkd_lds(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
// D[tmem] (+)= A[tmem] * B[smem desc], kind::tf32
__device__ __forceinline__ void pkd_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}

// sixteen values of one row -> two K-steps of an A operand in TMEM: per K-step hi (8 columns) then lo (8 columns)
// Returns the OR of the lo bit patterns.
__device__ __forceinline__ uint32_t pkd_split_store16(uint32_t taddr, const float (&x)[16]) {
  uint32_t lo_or = 0u;
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    uint32_t v[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float hi = pk_tf32_hi(x[8 * s + j]), lo = x[8 * s + j] - hi;
      v[j] = __float_as_uint(hi);
      v[8 + j] = __float_as_uint(lo);
      lo_or |= __float_as_uint(lo);
    }
    pkd_tmem_st16(taddr + 16u * (uint32_t)s, v);
  }
  return lo_or;
}

// debug timeline (pk_debug_set_trace): SM clock of CTA 0 at point p of iteration it < 128 of role 0 (warp 0),
// 1 (MMA thread), 2 (producer lane 0); trace == nullptr in normal operation
__device__ __forceinline__ void pkd_trace(unsigned long long* tr, int role, uint32_t it, int p) {
  if (tr != nullptr && it < 128u) tr[((size_t)role * 128 + it) * 8 + p] = (unsigned long long)clock64();
}

__global__ void __launch_bounds__(PKD_WARPS * 32, 1)
pk_dense_tc_kernel(const float* __restrict__ feat, int B, int A, const float* __restrict__ prep, PkLayout lay,
                   float* __restrict__ q, unsigned max_smem, unsigned long long* __restrict__ trace) {
  constexpr int HP = PKT_H;
  extern __shared__ __align__(1024) unsigned char smem[];
  if ((pk_smem_u32(smem) & 1023u) != 0u) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const PkdSmemPlan plan = pkd_smem_plan(lay, A, max_smem);
  const int XS = (int)plan.x_stages;
  const int nb = pkd_battles_per_tile(A), rows = nb * A;
  const int L = lay.C * A;
  const int NKB = (lay.C + PKD_KBF - 1) / PKD_KBF;
  const int NXB = (NKB + PKD_XK - 1) / PKD_XK;  // X blocks per tile
  const int ntiles = (B + nb - 1) / nb;
  const int n_hidden = lay.n_layers - 1;  // >= 1
  const int npairs = (ntiles + 1) >> 1;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + plan.bars);
  uint64_t* x_full = bars;
  uint64_t* x_empty = x_full + PKD_X_STAGES_MAX;
  uint64_t* a_full = x_empty + PKD_X_STAGES_MAX;
  uint64_t* a_empty = a_full + PKD_A_STAGES;
  uint64_t* w_full = a_empty + PKD_A_STAGES;
  uint64_t* w_empty = w_full + PKD_W_STAGES;
  uint64_t* acc0_full = w_empty + PKD_W_STAGES;
  uint64_t* hid_full = acc0_full + 1;
  uint64_t* hidacc_full = hid_full + 1;
  uint64_t* wbar = hidacc_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + 1);
  volatile uint32_t* lo_flag = reinterpret_cast<volatile uint32_t*>(smem + plan.flags);  // [A stage][phase]
  float* s_small = reinterpret_cast<float*>(smem + plan.small);
  float* s_part = reinterpret_cast<float*>(smem + plan.part);
  const uint32_t w_u32 = pk_smem_u32(smem + plan.w);
  const uint32_t x_u32 = pk_smem_u32(smem + plan.x);

  if (threadIdx.x == 0) {
    for (int i = 0; i < PKD_X_STAGES_MAX; ++i) {
      pk_mbar_init(x_full + i, PKD_XPW);  // each X producer warp's arrive.expect_tx with the byte count of its copies
      pk_mbar_init(x_empty + i, PKD_TW / 2);  // the eight warps of the tile's set
    }
    for (int i = 0; i < PKD_A_STAGES; ++i) {
      pk_mbar_init(a_full + i, PKD_TW / 2);
      pk_mbar_init(a_empty + i, 1);
    }
    for (int i = 0; i < PKD_W_STAGES; ++i) {
      pk_mbar_init(w_full + i, 1);
      pk_mbar_init(w_empty + i, 1);
    }
    pk_mbar_init(acc0_full, 1);
    pk_mbar_init(hid_full, PKD_TW);
    pk_mbar_init(hidacc_full, 1);
    pk_mbar_init(wbar, 1);
    for (int i = 0; i < 2 * PKD_A_STAGES; ++i) lo_flag[i] = 0u;
    pk_fence_mbar_init();
  }
  if (warp == 0) pkw_tmem_alloc(tmem_slot, PKD_TMEM_COLS);
  pkw_tc_fence_before();
  __syncthreads();
  pkw_tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == PKD_PRODUCER) {
    // ---- TMA producer: biases / head / hidden operands once, then the W0 k-blocks in consumption order ----
    if (lane == 0) {
      pk_mbar_expect_tx(wbar, (uint32_t)(lay.n_layers * HP + HP + 4) * 4u);
      for (int l = 0; l < lay.n_layers; ++l) pk_bulk_load(s_small + HP * l, prep + lay.b[l], HP * 4u, wbar);
      pk_bulk_load(s_small + HP * lay.n_layers, prep + lay.wout, HP * 4u, wbar);
      pk_bulk_load(s_small + HP * lay.n_layers + HP, prep + lay.bout, 16u, wbar);
      uint32_t wit = 0;
      unsigned long long* ptr = blockIdx.x == 0 ? trace : nullptr;
      for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
        for (int kb = 0; kb < NKB; ++kb, ++wit) {
          const int ws = (int)(wit % PKD_W_STAGES);
          const uint32_t wph = (wit / PKD_W_STAGES) & 1u;
          pkd_trace(ptr, 2, wit, 0);
          pk_mbar_wait(w_empty + ws, wph ^ 1u);
          pkd_trace(ptr, 2, wit, 1);
          pk_mbar_expect_tx(w_full + ws, PKD_W_STAGE);
          pk_bulk_load(smem + plan.w + (unsigned)ws * PKD_W_STAGE, prep + lay.tc0 + (size_t)kb * (PKD_W_STAGE >> 2),
                       PKD_W_STAGE, w_full + ws);
        }
        // the hidden matrices ride the same ring during the epilogue (one stage each; a single hidden layer's matrix
        // is loaded once for the pair's two tiles)
        for (int tl = 0; tl < 2 * n_hidden; ++tl) {
          if (n_hidden == 1 && tl == 1) break;
          const int ws = (int)(wit % PKD_W_STAGES);
          pk_mbar_wait(w_empty + ws, ((wit / PKD_W_STAGES) & 1u) ^ 1u);
          pk_mbar_expect_tx(w_full + ws, PKT_SET_BYTES);
          // PkLayout::tc holds [hi kb0][hi kb1][lo kb0][lo kb1]; the stage gets [hi kb0][lo kb0][hi kb1][lo kb1], so that
          // each k-block is one 128-row operand [W_hi ; W_lo] as for layer 0
          const float* src = prep + lay.tc[1 + tl % n_hidden];
          unsigned char* dst = smem + plan.w + (unsigned)ws * PKD_W_STAGE;
          for (int kbh = 0; kbh < 2; ++kbh)
            for (int part = 0; part < 2; ++part)
              pk_bulk_load(dst + (unsigned)(2 * kbh + part) * PKT_KB_BYTES,
                           src + (size_t)part * PK_TC_TILE_FLOATS + (size_t)kbh * (PKT_KB_BYTES >> 2), PKT_KB_BYTES, w_full + ws);
          ++wit;
        }
      }
    }
    __syncwarp();
  } else if (warp >= PKD_XPRODUCER && warp < PKD_XPRODUCER + PKD_XPW) {
    // ---- TMA producers of the X stream: a lane copies the 16-byte-aligned cover of one battle's k-block ----
    const int pb = PKD_XPW * lane + (warp - PKD_XPRODUCER);  // this lane's battle within a tile
    const unsigned long long feat_addr = (unsigned long long)feat;
    const unsigned long long arr_end16 = (feat_addr + (unsigned long long)B * L * 4ull) & ~15ull;
    uint32_t xit = 0;
    unsigned long long* ptr = (blockIdx.x == 0 && lane == 0 && warp == PKD_XPRODUCER) ? trace : nullptr;
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
      for (int kt = 0; kt < 2 * NXB; ++kt, ++xit) {
        const int kb = kt >> 1, tile = 2 * pair + (kt & 1);  // kb: the X block
        const long long b = (long long)tile * nb + pb;
        const bool mine = pb < nb && tile < ntiles && b < B;
        const int xs = (int)(xit % (uint32_t)XS);
        pkd_trace(ptr, 2, xit, 2);
        pk_mbar_wait(x_empty + xs, ((xit / (uint32_t)XS) & 1u) ^ 1u);
        pkd_trace(ptr, 2, xit, 3);
        uint32_t bytes = 0;
        unsigned long long src = 0;
        unsigned char* dst = smem + plan.x + (unsigned)xs * plan.x_stage + (unsigned)pb * plan.piece;
        if (mine) {
          const unsigned long long addr =
              feat_addr + 4ull * ((unsigned long long)b * L + (unsigned long long)kb * PKD_XKF * A);
          const uint32_t mis = (uint32_t)(addr & 15ull);
          const uint32_t len = (uint32_t)min(PKD_XKF, lay.C - kb * PKD_XKF) * (uint32_t)A * 4u;
          src = addr - mis;
          bytes = (mis + len + 15u) & ~15u;
          if (src + bytes > arr_end16) {
            // the cover would pass the end of the array: bulk-copy the whole 16-byte chunks, the rest by hand
            const uint32_t whole = arr_end16 > src ? (uint32_t)(arr_end16 - src) : 0u;
            for (uint32_t o = whole > mis ? whole : mis; o < mis + len; o += 4u)
              *reinterpret_cast<float*>(dst + o) = __ldg(reinterpret_cast<const float*>(src + o));
            bytes = whole;
          }
        }
        uint32_t total = bytes;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) total += __shfl_xor_sync(0xffffffffu, total, d);
        __syncwarp();  // the hand-copied floats above are ordered before lane 0's arrival
        if (lane == 0) pk_mbar_expect_tx(x_full + xs, total);
        __syncwarp();
        if (bytes) pk_bulk_load(dst, reinterpret_cast<const void*>(src), bytes, x_full + xs);
        pkd_trace(ptr, 2, xit, 4);
      }
    }
  } else if (warp == PKD_MMA) {
    // ---- MMA issuer ----
    // Measured: a tcgen05.mma whose A operand is in tensor memory issues every ~131 cycles whatever its N (64 or 128),
    // whether successive MMAs share an accumulator or alternate between two, and two issuing warps are slower (~200).
    if (lane == 0) {
      constexpr uint32_t idesc_hi = pkt_instr_desc(128, 128), idesc_lo = pkt_instr_desc(128, 64);
      uint32_t ait = 0, wit = 0, hfpar = 0;
      unsigned long long* mtr = blockIdx.x == 0 ? trace : nullptr;
      pk_mbar_wait(wbar, 0);
      for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
        for (int kt = 0; kt < 2 * NKB; ++kt, ++ait) {
          const int kb = kt >> 1, as = kt & 1;  // the tile of the pair = its A stage = its accumulator
          const uint32_t aph = (ait >> 1) & 1u;
          const int ws = (int)(wit % PKD_W_STAGES);
          pkd_trace(mtr, 1, ait, 0);
          pk_mbar_wait(a_full + as, aph);
          pkd_trace(mtr, 1, ait, 1);
          if (as == 0) pk_mbar_wait(w_full + ws, (wit / PKD_W_STAGES) & 1u);
          pkd_trace(mtr, 1, ait, 2);
          pkw_tc_fence_after();
          const uint32_t fl = lo_flag[as * 2 + (int)aph];
          lo_flag[as * 2 + (int)aph] = 0u;  // set again two uses of this stage later, after the commit below
          const uint32_t ta = tmem_base + PKD_COL_A + 128u * (uint32_t)as, tw = w_u32 + (uint32_t)ws * PKD_W_STAGE;
          const uint32_t tacc = tmem_base + PKD_COL_ACC0 + 128u * (uint32_t)as;
#pragma unroll
          for (int s = 0; s < PKD_KBF / 8; ++s) {  // UMMA_K = 8 fp32: 8 TMEM columns of A (hi; lo 64 further), 32 B of the swizzle atom of W
            const uint64_t bd = pkw_smem_desc_k128(tw + (uint32_t)(s >> 2) * PKD_W_TILE) + 2u * (uint64_t)(s & 3);
            pkd_mma_ts(tacc, ta + 8u * (uint32_t)s, bd, idesc_hi, (kb | s) != 0 ? 1u : 0u);
            if (fl) pkd_mma_ts(tacc, ta + 64u + 8u * (uint32_t)s, bd, idesc_lo, 1u);
          }
          pkd_trace(mtr, 1, ait, 3);
          if (as == 1) {
            pkw_mma_commit(w_empty + ws);
            ++wit;
          }
          pkw_mma_commit(a_empty + as);
          pkd_trace(mtr, 1, ait, 4);
        }
        pkw_mma_commit(acc0_full);
        for (int tl = 0; tl < 2 * n_hidden; ++tl) {  // the pair's two tiles in turn
          const int l = 1 + tl % n_hidden;
          pk_mbar_wait(hid_full, hfpar);
          hfpar ^= 1u;
          pkw_tc_fence_after();
          // D (128 x 128) = H_hi . [W_hi ; W_lo]^T, D[:, 0:64] += H_lo . W_hi^T  (H in TMEM)
          constexpr uint32_t idesc_h2 = pkt_instr_desc(128, 128), idesc_h = pkt_instr_desc(128, 64);
          const int ws = (int)(wit % PKD_W_STAGES);
          if (n_hidden > 1 || tl == 0) pk_mbar_wait(w_full + ws, (wit / PKD_W_STAGES) & 1u);
          const uint32_t tb = w_u32 + (uint32_t)ws * PKD_W_STAGE;
#pragma unroll
          for (int s = 0; s < 8; ++s) {
            const uint32_t th = tmem_base + PKD_COL_HA + 16u * (uint32_t)s;
            const uint64_t bd = pkw_smem_desc_k128(tb + (uint32_t)(s >> 2) * PKD_W_TILE) + 2u * (uint64_t)(s & 3);
            pkd_mma_ts(tmem_base + PKD_COL_HACC, th, bd, idesc_h2, s != 0 ? 1u : 0u);
            pkd_mma_ts(tmem_base + PKD_COL_HACC, th + 8u, bd, idesc_h, 1u);
          }
          pkw_mma_commit(hidacc_full);
          if (n_hidden > 1 || tl == 1) {
            pkw_mma_commit(w_empty + ws);
            ++wit;
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ---- loading / splitting / epilogue warps ----
    // Main loop: warps 0-7 feed tile 0 of the pair, warps 8-15 tile 1 (each its own A stage and every other X stage), so
    // the two sets drift apart and one set's shared-memory reads overlap the other's tensor-memory stores.  Within a set
    // warp (q, half) owns TMEM lane quarter q and the k-block's rows 32 half .. 32 half + 31.
    const int tset = warp >> 3, wq = warp & 3, khalf = (warp >> 2) & 1;
    const int cg = warp >> 2;                 // epilogue: this warp's 16-channel group (all sixteen warps, one tile at a time)
    const int row = 32 * wq + lane;           // this thread's accumulator row = TMEM lane
    const bool row_ok = row < rows;
    const int rbi = row / A;                  // its battle within the tile, its action column
    const int ra = row - rbi * A;
    const unsigned long long feat_addr = (unsigned long long)feat;
    const uint32_t tlane = tmem_base + ((uint32_t)(32 * wq) << 16);
    // element (k, ra) of this row's battle inside a stage: piece rbi, float index k * A + ra (+ the piece's misalignment)
    const uint32_t src_row = (uint32_t)rbi * plan.piece + (uint32_t)ra * 4u + (uint32_t)((PKD_KBF / 2) * khalf * A) * 4u;
    uint32_t tpar = 0, hpar = 0;
    bool first = true;
    unsigned long long* ttr = (blockIdx.x == 0 && threadIdx.x == 0) ? trace : nullptr;
    uint32_t sit = 0;  // this set's iteration count: (pair, k-block)
    uint32_t xsit = 0;  // X blocks this set has consumed
    bool lo_clean[2] = {false, false};  // this warp's two lo groups of its A stage hold zeros
    for (int pair = blockIdx.x; pair < npairs; pair += gridDim.x) {
      const int tile0 = 2 * pair;
      const uint32_t mis = (uint32_t)(
          (feat_addr + 4ull * (unsigned long long)((long long)(tile0 + tset) * nb + rbi) * (unsigned long long)L) & 15ull);
      for (int kb = 0; kb < NKB; ++kb, ++sit) {
        const uint32_t xit = 2u * xsit + (uint32_t)tset;  // X stages alternate between the pair's tiles
        const int xs = (int)(xit % (uint32_t)XS);
        const int sub = kb % PKD_XK;  // which 64 rows of the X block
        const uint32_t aph = sit & 1u;
        pkd_trace(ttr, 0, sit, 0);
        if (sub == 0) pk_mbar_wait(x_full + xs, (xit / (uint32_t)XS) & 1u);
        pkd_trace(ttr, 0, sit, 1);
        const uint32_t xb = x_u32 + (uint32_t)xs * plan.x_stage + src_row + mis + (uint32_t)(sub * PKD_KBF * A) * 4u;
        const int k0 = kb * PKD_KBF + (PKD_KBF / 2) * khalf;
        float x[PKD_KBF / 2];
#pragma unroll
        for (int j = 0; j < PKD_KBF / 2; ++j) x[j] = (row_ok && k0 + j < lay.C) ? pkd_lds(xb + (uint32_t)(j * A) * 4u) : 0.0f;
        pkd_trace(ttr, 0, sit, 2);
        __syncwarp();
        if (sub == PKD_XK - 1 || kb == NKB - 1) {  // the block's last rows are in registers: the stage may be refilled
          if (lane == 0) pkw_mbar_arrive(x_empty + xs);
          ++xsit;
        }
        pk_mbar_wait(a_empty + tset, aph ^ 1u);
        pkw_tc_fence_after();
        pkd_trace(ttr, 0, sit, 3);
        static_assert(PKD_KBF == 64, "thirty-two values per thread: four K-steps");
        // A stage: columns 0-63 the hi parts (column = k), 64-127 the lo parts.  One-hot rows have no lo part: a warp
        // whose sixteen-column group of lo is zero and already holds zeros ("clean") does not store it again, which
        // halves the tensor-memory stores of all but the first k-block.
        bool any_lo = false;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t vh[16], vl[16];
          uint32_t lo_or = 0u;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float hi = pk_tf32_hi(x[16 * hh + j]), lo = x[16 * hh + j] - hi;
            vh[j] = __float_as_uint(hi);
            vl[j] = __float_as_uint(lo);
            lo_or |= __float_as_uint(lo);
          }
          const uint32_t tcol = tlane + PKD_COL_A + 128u * (uint32_t)tset + 32u * (uint32_t)khalf + 16u * (uint32_t)hh;
          pkd_tmem_st16(tcol, vh);
          const bool nz = __any_sync(0xffffffffu, (lo_or << 1) != 0u);
          if (nz || !lo_clean[hh]) pkd_tmem_st16(tcol + 64u, vl);  // values, or the zeros that make the group clean
          lo_clean[hh] = !nz;
          any_lo |= nz;
        }
        pkd_tmem_st_wait();
        pkw_tc_fence_before();
        __syncwarp();
        pkd_trace(ttr, 0, sit, 4);
        if (lane == 0) {
          if (any_lo) lo_flag[tset * 2 + (int)aph] = 1u;
          pkw_mbar_arrive(a_full + tset);
        }
        pkd_trace(ttr, 0, sit, 5);
      }
      const uint32_t xit = sit;
      // ---- epilogue of the pair: relu(D[:, j] + D[:, 64 + j] + b0[j]) -> the hidden layer's A operand (TMEM) ----
      pkd_trace(ttr, 0, xit - 1u, 6);
      pk_mbar_wait(acc0_full, tpar);
      tpar ^= 1u;
      pkw_tc_fence_after();
      if (first) {
        pk_mbar_wait(wbar, 0);
        first = false;
      }
      for (int t = 0; t < 2; ++t) {
      const long long b0 = (long long)(tile0 + t) * nb;
      {
        uint32_t v1[16], v2[16];
        pkd_tmem_ld16(tlane + PKD_COL_ACC0 + 128u * (uint32_t)t + 16u * (uint32_t)cg, v1);
        pkd_tmem_ld16(tlane + PKD_COL_ACC0 + 128u * (uint32_t)t + 64u + 16u * (uint32_t)cg, v2);
        pkw_tmem_ld_wait();
        float h[16];
#pragma unroll
        for (int j = 0; j < 16; ++j)
          h[j] = fmaxf((__uint_as_float(v1[j]) + __uint_as_float(v2[j])) + s_small[16 * cg + j], 0.0f);
        pkd_split_store16(tlane + PKD_COL_HA + 32u * (uint32_t)cg, h);
      }
      pkd_tmem_st_wait();
      pkw_tc_fence_before();
      __syncwarp();
      if (lane == 0) pkw_mbar_arrive(hid_full);
      for (int l = 1; l <= n_hidden; ++l) {
        pk_mbar_wait(hidacc_full, hpar);
        hpar ^= 1u;
        pkw_tc_fence_after();
        uint32_t v[16];
        {
          uint32_t v2[16];
          pkd_tmem_ld16(tlane + PKD_COL_HACC + 16u * (uint32_t)cg, v);
          pkd_tmem_ld16(tlane + PKD_COL_HACC + 64u + 16u * (uint32_t)cg, v2);
          pkw_tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
        }
        const float* bl = s_small + HP * l + 16 * cg;
        if (l < n_hidden) {
          float h[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) h[j] = fmaxf(__uint_as_float(v[j]) + bl[j], 0.0f);
          pkd_split_store16(tlane + PKD_COL_HA + 32u * (uint32_t)cg, h);
          pkd_tmem_st_wait();
          pkw_tc_fence_before();
          __syncwarp();
          if (lane == 0) pkw_mbar_arrive(hid_full);
        } else {
          // head: this thread's 16 channels of q = w_out . relu(D + b) + b_out
          const float* wo = s_small + HP * lay.n_layers + 16 * cg;
          float p0 = 0.0f, p1 = 0.0f, p2 = 0.0f, p3 = 0.0f;
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 bv = *reinterpret_cast<const float4*>(bl + j), wv = *reinterpret_cast<const float4*>(wo + j);
            p0 = fmaf(wv.x, fmaxf(__uint_as_float(v[j]) + bv.x, 0.0f), p0);
            p1 = fmaf(wv.y, fmaxf(__uint_as_float(v[j + 1]) + bv.y, 0.0f), p1);
            p2 = fmaf(wv.z, fmaxf(__uint_as_float(v[j + 2]) + bv.z, 0.0f), p2);
            p3 = fmaf(wv.w, fmaxf(__uint_as_float(v[j + 3]) + bv.w, 0.0f), p3);
          }
          s_part[128 * cg + row] = (p0 + p1) + (p2 + p3);
          pkw_tc_fence_before();
        }
      }
      pkt_named_barrier(1, 32 * PKD_TW);
      pkd_trace(ttr, 0, xit - 1u, 7);
      if (cg == 0) {
        const long long o = b0 * A + row;
        if (row_ok && o < (long long)B * A)
          q[o] = ((s_part[row] + s_part[128 + row]) + (s_part[256 + row] + s_part[384 + row])) +
                 s_small[HP * lay.n_layers + HP];
      }
      pkt_named_barrier(2, 32 * PKD_TW);  // s_part is rewritten by the next tile's head
      }
      lo_clean[0] = lo_clean[1] = false;  // the hidden layers' A operand and accumulator were written over the A stages
    }
  }
  pkw_tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    pkw_tc_fence_after();
    pkw_tmem_dealloc(tmem_base, PKD_TMEM_COLS);
  }
}
