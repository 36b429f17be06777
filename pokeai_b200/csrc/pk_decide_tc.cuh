// The decision kernel for 64-channel networks with the hidden layers on the 5th-generation tensor cores
// (PK_DECIDE_TC / pk_debug_set_decide_tc; OFF by default -- see "measured" below).
//
// pk_decide_kernel (pk_decide.cuh) runs the 64 x 64 hidden layers per warp on the FP32 pipe: 576 FFMA2 and 352
// shared-memory loads per battle, 4.3 of the 12.4 us one battle takes (profiles/r02_trace_phases.txt).  Here the warps
// keep that kernel's structure -- one battle per warp, parse / feature fill / layer 0 in gather form / select exactly
// as there -- but the hidden layers of the EIGHT battles of a warp GROUP are one tensor-core contraction:
//
//   group     warps 8g .. 8g+7.  A group owns a 128-row A operand in shared memory and 64 TMEM columns.  A UMMA with
//             M = 128 keeps accumulator row i in TMEM lane i, and a warp may only read lanes 32 (w % 4) .. +31
//             (tcgen05.ld), so warp w writes relu(h0) of its A action columns into rows 32 (w % 4) + 16 ((w / 4) % 2) + c
//             and finds ITS OWN result rows in the lane quarter it is allowed to read.  The only coupling between
//             warps is a 256-thread named barrier per battle inside the group; the two groups of a CTA drift apart.
//   operands  fp32 values split x = hi + lo with hi = tf32(x) (cvt.rna) and lo = x - hi (exact); the product is
//             lo*hi + hi*lo + hi*hi accumulated in fp32 in TMEM by tcgen05.mma.kind::tf32 (the tensor core reads the
//             top 19 bits of lo: relative error 2^-21 per product, the dropped lo*lo term 2^-22), so Q stays inside
//             the 1e-5 fp32 tolerance of the parity tests.  Both operands are K-major with SWIZZLE_128B: A is written in
//             that layout straight from registers, B comes ready-made from the prepared weights (PkLayout::tc, kept
//             current by pk_clamp_adam) with one TMA bulk load per launch.
//   head      the lane that holds accumulator row c (64 columns) adds the bias, applies ReLU and takes the dot product
//             with the output weights: 64 FMAs; nine shuffles hand Q to the masked selection.
//
// Measured (B200, 4 096 states, features written; DESIGN.md 4.1c has the three designs that were built and timed):
//   * one 128-row tile per CTA, 14 warps, block-wide barrier per round of 14 battles: 26.6 - 27.8 us;
//   * two such tiles, 28 warps, one round: 38.8 us (all 92 MB of fills issued at once, every L2 read behind them);
//   * four groups of four warps with M = 64 tiles: 29.8 us;  this version (two groups of eight, M = 128): 28.0 us;
//   * pk_decide_kernel (FP32 pipe): 22.9 us.
// Why it does not win at this size: a hidden layer is 24 DEPENDENT K = 8 MMAs (3 split terms x 64 / 8) into one
// accumulator, i.e. tensor-core LATENCY, plus a group barrier, a TMEM round trip and the head -- 3.9 us per battle
// against 4.3 us for FFMA2 -- while the 160 KB of operand tiles leave 35 KB of L1 instead of 146 KB for the layer-0
// gathers (2.1 against 1.4 us), and every launch stages 46 KB instead of 30 KB.  What the exercise did show: global
// READS issued behind a burst of feature fills wait for the burst to drain (the weight loads moved in front of the
// prologue barrier for that reason, in both kernels).
//
// Reference semantics restated are those of pk_decide.cuh (same parse / feature / select helpers):
//   FeatureExtractor.transform feature_extractor.py:40-52, MLPModel.forward mlp_model.py:21-30,
//   Agent._act_by_model / _act_random agent.py:32-45.
#pragma once

#include "pk_wide.cuh"

#define PKT_H 64                   // channels (prepared HP == 64)
#define PKT_WARPS 16               // two groups of eight warps
#define PKT_GROUP 8                // warps per group
#define PKT_NGROUPS (PKT_WARPS / PKT_GROUP)
#define PKT_A_KB_BYTES 16384u      // one k-block (32 fp32 = 128 B) of the 128-row A operand tile
#define PKT_A_TILE_BYTES 32768u    // two k-blocks: K = 64
#define PKT_A_SET_BYTES 65536u     // hi tile + lo tile: one group's A operand
#define PKT_KB_BYTES 8192u         // one k-block of a 64-row B operand tile
#define PKT_TILE_BYTES 16384u
#define PKT_SET_BYTES 32768u       // hi tile + lo tile: one hidden matrix's B operand
#define PKT_ZERO_TILE_SMALL 4096u  // zero tile when the whole zero run does not fit beside two B operands

struct PktSmemPlan {
  unsigned a_tiles;      // per group: hi (32 KB), lo (32 KB); 1024-byte aligned
  unsigned b_tiles;      // per hidden layer: hi, lo (TMA-loaded from PkLayout::tc)
  unsigned w1s;          // the 50 state rows of W0^T
  unsigned small;        // b[0..n_layers-1] (64 floats each), wout (64), bout (4)
  unsigned warp_area;    // per warp: record, state values, gather rows
  unsigned zero_tile, zero_bytes;
  unsigned bars;         // mbarriers: weights + one per group; TMEM base slot
  unsigned total;        // including 1024 bytes of alignment slack
};

__host__ __device__ inline PktSmemPlan pkt_smem_plan(const PkLayout& lay, int A, bool zero_tile, unsigned max_smem) {
  PktSmemPlan p;
  unsigned o = 0;
  p.a_tiles = o;
  o += PKT_A_SET_BYTES * PKT_NGROUPS;
  p.b_tiles = o;
  o += PKT_SET_BYTES * (unsigned)(lay.n_layers - 1);
  p.w1s = o;
  o += (unsigned)PK_N_STATE * PKT_H * 4u;
  p.small = o;
  o += ((unsigned)lay.n_layers * PKT_H + PKT_H + 4u) * 4u;
  p.warp_area = o;
  o += (unsigned)PK_WARP_FIXED_BYTES * PKT_WARPS;
  o = (o + 127u) & ~127u;
  p.zero_tile = o;
  p.zero_bytes = 0u;
  if (zero_tile) {
    const unsigned full = (((unsigned)lay.C * A * 4u) + 15u) & ~15u;
    p.zero_bytes = (o + full + 64u + 1024u <= max_smem) ? full : PKT_ZERO_TILE_SMALL;
  }
  o += p.zero_bytes;
  o = (o + 15u) & ~15u;
  p.bars = o;
  o += 64u;
  p.total = o + 1024u;
  return p;
}

// byte offset of element (row, k) of the 128-row K-major SWIZZLE_128B fp32 A tile (pk_tc_tile_off, in bytes)
__device__ __forceinline__ uint32_t pkt_tile_off(int row, int k) {
  return pk_tc_tile_off((unsigned)row, (unsigned)k, PKT_A_KB_BYTES >> 2) << 2;
}

__device__ __forceinline__ void pkt_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// kind::tf32, fp32 accumulate, both operands K-major, M x N
__host__ __device__ constexpr uint32_t pkt_instr_desc(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void pkt_named_barrier(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// D (128 x 64, TMEM) = A (128 x 64) . W^T (64 x 64):  lo*hi + hi*lo + hi*hi, small terms first.  One thread.
__device__ __forceinline__ void pkt_issue_layer(uint32_t d_tmem, uint32_t a_hi, uint32_t b_hi) {
  constexpr uint32_t idesc = pkt_instr_desc(128, 64);
#pragma unroll 1
  for (int t = 0; t < 3; ++t) {
    const uint32_t ta = t == 0 ? a_hi + PKT_A_TILE_BYTES : a_hi, tb = t == 1 ? b_hi + PKT_TILE_BYTES : b_hi;
#pragma unroll
    for (int s = 0; s < 8; ++s) {  // UMMA_K = 8 fp32 = 32 B inside the swizzle atom (+2 in the 16-byte address field)
      const uint64_t ad = pkw_smem_desc_k128(ta + (uint32_t)(s >> 2) * PKT_A_KB_BYTES) + 2u * (uint64_t)(s & 3);
      const uint64_t bd = pkw_smem_desc_k128(tb + (uint32_t)(s >> 2) * PKT_KB_BYTES) + 2u * (uint64_t)(s & 3);
      pkt_mma_tf32(d_tmem, ad, bd, idesc, (t | s) != 0 ? 1u : 0u);
    }
  }
}

template <int A>
__global__ void __launch_bounds__(PKT_WARPS * 32, 1)
pk_decide_tc_kernel(const uint4* __restrict__ packed, int B, const float* __restrict__ prep, PkLayout lay,
                    float* __restrict__ feat, int* __restrict__ mask, float* __restrict__ q, int* __restrict__ action,
                    PkSelect sel, int zero_by_tma, unsigned max_smem) {
  constexpr int CPL = 2, HP = PKT_H, AT = A, NW = PKT_WARPS;
  // SWIZZLE_128B operand tiles need 1024-byte alignment; asking for it in the declaration (instead of rounding a pointer
  // up by hand) keeps the accesses in the shared address space (LDS / STS, not generic LD / ST)
  extern __shared__ __align__(1024) unsigned char smem[];
  if ((pk_smem_u32(smem) & 1023u) != 0u) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // group; TMEM lane quarter (warp % 4: the lanes this warp may read); the warp's 16-row half of that quarter
  const int grp = warp / PKT_GROUP, wq = warp & 3, wh = (warp / 4) & 1;
  const int row0 = 32 * wq + 16 * wh;  // accumulator rows row0 .. row0 + A - 1 = TMEM lanes of the same numbers
  const int L = lay.C * A;
  const bool use_tma = feat != nullptr && (zero_by_tma & 1) != 0;
  const PktSmemPlan plan = pkt_smem_plan(lay, A, use_tma, max_smem);
  const int n_hidden = lay.n_layers - 1;  // >= 1

  pk_grid_launch_dependents();
  unsigned long long* tr = sel.trace;
  const int tslot = blockIdx.x * NW + warp;
  pk_trace(tr, tslot, 0);

  unsigned char* wa = smem + plan.warp_area + warp * PK_WARP_FIXED_BYTES;
  uint4* rec = reinterpret_cast<uint4*>(wa);
  const uint8_t* r = reinterpret_cast<const uint8_t*>(rec);
  float* sv = reinterpret_cast<float*>(wa + PK_WARP_REC_BYTES);
  int* idx = reinterpret_cast<int*>(wa + PK_WARP_REC_BYTES + PK_WARP_SV_BYTES);
  const float* s_w1s = reinterpret_cast<const float*>(smem + plan.w1s);
  float* s_small = reinterpret_cast<float*>(smem + plan.small);  // b[l] at 64*l, wout at 64*n_layers, bout behind
  uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + plan.bars);
  uint64_t* gbar = wbar + 1 + grp;  // completion of this group's MMAs
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wbar + 1 + PKT_NGROUPS);
  const float* w1t = prep + lay.w1t;
  const int zero_row = lay.C;
  unsigned char* a_hi = smem + plan.a_tiles + (unsigned)grp * PKT_A_SET_BYTES;  // this group's A operand (lo behind hi)
  const uint32_t a_u32 = pk_smem_u32(a_hi), b_u32 = pk_smem_u32(smem + plan.b_tiles);

  // battles are dealt round-robin over every warp of the grid; the four warps of a group iterate together
  const long long stride = (long long)gridDim.x * NW;
  long long b = (long long)warp * gridDim.x + blockIdx.x;
  const long long b_first = (long long)(PKT_GROUP * grp) * gridDim.x + blockIdx.x;  // the group's first warp has the most battles
  const int n_iter = b_first < B ? (int)(((long long)B - b_first + stride - 1) / stride) : 0;
  uint4 cur = make_uint4(0u, 0u, 0u, 0u);
  if (b < B && lane < 9) cur = __ldg(packed + b * 9 + lane);

  if (threadIdx.x == 0) {
    pk_mbar_init(wbar, 1);
    for (int i = 0; i < PKT_NGROUPS; ++i) pk_mbar_init(wbar + 1 + i, 1);
    pk_fence_mbar_init();
  }
  if (warp == 0) pkw_tmem_alloc(tmem_slot, 64 * PKT_NGROUPS);
  const float* zero_tile = nullptr;
  if (use_tma) {
    float4* z = reinterpret_cast<float4*>(smem + plan.zero_tile);
    for (int i = threadIdx.x; i < (int)(plan.zero_bytes >> 4); i += blockDim.x) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pk_fence_proxy_async_smem();
    zero_tile = reinterpret_cast<const float*>(z);
  }

  auto take_and_parse = [&]() {
    if (lane < 9) rec[lane] = cur;
    if (b + stride < B && lane < 9) cur = __ldg(packed + (b + stride) * 9 + lane);  // prefetch, consumed next turn
    __syncwarp();
    const int na = min((int)r[PK_OFF_NACT], A);
    sv[lane] = pk_state_feature(r, lane);
    sv[lane + 32] = pk_state_feature(r, lane + 32);
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const int e = lane + 32 * it;
      if (e < 8 * A) {
        const int c = e >> 3;
        const int row = c < na ? pk_choice_row(r, c, e & 7) : -1;
        idx[e] = row < 0 ? zero_row : row;
      }
    }
    __syncwarp();
  };
  if (b < B) take_and_parse();
  pk_trace(tr, tslot, 1);
  // Everything above reads only the packed records and fills shared memory (programmatic dependent launch, see
  // pk_decide.cuh): weights are read and every global write happens below griddepcontrol.wait.
  if (warp == 0) {
    // Weights: the 50 state rows of W0^T, biases, head, and the hidden matrices as ready-made tensor-core operand
    // tiles -- TMA bulk loads against one mbarrier, issued BEFORE the prologue barrier, i.e. before any warp's feature
    // fill: the SM's TMA unit serves requests in order, and loads queued behind sixteen 20 KB fills were measured to
    // land 5 us later.  Waited for just before layer 0.
    pk_grid_dependency_wait();
    if (lane == 0) {
      const uint32_t w1s_bytes = (uint32_t)PK_N_STATE * HP * 4u;
      pk_mbar_expect_tx(wbar, w1s_bytes + (uint32_t)(lay.n_layers * HP + HP + 4) * 4u + (uint32_t)n_hidden * PKT_SET_BYTES);
      pk_bulk_load(smem + plan.w1s, prep + lay.w1t, w1s_bytes, wbar);
      for (int l = 0; l < lay.n_layers; ++l) pk_bulk_load(s_small + HP * l, prep + lay.b[l], HP * 4u, wbar);
      pk_bulk_load(s_small + HP * lay.n_layers, prep + lay.wout, HP * 4u, wbar);
      pk_bulk_load(s_small + HP * lay.n_layers + HP, prep + lay.bout, 16u, wbar);
      for (int l = 1; l <= n_hidden; ++l)
        pk_bulk_load(smem + plan.b_tiles + (unsigned)(l - 1) * PKT_SET_BYTES, prep + lay.tc[l], PKT_SET_BYTES, wbar);
    }
    __syncwarp();
  }
  pkw_tc_fence_before();
  __syncthreads();
  pkw_tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pk_trace(tr, tslot, 2);
  pk_grid_dependency_wait();  // returns at once: warp 0 waited above and the barrier is behind it
  pk_trace(tr, tslot, 3);
  float b1r[CPL] = {0.0f, 0.0f};
  uint32_t gphase = 0;

  for (int it = 0; it < n_iter; ++it, b += stride) {
    const bool valid = b < B;
    if (it > 0 && valid) {
      __syncwarp();  // the previous battle's readers of rec / sv / idx are done
      take_and_parse();
    }
    const int n_act = valid ? min((int)r[PK_OFF_NACT], A) : 0;
    float* gfeat = (feat != nullptr && valid) ? feat + (size_t)b * L : nullptr;
    if (valid) {
      if (mask != nullptr && lane < A) mask[b * A + lane] = lane < n_act ? 1 : 0;  // feature_extractor.py:50-51
      if (gfeat != nullptr) {
        pk_store_feature_block<A>(gfeat, L, sv, zero_tile, lane, plan.zero_bytes);
        if (!use_tma) {
          __syncwarp();  // orders the fill before the pokes of other lanes (same addresses)
          pk_poke_feature_block<A>(gfeat, idx, zero_row, lane);
        }
      }
    }
    pk_trace(tr, tslot, 4);
    if (it == 0) {
      pk_mbar_wait(wbar, 0);  // the staged weights have landed (async proxy -> visible after the wait)
      pk_lds_vec<CPL>(s_small, lane, b1r);
    }
    if (valid) {
      // ---- layer 0 in gather form (as pk_decide_kernel): the first GD columns' rows are requested before the
      //      state part so that their L2 latency sits under the shared-memory FMAs ----
      constexpr int GD = AT < 4 ? AT : 4;
      float gr[GD][8][CPL];
#pragma unroll
      for (int c = 0; c < GD; ++c) pk_gather_column<CPL>(w1t, idx + 8 * c, zero_row, HP, lane, gr[c]);
      float base[CPL], base2[CPL];
#pragma unroll
      for (int i = 0; i < CPL; ++i) {
        base[i] = b1r[i];
        base2[i] = 0.0f;
      }
#pragma unroll 3
      for (int k4 = 0; k4 < 48; k4 += 4) {
        const float4 s4 = *reinterpret_cast<const float4*>(sv + k4);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const float x = kk == 0 ? s4.x : kk == 1 ? s4.y : kk == 2 ? s4.z : s4.w;
          float w[CPL];
          pk_lds_vec<CPL>(s_w1s + (k4 + kk) * HP, lane, w);
          if (kk & 1) pk_fma_vec<CPL>(base2, w, x);
          else pk_fma_vec<CPL>(base, w, x);
        }
      }
      {
        const float2 s2 = *reinterpret_cast<const float2*>(sv + 48);
        float w[CPL];
        pk_lds_vec<CPL>(s_w1s + 48 * HP, lane, w);
        pk_fma_vec<CPL>(base, w, s2.x);
        pk_lds_vec<CPL>(s_w1s + 49 * HP, lane, w);
        pk_fma_vec<CPL>(base2, w, s2.y);
      }
#pragma unroll
      for (int i = 0; i < CPL; ++i) base[i] += base2[i];
      pk_trace(tr, tslot, 5);
      // relu(h0) of column c -> row row0 + c of the group's A operand (hi and lo), channels 2*lane, 2*lane+1
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        float hc[CPL];
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
          const float s01 = gr[c % GD][0][i] + gr[c % GD][1][i], s23 = gr[c % GD][2][i] + gr[c % GD][3][i];
          const float s45 = gr[c % GD][4][i] + gr[c % GD][5][i], s67 = gr[c % GD][6][i] + gr[c % GD][7][i];
          hc[i] = fmaxf(base[i] + ((s01 + s23) + (s45 + s67)), 0.0f);
        }
        const float h0 = pk_tf32_hi(hc[0]), h1 = pk_tf32_hi(hc[1]);
        const uint32_t off = pkt_tile_off(row0 + c, 2 * lane);
        *reinterpret_cast<float2*>(a_hi + off) = make_float2(h0, h1);
        *reinterpret_cast<float2*>(a_hi + PKT_A_TILE_BYTES + off) = make_float2(hc[0] - h0, hc[1] - h1);
        if (c + GD < AT) pk_gather_column<CPL>(w1t, idx + 8 * (c + GD), zero_row, HP, lane, gr[c % GD]);
      }
    }
    pk_trace(tr, tslot, 6);
    // ---- hidden layers: one tensor-core contraction per layer for the eight battles of the group ----
    float qv[AT];
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * wq) << 16) + 64u * (uint32_t)grp;
    for (int l = 1; l <= n_hidden; ++l) {
      pk_fence_proxy_async_smem();  // the tensor core (async proxy) reads what the generic proxy just wrote
      pkw_tc_fence_before();
      pkt_named_barrier(1 + grp, 32 * PKT_GROUP);
      if (warp == grp * PKT_GROUP) {
        pkw_tc_fence_after();
        if (lane == 0) {
          pkt_issue_layer(tmem_base + 64u * (uint32_t)grp, a_u32, b_u32 + (uint32_t)(l - 1) * PKT_SET_BYTES);
          pkw_mma_commit(gbar);
        }
        __syncwarp();
      }
      pk_mbar_wait(gbar, gphase);
      gphase ^= 1u;
      pkw_tc_fence_after();
      const float* bl = s_small + HP * l;
      if (l == n_hidden) {
        // head: lane 16 wh + c holds accumulator row row0 + c; q = w_out . relu(D + b) + b_out
        const float* wo = s_small + HP * lay.n_layers;
        float p0 = 0.0f, p1 = 0.0f, p2 = 0.0f, p3 = 0.0f;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          pkw_tmem_ld32(taddr + 32u * (uint32_t)half, v);
          pkw_tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bv = *reinterpret_cast<const float4*>(bl + 32 * half + j),
                         wv = *reinterpret_cast<const float4*>(wo + 32 * half + j);
            p0 = fmaf(wv.x, fmaxf(__uint_as_float(v[j]) + bv.x, 0.0f), p0);
            p1 = fmaf(wv.y, fmaxf(__uint_as_float(v[j + 1]) + bv.y, 0.0f), p1);
            p2 = fmaf(wv.z, fmaxf(__uint_as_float(v[j + 2]) + bv.z, 0.0f), p2);
            p3 = fmaf(wv.w, fmaxf(__uint_as_float(v[j + 3]) + bv.w, 0.0f), p3);
          }
        }
        const float qrow = ((p0 + p1) + (p2 + p3)) + wo[HP];
#pragma unroll
        for (int c = 0; c < AT; ++c) qv[c] = __shfl_sync(0xffffffffu, qrow, 16 * wh + c);
      } else {
        // relu(D + b) is the next hidden layer's input: the lane that holds a row writes it into the A operand
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          pkw_tmem_ld32(taddr + 32u * (uint32_t)half, v);
          pkw_tmem_ld_wait();
          if (lane >= 16 * wh && lane < 16 * wh + A) {  // the lanes holding this warp's rows
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 bv = *reinterpret_cast<const float4*>(bl + 32 * half + j);
              float x[4], hi[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                x[i] = fmaxf(__uint_as_float(v[j + i]) + (i == 0 ? bv.x : i == 1 ? bv.y : i == 2 ? bv.z : bv.w), 0.0f);
                hi[i] = pk_tf32_hi(x[i]);
              }
              const uint32_t off = pkt_tile_off(32 * wq + lane, 32 * half + j);
              *reinterpret_cast<float4*>(a_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
              *reinterpret_cast<float4*>(a_hi + PKT_A_TILE_BYTES + off) =
                  make_float4(x[0] - hi[0], x[1] - hi[1], x[2] - hi[2], x[3] - hi[3]);
            }
          }
        }
      }
      pkw_tc_fence_before();  // orders the tcgen05.ld above before the next MMA into the same columns
    }
    pk_trace(tr, tslot, 7);
    if (valid) {
      if (q != nullptr && lane < A) {
        float myq = 0.0f;
#pragma unroll
        for (int c = 0; c < AT; ++c)
          if (lane == c) myq = qv[c];
        q[b * A + lane] = myq;
      }
      if (action != nullptr) {
        const int act = pk_select_dev<AT>(qv, (1u << n_act) - 1u, A, sel, b, lane == 0);
        if (lane == 0) action[b] = act;
      }
      pk_trace(tr, tslot, 8);
      if (gfeat != nullptr && use_tma) {  // the zeros have landed by now; finish the block
        if (lane == 0) pk_bulk_wait_all();
        __syncwarp();
        pk_poke_feature_block<A>(gfeat, idx, zero_row, lane);
      }
      pk_trace(tr, tslot, 9);
    }
    tr = nullptr;  // first battle of the warp only
  }
  pkw_tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    pkw_tc_fence_after();
    pkw_tmem_dealloc(tmem_base, 64 * PKT_NGROUPS);
  }
}
