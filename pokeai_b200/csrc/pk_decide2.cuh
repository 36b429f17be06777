// The paired decision kernel: one warp takes TWO battles at a time, one per half-warp in the MLP.
//
// Why (profiles/r01_trace_*.log): with one battle per warp the hidden layer was bound by shared-memory
// wavefronts -- per k a warp fetched one 256-byte weight row (2 wavefronts) plus 9 broadcast activations for
// only 9 FFMA2.  Here the lane owns CH = HP/16 adjacent output channels of ITS half's battle, so the weight row
// (same address in both halves, still 2 wavefronts) now feeds 18 columns, and the activations of 4 k's of one
// column arrive as one LDS.128 with two distinct addresses (one per half): 17 wavefronts per 72 FFMA2 instead of
// 26 per 36.  The same pairing halves the wavefronts of the layer-0 state part; gathers of the all-zero row
// (empty slots, illegal columns) are predicated off instead of fetched.
//
// Parse, feature stores, selection and the reference semantics are those of pk_decide.cuh (same helpers).
#pragma once

#include "pk_decide.cuh"

#define PK2_REC_BYTES (2 * PK_WARP_REC_BYTES)
#define PK2_SV_BYTES (2 * PK_WARP_SV_BYTES)
#define PK2_IDX_BYTES (2 * PK_WARP_IDX_BYTES)
#define PK2_FIXED_BYTES (PK2_REC_BYTES + PK2_SV_BYTES + PK2_IDX_BYTES)

// `dual`: a second network (the target network of Double DQN) is staged behind the first; its tail / state rows
// start PK2_NET_BYTES(lay) after the first one's
#define PK2_NET_BYTES(lay) (((lay).total - (lay).b[0]) * 4u + (unsigned)PK_N_STATE * (lay).HP * 4u)
__host__ __device__ inline PkSmemPlan pk2_smem_plan(const PkLayout& lay, int A, int nwarps, bool zero_tile,
                                                    bool dual = false) {
  PkSmemPlan p;
  unsigned o = 0;
  p.tail = o;
  o += (lay.total - lay.b[0]) * 4u;
  p.w1s = o;
  o += (unsigned)PK_N_STATE * lay.HP * 4u;
  if (dual) o += PK2_NET_BYTES(lay);
  p.warp_area = o;
  p.warp_stride = PK2_FIXED_BYTES + 2u * (unsigned)A * lay.HP * 4u;
  o += p.warp_stride * nwarps;
  o = (o + 127u) & ~127u;
  p.zero_tile = o;
  p.zero_bytes = zero_tile ? ((((unsigned)lay.C * A * 4u) + 15u) & ~15u) : 0u;
  o += p.zero_bytes;
  p.wfrag = o;
  o = (o + 15u) & ~15u;
  p.mbar = o;
  o += 16u;
  p.total = o;
  return p;
}

// CH adjacent floats (CH = 2, 4 or 8), 8/16-byte vector accesses
template <int CH>
__device__ __forceinline__ void pk2_lds(const float* p, float (&v)[CH]) {
  if constexpr (CH == 2) {
    const float2 t = *reinterpret_cast<const float2*>(p);
    v[0] = t.x;
    v[1] = t.y;
  } else {
#pragma unroll
    for (int i = 0; i < CH; i += 4) {
      const float4 t = *reinterpret_cast<const float4*>(p + i);
      v[i] = t.x;
      v[i + 1] = t.y;
      v[i + 2] = t.z;
      v[i + 3] = t.w;
    }
  }
}
template <int CH>
__device__ __forceinline__ void pk2_sts(float* p, const float (&v)[CH]) {
  if constexpr (CH == 2) {
    *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
  } else {
#pragma unroll
    for (int i = 0; i < CH; i += 4) *reinterpret_cast<float4*>(p + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
  }
}
// gather of one W0^T row slice; rows equal to zero_row are not fetched (they are all zeros)
template <int CH>
__device__ __forceinline__ void pk2_gather(const float* __restrict__ w1t, int row, int zero_row, int off, int HP,
                                           float (&v)[CH]) {
#pragma unroll
  for (int i = 0; i < CH; ++i) v[i] = 0.0f;
  if (row != zero_row) {
    const float* p = w1t + (size_t)row * HP + off;
    if constexpr (CH == 2) {
      const float2 t = __ldg(reinterpret_cast<const float2*>(p));
      v[0] = t.x;
      v[1] = t.y;
    } else {
#pragma unroll
      for (int i = 0; i < CH; i += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p + i));
        v[i] = t.x;
        v[i + 1] = t.y;
        v[i + 2] = t.z;
        v[i + 3] = t.w;
      }
    }
  }
}

// hidden layer for the pair: input relu(h) in hb (this half's [A][HP] block), pre-activation out in acc
template <int CH, int A>
__device__ __forceinline__ void pk2_hidden(const float* __restrict__ WT, const float* __restrict__ bias,
                                           const float* hb, float (&acc)[A][CH], int hl) {
  constexpr int HP = 16 * CH;
  {
    float bv[CH];
    pk2_lds<CH>(bias + CH * hl, bv);
#pragma unroll
    for (int c = 0; c < A; ++c)
#pragma unroll
      for (int i = 0; i < CH; ++i) acc[c][i] = bv[i];
  }
  const float* wp = WT + CH * hl;
#pragma unroll 1
  for (int k = 0; k < HP; k += 4) {
    float4 hv[A];
#pragma unroll
    for (int c = 0; c < A; ++c) hv[c] = *reinterpret_cast<const float4*>(hb + c * HP + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float w[CH];
      pk2_lds<CH>(wp + (k + kk) * HP, w);
#pragma unroll
      for (int c = 0; c < A; ++c) {
        const float x = kk == 0 ? hv[c].x : kk == 1 ? hv[c].y : kk == 2 ? hv[c].z : hv[c].w;
        pk_fma_vec<CH>(acc[c], w, x);
      }
    }
  }
}

// DUAL (the next-state pass of the Double-DQN update, trainer.py:169-177): a second network `prep2` (the target) is
// staged too, and after the masked argmax a* of the first (online) network's Q the kernel evaluates the target network
// for column a* ONLY and stores v_out[b] = Q_target(s', a*) -- the records are parsed once, and the target pass costs
// one column instead of nine.
template <int CH, int A, bool DUAL = false>
__global__ void __launch_bounds__(512, 1)
pk_decide2_kernel(const uint4* __restrict__ packed, int B, const float* __restrict__ prep, PkLayout lay,
                  float* __restrict__ feat, int* __restrict__ mask, float* __restrict__ q,
                  int* __restrict__ action, PkSelect sel, int zero_by_tma, const float* __restrict__ prep2 = nullptr,
                  float* __restrict__ v_out = nullptr) {
  constexpr int HP = 16 * CH;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int half = lane >> 4, hl = lane & 15;
  const int L = lay.C * A;
  const bool use_tma = feat != nullptr && (zero_by_tma & 1) != 0;
  const PkSmemPlan plan = pk2_smem_plan(lay, A, nwarps, use_tma, DUAL);

  unsigned long long* tr = sel.trace;
  const int tslot = blockIdx.x * nwarps + warp;
  pk_trace(tr, tslot, 0);
  uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + plan.mbar);
  if (threadIdx.x == 0) {  // weights staged by two TMA bulk loads, waited for after the prologue barrier
    const uint32_t tail_bytes = (lay.total - lay.b[0]) * 4u, w1s_bytes = (uint32_t)PK_N_STATE * lay.HP * 4u;
    pk_mbar_init(wbar, 1);
    pk_fence_mbar_init();
    pk_mbar_expect_tx(wbar, (tail_bytes + w1s_bytes) * (DUAL ? 2u : 1u));
    pk_bulk_load(smem + plan.tail, prep + lay.b[0], tail_bytes, wbar);
    pk_bulk_load(smem + plan.w1s, prep + lay.w1t, w1s_bytes, wbar);
    if (DUAL) {
      pk_bulk_load(smem + plan.tail + PK2_NET_BYTES(lay), prep2 + lay.b[0], tail_bytes, wbar);
      pk_bulk_load(smem + plan.w1s + PK2_NET_BYTES(lay), prep2 + lay.w1t, w1s_bytes, wbar);
    }
  }
  const float* zero_tile = nullptr;
  if (use_tma) {
    float4* z = reinterpret_cast<float4*>(smem + plan.zero_tile);
    for (int i = threadIdx.x; i < (int)(plan.zero_bytes >> 4); i += blockDim.x) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pk_fence_proxy_async_smem();
    zero_tile = reinterpret_cast<const float*>(z);
  }

  unsigned char* wa = smem + plan.warp_area + warp * plan.warp_stride;
  uint4* rec = reinterpret_cast<uint4*>(wa);                                   // two 144-byte records
  float* sv2 = reinterpret_cast<float*>(wa + PK2_REC_BYTES);                   // [2][64]
  int* idx2 = reinterpret_cast<int*>(wa + PK2_REC_BYTES + PK2_SV_BYTES);       // [2][80]
  float* hbuf2 = reinterpret_cast<float*>(wa + PK2_FIXED_BYTES);               // [2][A][HP]
  const float* s_tail = reinterpret_cast<const float*>(smem + plan.tail);
  const float* s_w1s = reinterpret_cast<const float*>(smem + plan.w1s);
  const float* w1t = prep + lay.w1t;
  const int zero_row = lay.C;
  const unsigned tb = lay.b[0];

  // pairs of adjacent battles (2p, 2p+1) are dealt round-robin over every warp of the grid
  const long long NP = ((long long)B + 1) >> 1;
  const long long n16 = (long long)B * 9;  // 16-byte words of the packed array
  const long long stride = (long long)gridDim.x * nwarps;
  long long p = (long long)warp * gridDim.x + blockIdx.x;
  uint4 cur = make_uint4(0u, 0u, 0u, 0u);
  if (p < NP && lane < 18 && p * 18 + lane < n16) cur = __ldg(packed + p * 18 + lane);
  __syncthreads();  // the mbarrier is initialised and the zero tile is visible
  pk_mbar_wait(wbar, 0);  // staged weights have landed
  pk_trace(tr, tslot, 1);

  const uint8_t* r_own = reinterpret_cast<const uint8_t*>(rec) + PK_WARP_REC_BYTES * half;
  const float* sv_own = sv2 + 64 * half;
  const int* idx_own = idx2 + 80 * half;
  float* hb_own = hbuf2 + half * (A * HP);

  for (; p < NP; p += stride) {
    __syncwarp();
    if (lane < 18) rec[lane] = cur;
    {
      const long long pn = p + stride;
      cur = make_uint4(0u, 0u, 0u, 0u);
      if (pn < NP && lane < 18 && pn * 18 + lane < n16) cur = __ldg(packed + pn * 18 + lane);
    }
    __syncwarp();
    pk_trace(tr, tslot, 2);
    const long long b_own = 2 * p + half;
    const bool valid_own = b_own < B;
    const bool valid1 = 2 * p + 1 < B;  // warp-uniform: the second battle of the pair exists
    const int n_act_own = valid_own ? min((int)r_own[PK_OFF_NACT], A) : 0;
    if (mask != nullptr && valid_own && hl < A) mask[b_own * A + hl] = hl < n_act_own ? 1 : 0;
#pragma unroll
    for (int hb = 0; hb < 2; ++hb) {
      const uint8_t* r = reinterpret_cast<const uint8_t*>(rec) + PK_WARP_REC_BYTES * hb;
      const int n_act = (hb == 0 || valid1) ? min((int)r[PK_OFF_NACT], A) : 0;
      sv2[64 * hb + lane] = pk_state_feature(r, lane);
      sv2[64 * hb + lane + 32] = pk_state_feature(r, lane + 32);
#pragma unroll
      for (int it = 0; it < 3; ++it) {
        const int e = lane + 32 * it;
        if (e < 8 * A) {
          const int c = e >> 3;
          const int row = c < n_act ? pk_choice_row(r, c, e & 7) : -1;
          idx2[80 * hb + e] = row < 0 ? zero_row : row;
        }
      }
    }
    __syncwarp();
    pk_trace(tr, tslot, 3);

    float* gfeat0 = feat != nullptr ? feat + (size_t)(2 * p) * L : nullptr;
    if (gfeat0 != nullptr) {
      pk_store_feature_block<A>(gfeat0, L, sv2, zero_tile, lane);
      if (valid1) pk_store_feature_block<A>(gfeat0 + L, L, sv2 + 64, zero_tile, lane);
      if (!use_tma) {
        __syncwarp();
        pk_poke_feature_block<A>(gfeat0, idx2, zero_row, lane);
        if (valid1) pk_poke_feature_block<A>(gfeat0 + L, idx2 + 80, zero_row, lane);
      }
    }
    pk_trace(tr, tslot, 4);

    // ---- layer 0, state part of the own battle: b + sum_k W0^T[k] s_k (two chains) ------------------------
    float base[CH], base2[CH];
    pk2_lds<CH>(s_tail + CH * hl, base);  // lay.b[0] is the start of the tail
#pragma unroll
    for (int i = 0; i < CH; ++i) base2[i] = 0.0f;
    {
      const float* wp = s_w1s + CH * hl;
#pragma unroll 3
      for (int k4 = 0; k4 < 48; k4 += 4) {
        const float4 s4 = *reinterpret_cast<const float4*>(sv_own + k4);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const float x = kk == 0 ? s4.x : kk == 1 ? s4.y : kk == 2 ? s4.z : s4.w;
          float w[CH];
          pk2_lds<CH>(wp + (k4 + kk) * HP, w);
          if (kk & 1) pk_fma_vec<CH>(base2, w, x);
          else pk_fma_vec<CH>(base, w, x);
        }
      }
      const float2 s2 = *reinterpret_cast<const float2*>(sv_own + 48);
      float w[CH];
      pk2_lds<CH>(wp + 48 * HP, w);
      pk_fma_vec<CH>(base, w, s2.x);
      pk2_lds<CH>(wp + 49 * HP, w);
      pk_fma_vec<CH>(base2, w, s2.y);
    }
#pragma unroll
    for (int i = 0; i < CH; ++i) base[i] += base2[i];
    pk_trace(tr, tslot, 5);

    // ---- layer 0, choice part: <= 8 gathered rows per column, next column's loads in flight ---------------
    {
      float gc[8][CH], gn[8][CH];
      {
        const int4 ia = *reinterpret_cast<const int4*>(idx_own);
        const int4 ib = *reinterpret_cast<const int4*>(idx_own + 4);
        pk2_gather<CH>(w1t, ia.x, zero_row, CH * hl, HP, gc[0]);
        pk2_gather<CH>(w1t, ia.y, zero_row, CH * hl, HP, gc[1]);
        pk2_gather<CH>(w1t, ia.z, zero_row, CH * hl, HP, gc[2]);
        pk2_gather<CH>(w1t, ia.w, zero_row, CH * hl, HP, gc[3]);
        pk2_gather<CH>(w1t, ib.x, zero_row, CH * hl, HP, gc[4]);
        pk2_gather<CH>(w1t, ib.y, zero_row, CH * hl, HP, gc[5]);
        pk2_gather<CH>(w1t, ib.z, zero_row, CH * hl, HP, gc[6]);
        pk2_gather<CH>(w1t, ib.w, zero_row, CH * hl, HP, gc[7]);
      }
#pragma unroll
      for (int c = 0; c < A; ++c) {
        if (c + 1 < A) {
          const int4 ia = *reinterpret_cast<const int4*>(idx_own + 8 * (c + 1));
          const int4 ib = *reinterpret_cast<const int4*>(idx_own + 8 * (c + 1) + 4);
          pk2_gather<CH>(w1t, ia.x, zero_row, CH * hl, HP, gn[0]);
          pk2_gather<CH>(w1t, ia.y, zero_row, CH * hl, HP, gn[1]);
          pk2_gather<CH>(w1t, ia.z, zero_row, CH * hl, HP, gn[2]);
          pk2_gather<CH>(w1t, ia.w, zero_row, CH * hl, HP, gn[3]);
          pk2_gather<CH>(w1t, ib.x, zero_row, CH * hl, HP, gn[4]);
          pk2_gather<CH>(w1t, ib.y, zero_row, CH * hl, HP, gn[5]);
          pk2_gather<CH>(w1t, ib.z, zero_row, CH * hl, HP, gn[6]);
          pk2_gather<CH>(w1t, ib.w, zero_row, CH * hl, HP, gn[7]);
        }
        float hc[CH];
#pragma unroll
        for (int i = 0; i < CH; ++i) {
          const float s01 = gc[0][i] + gc[1][i], s23 = gc[2][i] + gc[3][i];
          const float s45 = gc[4][i] + gc[5][i], s67 = gc[6][i] + gc[7][i];
          hc[i] = fmaxf(base[i] + ((s01 + s23) + (s45 + s67)), 0.0f);
        }
        pk2_sts<CH>(hb_own + c * HP + CH * hl, hc);
        if (c + 1 < A) {
#pragma unroll
          for (int t = 0; t < 8; ++t)
#pragma unroll
            for (int i = 0; i < CH; ++i) gc[t][i] = gn[t][i];
        }
      }
    }
    __syncwarp();
    pk_trace(tr, tslot, 6);

    // ---- hidden layers + head --------------------------------------------------------------------------
    float qv[A];
    {
      float acc[A][CH];
      for (int l = 1; l < lay.n_layers; ++l) {
        pk2_hidden<CH, A>(s_tail + (lay.wl[l] - tb), s_tail + (lay.b[l] - tb), hb_own, acc, hl);
        if (l + 1 < lay.n_layers) {
          __syncwarp();
#pragma unroll
          for (int c = 0; c < A; ++c) {
            float hc[CH];
#pragma unroll
            for (int i = 0; i < CH; ++i) hc[i] = fmaxf(acc[c][i], 0.0f);
            pk2_sts<CH>(hb_own + c * HP + CH * hl, hc);
          }
          __syncwarp();
        }
      }
      if (lay.n_layers == 1) {
#pragma unroll
        for (int c = 0; c < A; ++c) pk2_lds<CH>(hb_own + c * HP + CH * hl, acc[c]);
      }
      float wo[CH];
      pk2_lds<CH>(s_tail + (lay.wout - tb) + CH * hl, wo);
      const float bo = s_tail[lay.bout - tb];
#pragma unroll
      for (int c = 0; c < A; ++c) {
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < CH; ++i) s = fmaf(wo[i], fmaxf(acc[c][i], 0.0f), s);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);  // stays inside the half-warp
        qv[c] = s + bo;
      }
    }
    __syncwarp();
    pk_trace(tr, tslot, 7);
    float myq = 0.0f;
#pragma unroll
    for (int c = 0; c < A; ++c)
      if (hl == c) myq = qv[c];
    if (q != nullptr && valid_own && hl < A) q[b_own * A + hl] = myq;
    if (action != nullptr || DUAL) {
      const bool writer = valid_own && hl == 0;
      const int act = pk_select_dev<A>(qv, (1u << n_act_own) - 1u, A, sel, valid_own ? b_own : 0, writer);
      if (writer && action != nullptr) action[b_own] = act;
      if constexpr (DUAL) {
        // ---- target network, column a* only: v = Q_target(s', a*) (trainer.py:175-177) ----
        const float* t_tail = reinterpret_cast<const float*>(smem + plan.tail + PK2_NET_BYTES(lay));
        const float* t_w1s = reinterpret_cast<const float*>(smem + plan.w1s + PK2_NET_BYTES(lay));
        const float* w1t2 = prep2 + lay.w1t;
        float gt[8][CH];
        {
          const int4 ia = *reinterpret_cast<const int4*>(idx_own + 8 * act);
          const int4 ib = *reinterpret_cast<const int4*>(idx_own + 8 * act + 4);
          pk2_gather<CH>(w1t2, ia.x, zero_row, CH * hl, HP, gt[0]);
          pk2_gather<CH>(w1t2, ia.y, zero_row, CH * hl, HP, gt[1]);
          pk2_gather<CH>(w1t2, ia.z, zero_row, CH * hl, HP, gt[2]);
          pk2_gather<CH>(w1t2, ia.w, zero_row, CH * hl, HP, gt[3]);
          pk2_gather<CH>(w1t2, ib.x, zero_row, CH * hl, HP, gt[4]);
          pk2_gather<CH>(w1t2, ib.y, zero_row, CH * hl, HP, gt[5]);
          pk2_gather<CH>(w1t2, ib.z, zero_row, CH * hl, HP, gt[6]);
          pk2_gather<CH>(w1t2, ib.w, zero_row, CH * hl, HP, gt[7]);
        }
        float tb0[CH], tb1[CH];
        pk2_lds<CH>(t_tail + CH * hl, tb0);
#pragma unroll
        for (int i = 0; i < CH; ++i) tb1[i] = 0.0f;
        {
          const float* wp = t_w1s + CH * hl;
#pragma unroll 3
          for (int k4 = 0; k4 < 48; k4 += 4) {
            const float4 s4 = *reinterpret_cast<const float4*>(sv_own + k4);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const float x = kk == 0 ? s4.x : kk == 1 ? s4.y : kk == 2 ? s4.z : s4.w;
              float w[CH];
              pk2_lds<CH>(wp + (k4 + kk) * HP, w);
              if (kk & 1) pk_fma_vec<CH>(tb1, w, x);
              else pk_fma_vec<CH>(tb0, w, x);
            }
          }
          const float2 s2 = *reinterpret_cast<const float2*>(sv_own + 48);
          float w[CH];
          pk2_lds<CH>(wp + 48 * HP, w);
          pk_fma_vec<CH>(tb0, w, s2.x);
          pk2_lds<CH>(wp + 49 * HP, w);
          pk_fma_vec<CH>(tb1, w, s2.y);
        }
        float ht[CH];
#pragma unroll
        for (int i = 0; i < CH; ++i) {
          const float s01 = gt[0][i] + gt[1][i], s23 = gt[2][i] + gt[3][i];
          const float s45 = gt[4][i] + gt[5][i], s67 = gt[6][i] + gt[7][i];
          ht[i] = fmaxf((tb0[i] + tb1[i]) + ((s01 + s23) + (s45 + s67)), 0.0f);
        }
        for (int l = 1; l < lay.n_layers; ++l) {
          __syncwarp();  // every lane is done with this half's hbuf (online pass / previous layer)
          pk2_sts<CH>(hb_own + CH * hl, ht);
          __syncwarp();
          float at[CH];
          pk2_lds<CH>(t_tail + (lay.b[l] - tb) + CH * hl, at);
          const float* wp = t_tail + (lay.wl[l] - tb) + CH * hl;
#pragma unroll 4
          for (int k = 0; k < HP; k += 4) {
            const float4 hv = *reinterpret_cast<const float4*>(hb_own + k);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              float w[CH];
              pk2_lds<CH>(wp + (k + kk) * HP, w);
              pk_fma_vec<CH>(at, w, kk == 0 ? hv.x : kk == 1 ? hv.y : kk == 2 ? hv.z : hv.w);
            }
          }
#pragma unroll
          for (int i = 0; i < CH; ++i) ht[i] = fmaxf(at[i], 0.0f);
        }
        float wo[CH];
        pk2_lds<CH>(t_tail + (lay.wout - tb) + CH * hl, wo);
        float sacc = 0.0f;
#pragma unroll
        for (int i = 0; i < CH; ++i) sacc = fmaf(wo[i], ht[i], sacc);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) sacc += __shfl_xor_sync(0xffffffffu, sacc, o);
        if (writer) v_out[b_own] = sacc + t_tail[lay.bout - tb];
      }
    }
    pk_trace(tr, tslot, 8);
    if (gfeat0 != nullptr && use_tma) {
      if (lane == 0) pk_bulk_wait_all();
      __syncwarp();
      pk_poke_feature_block<A>(gfeat0, idx2, zero_row, lane);
      if (valid1) pk_poke_feature_block<A>(gfeat0 + L, idx2 + 80, zero_row, lane);
    }
    pk_trace(tr, tslot, 9);
    tr = nullptr;
  }
}
