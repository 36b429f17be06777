// The decision kernel: packed observations -> [features] [mask] -> Q values -> action.
//
// One persistent CTA per SM; every warp takes one battle at a time and is independent of the others
// (no block-level barrier after the prologue):
//
//   parse     the 144-byte record is read once (nine 16-byte loads, prefetched one battle ahead),
//             the 50 state values and the <= 72 one-hot rows are derived branch-free into the warp's
//             shared-memory scratch.
//   features  (optional) the (C, A) fp32 block of the battle is a 22 KB run of zeros with ~450 state
//             values in front and <= 72 ones scattered behind.  The warp streams it straight from
//             registers with coalesced 16-byte stores (zeros need no staging), then pokes the ones.
//             HBM-write bound: 22 212 B per battle.
//   MLP       never reads the dense features.  Layer 0 is evaluated in gather form
//             (b + sum_k W[:,k] s_k over the non-zero state values + the <= 8 one-hot rows of each
//             action column, fetched from the transposed weights through L1); the hidden layers run out
//             of shared-memory weights with the lane owning adjacent output channels and all A columns,
//             using packed fp32x2 FMAs (FFMA2); the head is a warp butterfly; the masked argmax /
//             epsilon-greedy / softmax draw happens in registers.
//
// Reference semantics being restated (bit-exact for features and masks):
//   FeatureExtractor.transform           feature_extractor.py:40-52
//   StateFeatureExtractor._transform_*   state_feature_extractor.py:113-197
//   ChoiceToVec.transform                choice_to_vec.py:40-61
//   MLPModel.forward                     mlp_model.py:21-30
//   Agent._act_by_model / _act_random    agent.py:32-45, agent_train.py:30-33
#pragma once

#include <math_constants.h>

#include "pk_common.cuh"

__constant__ PkDeviceTables c_tab;

#define PK_WARP_REC_BYTES 144   // the record being processed
#define PK_WARP_SV_BYTES 256    // 64 floats: the 50 state values, zero padded
#define PK_WARP_IDX_BYTES 320   // 80 ints: gather rows, 8 per action column
#define PK_WARP_FIXED_BYTES (PK_WARP_REC_BYTES + PK_WARP_SV_BYTES + PK_WARP_IDX_BYTES)

// ---- state features (state_feature_extractor.py:113-197) ----------------------------------------
// r points at one 144-byte record in shared memory.  Every feature is num/den of two small integers
// (den = 1 for the one-hots).  Python computes a/b in double and numpy stores it as float32; IEEE fp32
// division of two integers < 2^24 gives the same bits (double rounding is innocuous for division when
// 53 >= 2*24+2), so one __fdiv_rn is exact parity -- never __fdividef.
__device__ __forceinline__ float pk_state_feature(const uint8_t* r, int k) {
  int num = 0, den = 1;
  if (k < PK_ST_TYPE) {  // :113-123 remaining / total
    num = r[PK_OFF_REMAINING + k];
    den = r[PK_OFF_TOTAL + k];
  } else if (k < PK_ST_HP) {  // :125-138 opponent active's types
    num = (int)((c_tab.type_mask[r[PK_OFF_OPP]] >> (k - PK_ST_TYPE)) & 1u);
  } else if (k < PK_ST_NV) {  // :140-151 hp ratio
    num = reinterpret_cast<const uint16_t*>(r + PK_OFF_HPCUR)[k - PK_ST_HP];
    den = reinterpret_cast<const uint16_t*>(r + PK_OFF_HPMAX)[k - PK_ST_HP];
  } else if (k < PK_ST_RANK) {  // :153-167 non-volatile condition one-hot
    const int side = (k >= PK_ST_NV + 6) ? 1 : 0;
    num = r[PK_OFF_STATUS + side] == (k - PK_ST_NV - 6 * side) + 1 ? 1 : 0;
  } else if (k < PK_ST_WEATHER) {  // :169-182 (rank + 6) / 12
    num = (int)(int8_t)r[PK_OFF_BOOST + (k - PK_ST_RANK)] + 6;
    den = 12;
  } else if (k < PK_N_STATE) {  // :184-197 weather one-hot
    num = r[PK_OFF_WEATHER] == (k - PK_ST_WEATHER) + 1 ? 1 : 0;
  }
  return __fdiv_rn((float)num, (float)den);
}

// ---- choice rows (choice_to_vec.py:40-61) -----------------------------------------------------
// Slot t of action column c: 0 switch flag, 1 force_switch flag, 2 species, 3 item, 4..7 moves.
// Returns the absolute feature row, or -1 when the slot is empty / out of table range / a repeat
// of an earlier move of the same action (the reference sets the element to 1, it does not add).
// Straight-line selects: the three action words are the same for the 8 lanes of a column.
__device__ __forceinline__ int pk_choice_row(const uint8_t* r, int c, int t) {
  const uint32_t* aw = reinterpret_cast<const uint32_t*>(r + PK_OFF_ACT) + 3 * c;
  const uint32_t w0 = aw[0], w1 = aw[1], w2 = aw[2];
  const uint32_t flags = w0 & 0xffu, sp = (w0 >> 8) & 0xffu, it = (w0 >> 16) & 0xffu, nm = w0 >> 24;
  const uint32_t m0 = w1 & 0xffffu, m1 = w1 >> 16, m2 = w2 & 0xffffu, m3 = w2 >> 16;
  const uint32_t mv = t == 4 ? m0 : t == 5 ? m1 : t == 6 ? m2 : m3;
  const bool dup = (t >= 5 && mv == m0) || (t >= 6 && mv == m1) || (t == 7 && mv == m2);
  const bool mv_ok = (uint32_t)(t - 4) < nm && mv < (uint32_t)c_tab.n_moves && !dup;
  int row = mv_ok ? c_tab.row_move + (int)mv : -1;                                           // t >= 4
  row = t == 0 ? ((flags & PK_FLAG_SWITCH) ? PK_N_STATE + PK_CH_SWITCH : -1) : row;
  row = t == 1 ? ((flags & PK_FLAG_FORCE) ? PK_N_STATE + PK_CH_FORCE : -1) : row;
  row = t == 2 ? (sp < (uint32_t)c_tab.n_species ? PK_N_STATE + PK_CH_SPECIES + (int)sp : -1) : row;
  row = t == 3 ? (it < (uint32_t)c_tab.n_items ? c_tab.row_item + (int)it : -1) : row;
  return row;
}

// ---- action selection (agent.py:32-45, agent_train.py:30-33) --------------------------------------
template <int AT>
__device__ __forceinline__ int pk_select_dev(const float (&qv)[AT], uint32_t legal, int A, const PkSelect& sel,
                                             long long b, bool writer) {
  legal &= (1u << A) - 1u;
  // q[mask == 0] = -inf; np.argmax: first maximum wins, a NaN counts as the maximum
  float best = -CUDART_INF_F;
  int idx = 0;
  bool best_nan = false;
#pragma unroll
  for (int c = 0; c < AT; ++c) {
    if (c < A) {
      const float v = ((legal >> c) & 1u) ? qv[c] : -CUDART_INF_F;
      if (!best_nan && (v > best || v != v)) {
        best = v;
        idx = c;
        best_nan = (v != v);
      }
    }
  }
  const int n = __popc(legal);
  if (sel.mode == PK_SELECT_EPSGREEDY) {
    // np.random.random() < eps -> np.random.choice(np.flatnonzero(mask)), with the two uniforms supplied
    if (n > 0 && sel.u_explore[b] < sel.eps) {
      int k = (int)__fmul_rn(sel.u_pick[b], (float)n);
      k = min(k, n - 1);
      uint32_t m = legal;
      for (int i = 0; i < k; ++i) m &= m - 1u;
      idx = __ffs(m) - 1;
    }
  } else if (sel.mode == PK_SELECT_SOFTMAX) {
    // extension (no reference counterpart): p = softmax(q / T) over legal slots, inverse-CDF draw
    float e[AT];
    float total = 0.0f;
    if (n > 0) {
      float mx = -CUDART_INF_F;
#pragma unroll
      for (int c = 0; c < AT; ++c)
        if (c < A && ((legal >> c) & 1u)) mx = fmaxf(mx, qv[c]);
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        e[c] = 0.0f;
        if (c < A && ((legal >> c) & 1u)) {
          e[c] = expf(__fdiv_rn(__fsub_rn(qv[c], mx), sel.temperature));
          total = __fadd_rn(total, e[c]);
        }
      }
      const float thr = __fmul_rn(sel.u_pick[b], total);
      float run = 0.0f;
      int pick = 31 - __clz(legal);
      bool found = false;
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        if (c < A && ((legal >> c) & 1u)) {
          run = __fadd_rn(run, e[c]);
          if (!found && run > thr) {
            pick = c;
            found = true;
          }
        }
      }
      idx = pick;
    } else {
      idx = 0;
#pragma unroll
      for (int c = 0; c < AT; ++c) e[c] = 0.0f;
    }
    if (sel.probs != nullptr && writer) {
#pragma unroll
      for (int c = 0; c < AT; ++c)
        if (c < A) sel.probs[b * A + c] = (n > 0 && ((legal >> c) & 1u)) ? __fdiv_rn(e[c], total) : 0.0f;
    }
  }
  return idx;
}

// ---- vector helpers: the lane owns CPL adjacent channels, j = CPL * lane + i -------------------------
template <int CPL>
__device__ __forceinline__ void pk_ldg_vec(const float* __restrict__ row, int lane, float (&v)[CPL]) {
  if constexpr (CPL == 1) {
    v[0] = __ldg(row + lane);
  } else if constexpr (CPL == 2) {
    const float2 t = __ldg(reinterpret_cast<const float2*>(row) + lane);
    v[0] = t.x;
    v[1] = t.y;
  } else {
    static_assert(CPL == 4, "CPL must be 1, 2 or 4");
    const float4 t = __ldg(reinterpret_cast<const float4*>(row) + lane);
    v[0] = t.x;
    v[1] = t.y;
    v[2] = t.z;
    v[3] = t.w;
  }
}
template <int CPL>
__device__ __forceinline__ void pk_lds_vec(const float* row, int lane, float (&v)[CPL]) {
  if constexpr (CPL == 1) {
    v[0] = row[lane];
  } else if constexpr (CPL == 2) {
    const float2 t = reinterpret_cast<const float2*>(row)[lane];
    v[0] = t.x;
    v[1] = t.y;
  } else {
    const float4 t = reinterpret_cast<const float4*>(row)[lane];
    v[0] = t.x;
    v[1] = t.y;
    v[2] = t.z;
    v[3] = t.w;
  }
}
template <int CPL>
__device__ __forceinline__ void pk_sts_vec(float* row, int lane, const float (&v)[CPL]) {
  if constexpr (CPL == 1) {
    row[lane] = v[0];
  } else if constexpr (CPL == 2) {
    reinterpret_cast<float2*>(row)[lane] = make_float2(v[0], v[1]);
  } else {
    reinterpret_cast<float4*>(row)[lane] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// acc[0..CPL) += w[0..CPL) * x   -- packed fp32x2 FMA (FFMA2, sm_100+) when the lane owns channel pairs;
// each half is an IEEE fma.rn, identical to two scalar fmaf.
template <int CPL>
__device__ __forceinline__ void pk_fma_vec(float (&acc)[CPL], const float (&w)[CPL], float x) {
  if constexpr (CPL == 1) {
    acc[0] = fmaf(w[0], x, acc[0]);
  } else {
#pragma unroll
    for (int p = 0; p < CPL; p += 2) {
      unsigned long long a, ww, xx, d;
      asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(acc[p]), "f"(acc[p + 1]));
      asm("mov.b64 %0, {%1, %2};" : "=l"(ww) : "f"(w[p]), "f"(w[p + 1]));
      asm("mov.b64 %0, {%1, %1};" : "=l"(xx) : "f"(x));
      asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(ww), "l"(xx), "l"(a));
      asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[p]), "=f"(acc[p + 1]) : "l"(d));
    }
  }
}

// the 8 row slots of one action column: eight independent loads of W0^T row slices.  Empty slots and illegal
// columns carry zero_row (all zeros): they are not fetched (warp-uniform predicate) -- the gathers are bound by L1
// bandwidth (256 B per row and warp), and a typical state uses 47 of its 72 slots.
template <int CPL>
__device__ __forceinline__ void pk_gather_row(const float* __restrict__ w1t, int row, int zero_row, int HP, int lane,
                                              float (&v)[CPL]) {
#pragma unroll
  for (int i = 0; i < CPL; ++i) v[i] = 0.0f;
  if (row != zero_row) pk_ldg_vec<CPL>(w1t + (size_t)row * HP, lane, v);
}
template <int CPL>
__device__ __forceinline__ void pk_gather_column(const float* __restrict__ w1t, const int* idx8, int zero_row, int HP,
                                                 int lane, float (&g)[8][CPL]) {
  const int4 ia = *reinterpret_cast<const int4*>(idx8);
  const int4 ib = *reinterpret_cast<const int4*>(idx8 + 4);
  pk_gather_row<CPL>(w1t, ia.x, zero_row, HP, lane, g[0]);
  pk_gather_row<CPL>(w1t, ia.y, zero_row, HP, lane, g[1]);
  pk_gather_row<CPL>(w1t, ia.z, zero_row, HP, lane, g[2]);
  pk_gather_row<CPL>(w1t, ia.w, zero_row, HP, lane, g[3]);
  pk_gather_row<CPL>(w1t, ib.x, zero_row, HP, lane, g[4]);
  pk_gather_row<CPL>(w1t, ib.y, zero_row, HP, lane, g[5]);
  pk_gather_row<CPL>(w1t, ib.z, zero_row, HP, lane, g[6]);
  pk_gather_row<CPL>(w1t, ib.w, zero_row, HP, lane, g[7]);
}

// ---- hidden layer on the warp-level tensor path (64 channels): 3xTF32 split, fp32 accuracy --------------------
// D (64 channels x <= 16 columns) = W (64 x 64) . h (64 x columns) with mma.sync.m16n8k8 tf32.  Every fp32 operand is
// split x = hi + lo (hi = tf32(x), lo = tf32(x - hi)) and the product is hi*hi + lo*hi + hi*lo with fp32
// accumulation: the dropped lo*lo term is 2^-22 relative, so Q stays inside the 1e-5 fp32 tolerance while the
// 36 864 FMAs per battle leave the FP32 pipe (profiles/r01_trace_phases.txt: that pipe bounded the kernel).
// The weight fragments are laid out once per CTA in the order the lanes consume them: wfrag[kstep][mtile][hi/lo][lane].
#define PK_TC_HS 68                // hbuf row stride in floats: 4 * col + k makes the B-fragment loads conflict-free
#define PK_TC_WFRAG_BYTES 32768u   // 8 k-steps x 4 m-tiles x {hi, lo} x 32 lanes x 16 B
__device__ __forceinline__ uint32_t pk_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void pk_mma_tf32(float (&d)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
               "{%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}
// prologue: W^T [k][64] (prepared layout) -> fragment order, split into hi / lo
__device__ __forceinline__ void pk_tc_build_wfrag(const float* __restrict__ wt, uint4* wfrag) {
  for (int e = threadIdx.x; e < 8 * 4 * 32; e += blockDim.x) {
    const int ln = e & 31, mt = (e >> 5) & 3, s = e >> 7;
    const int j = 16 * mt + (ln >> 2), k = 8 * s + (ln & 3);
    const float w[4] = {__ldg(wt + k * 64 + j), __ldg(wt + k * 64 + j + 8), __ldg(wt + (k + 4) * 64 + j),
                        __ldg(wt + (k + 4) * 64 + j + 8)};
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      hi[i] = pk_tf32(w[i]);
      lo[i] = pk_tf32(w[i] - __uint_as_float(hi[i]));
    }
    wfrag[((s * 4 + mt) * 2 + 0) * 32 + ln] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    wfrag[((s * 4 + mt) * 2 + 1) * 32 + ln] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}
// d[mt][nt][0..3]: channel 16*mt + lane/4 (+8 for [2],[3]), column 8*nt + 2*(lane%4) (+1 for [1],[3]); pre-activation
template <int A>
__device__ __forceinline__ void pk_hidden_tc(const uint4* wfrag, const float* bias, const float* hb,
                                             float (&d)[4][(A + 7) / 8][4], int lane) {
  constexpr int NT = (A + 7) / 8;
  const int g4 = lane >> 2, t4 = lane & 3;
#pragma unroll
  for (int mt = 0; mt < 4; ++mt) {
    const float b0 = bias[16 * mt + g4], b1 = bias[16 * mt + g4 + 8];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      d[mt][nt][0] = b0;
      d[mt][nt][1] = b0;
      d[mt][nt][2] = b1;
      d[mt][nt][3] = b1;
    }
  }
#pragma unroll 2
  for (int s = 0; s < 8; ++s) {
    uint32_t bh[NT][2], bl[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int col = 8 * nt + g4;
      float x0 = 0.0f, x1 = 0.0f;
      if (col < A) {
        x0 = hb[col * PK_TC_HS + 8 * s + t4];
        x1 = hb[col * PK_TC_HS + 8 * s + t4 + 4];
      }
      bh[nt][0] = pk_tf32(x0);
      bh[nt][1] = pk_tf32(x1);
      bl[nt][0] = pk_tf32(x0 - __uint_as_float(bh[nt][0]));
      bl[nt][1] = pk_tf32(x1 - __uint_as_float(bh[nt][1]));
    }
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      const uint4 ahi = wfrag[((s * 4 + mt) * 2 + 0) * 32 + lane];
      const uint4 alo = wfrag[((s * 4 + mt) * 2 + 1) * 32 + lane];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        pk_mma_tf32(d[mt][nt], alo, bh[nt][0], bh[nt][1]);  // small terms first
        pk_mma_tf32(d[mt][nt], ahi, bl[nt][0], bl[nt][1]);
        pk_mma_tf32(d[mt][nt], ahi, bh[nt][0], bh[nt][1]);
      }
    }
  }
}

// ---- shared-memory plan ---------------------------------------------------------------------------
struct PkSmemPlan {
  unsigned tail;         // prepared floats from lay.b[0] to lay.total (biases, hidden weights, head)
  unsigned w1s;          // first PK_N_STATE rows of W1^T
  unsigned warp_area;    // per-warp scratch
  unsigned warp_stride;
  unsigned zero_tile;    // block-wide tile of zeros: source of the TMA zero fill
  unsigned zero_bytes;
  unsigned wfrag;        // tensor-path weight fragments of the hidden layers (PK_TC_WFRAG_BYTES each), or unused
  unsigned mbar;         // one mbarrier: completion of the TMA bulk loads that stage the weights
  unsigned total;
};

__host__ __device__ inline PkSmemPlan pk_smem_plan(const PkLayout& lay, int AT, int nwarps, bool with_mlp,
                                                   bool zero_tile, bool tc = false) {
  PkSmemPlan p;
  unsigned o = 0;
  p.tail = o;
  if (with_mlp) o += (lay.total - lay.b[0]) * 4u;
  p.w1s = o;
  if (with_mlp) o += (unsigned)PK_N_STATE * lay.HP * 4u;
  p.warp_area = o;
  p.warp_stride = PK_WARP_FIXED_BYTES + (with_mlp ? (unsigned)AT * (tc ? (unsigned)PK_TC_HS : lay.HP) * 4u : 0u);
  o += p.warp_stride * nwarps;
  o = (o + 127u) & ~127u;
  p.zero_tile = o;
  p.zero_bytes = zero_tile ? ((((unsigned)lay.C * AT * 4u) + 15u) & ~15u) : 0u;
  o += p.zero_bytes;
  p.wfrag = o;
  if (tc) o += PK_TC_WFRAG_BYTES * (unsigned)(lay.n_layers - 1);
  o = (o + 15u) & ~15u;
  p.mbar = o;
  o += 16u;
  p.total = o;
  return p;
}

// ---- hidden layer: h (registers) -> relu(W h + b) (registers), weights in shared memory -------------------
template <int CPL, int AT>
__device__ __forceinline__ void pk_hidden_layer(const float* __restrict__ WT, const float* __restrict__ bias,
                                                float* hbuf, float (&h)[AT][CPL], int lane) {
  constexpr int HP = 32 * CPL;
#pragma unroll
  for (int c = 0; c < AT; ++c) pk_sts_vec<CPL>(hbuf + c * HP, lane, h[c]);
  __syncwarp();
  float acc[AT][CPL];
  {
    float bv[CPL];
    pk_lds_vec<CPL>(bias, lane, bv);
#pragma unroll
    for (int c = 0; c < AT; ++c)
#pragma unroll
      for (int i = 0; i < CPL; ++i) acc[c][i] = bv[i];
  }
#pragma unroll 2
  for (int k = 0; k < HP; k += 4) {
    float4 hv[AT];
#pragma unroll
    for (int c = 0; c < AT; ++c) hv[c] = *reinterpret_cast<const float4*>(hbuf + c * HP + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float w[CPL];
      pk_lds_vec<CPL>(WT + (k + kk) * HP, lane, w);
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        const float x = kk == 0 ? hv[c].x : kk == 1 ? hv[c].y : kk == 2 ? hv[c].z : hv[c].w;
        pk_fma_vec<CPL>(acc[c], w, x);
      }
    }
  }
  __syncwarp();  // every lane is done reading hbuf before the next layer overwrites it
#pragma unroll
  for (int c = 0; c < AT; ++c)
#pragma unroll
    for (int i = 0; i < CPL; ++i) h[c][i] = fmaxf(acc[c][i], 0.0f);
}

// register-lean variant: the layer input is already in hbuf (AT columns x HP); returns the pre-activation in acc
template <int CPL, int AT>
__device__ __forceinline__ void pk_hidden_from_smem(const float* __restrict__ WT, const float* __restrict__ bias,
                                                    const float* hbuf, float (&acc)[AT][CPL], int lane) {
  constexpr int HP = 32 * CPL;
  {
    float bv[CPL];
    pk_lds_vec<CPL>(bias, lane, bv);
#pragma unroll
    for (int c = 0; c < AT; ++c)
#pragma unroll
      for (int i = 0; i < CPL; ++i) acc[c][i] = bv[i];
  }
  if constexpr (CPL == 2) {
    // 64 channels: four k per step -- the activations of a column arrive as ONE 16-byte broadcast load instead of two
    // 8-byte ones (the loop is bound by shared-memory instructions: 208 instead of 352 per battle at A = 9)
#pragma unroll 2
    for (int k = 0; k < HP; k += 4) {
      float w0[CPL], w1[CPL], w2[CPL], w3[CPL];
      pk_lds_vec<CPL>(WT + k * HP, lane, w0);
      pk_lds_vec<CPL>(WT + (k + 1) * HP, lane, w1);
      pk_lds_vec<CPL>(WT + (k + 2) * HP, lane, w2);
      pk_lds_vec<CPL>(WT + (k + 3) * HP, lane, w3);
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        const float4 hv = *reinterpret_cast<const float4*>(hbuf + c * HP + k);
        pk_fma_vec<CPL>(acc[c], w0, hv.x);
        pk_fma_vec<CPL>(acc[c], w1, hv.y);
        pk_fma_vec<CPL>(acc[c], w2, hv.z);
        pk_fma_vec<CPL>(acc[c], w3, hv.w);
      }
    }
  } else {
#pragma unroll 2
    for (int k = 0; k < HP; k += 2) {
      float w0[CPL], w1[CPL];
      pk_lds_vec<CPL>(WT + k * HP, lane, w0);
      pk_lds_vec<CPL>(WT + (k + 1) * HP, lane, w1);
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        const float2 hv = *reinterpret_cast<const float2*>(hbuf + c * HP + k);
        pk_fma_vec<CPL>(acc[c], w0, hv.x);
        pk_fma_vec<CPL>(acc[c], w1, hv.y);
      }
    }
  }
}

// hidden layers 1..n_layers-1 and the 1-channel head; h holds relu(layer 0) on entry, qv on exit is
// the Q value of every column in every lane
template <int CPL, int AT>
__device__ __forceinline__ void pk_mlp_tail(const float* s_tail, const PkLayout& lay, float* hbuf,
                                            float (&h)[AT][CPL], float (&qv)[AT], int lane) {
  const unsigned base = lay.b[0];
  for (int l = 1; l < lay.n_layers; ++l)
    pk_hidden_layer<CPL, AT>(s_tail + (lay.wl[l] - base), s_tail + (lay.b[l] - base), hbuf, h, lane);
  float wo[CPL];
  pk_lds_vec<CPL>(s_tail + (lay.wout - base), lane, wo);
  const float bo = s_tail[lay.bout - base];
#pragma unroll
  for (int c = 0; c < AT; ++c) {
    float p = 0.0f;
#pragma unroll
    for (int i = 0; i < CPL; ++i) p = fmaf(wo[i], h[c][i], p);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
    qv[c] = p + bo;
  }
}

// cooperative prologue copy of the network's small tensors into shared memory
__device__ __forceinline__ void pk_stage_weights(unsigned char* smem, const PkSmemPlan& plan, const float* prep,
                                                 const PkLayout& lay) {
  float4* dt = reinterpret_cast<float4*>(smem + plan.tail);
  const float4* st = reinterpret_cast<const float4*>(prep + lay.b[0]);
  const int nt = (int)((lay.total - lay.b[0]) >> 2);
  for (int i = threadIdx.x; i < nt; i += blockDim.x) dt[i] = __ldg(st + i);
  float4* dw = reinterpret_cast<float4*>(smem + plan.w1s);
  const float4* sw = reinterpret_cast<const float4*>(prep + lay.w1t);
  const int nw = (PK_N_STATE * lay.HP) >> 2;
  for (int i = threadIdx.x; i < nw; i += blockDim.x) dw[i] = __ldg(sw + i);
}

// ---- feature block of one battle ---------------------------------------------------------------------
// g = start of the battle's (C, A) block (4-byte aligned only: C*A*4 is not a multiple of 16).
// sv = the 50 state values in shared memory (padded with zeros to 64).
// The ~450 state values go out as coalesced 16-byte stores from registers.  The zeros behind them
// (about 20 KB) are either stored the same way (zero_tile == nullptr) or handed to the TMA engine as one
// bulk store from a block-wide tile of zeros in shared memory; the caller pokes the ones afterwards
// (pk_poke_feature_block), for the TMA variant only after cp.async.bulk.wait_group has returned.
template <int A>
__device__ __forceinline__ void pk_store_feature_block(float* __restrict__ g, int L, const float* sv,
                                                       const float* zero_tile, int lane, uint32_t zero_tile_bytes = 0) {
  const int m = (int)((reinterpret_cast<uintptr_t>(g) >> 2) & 3u);  // phase of the block inside a 16 B line
  const int hc = (4 - m) & 3;                                       // floats before the first 16 B boundary
  const int nchunks = (L - hc) >> 2;                                // whole 16 B chunks
  constexpr int ns = PK_N_STATE * A;                                // state rows fill every column (:48)
  float4* g4 = reinterpret_cast<float4*>(g + hc);
  // chunks that overlap the state region carry values; everything behind is zeros until poked
  const int n_state_chunks = min(nchunks, (ns - hc + 3) >> 2);
  if (zero_tile != nullptr && lane == 0 && nchunks > n_state_chunks) {
    uint32_t left = (uint32_t)(nchunks - n_state_chunks) * 16u;
    if (zero_tile_bytes == 0 || zero_tile_bytes >= left) {
      pk_bulk_store(g4 + n_state_chunks, zero_tile, left);
    } else {  // a tile smaller than the zero run: several bulk stores from the same tile, one group
      unsigned char* dst = reinterpret_cast<unsigned char*>(g4 + n_state_chunks);
      while (left != 0) {
        const uint32_t n = left < zero_tile_bytes ? left : zero_tile_bytes;
        pk_bulk_store(dst, zero_tile, n);
        dst += n;
        left -= n;
      }
    }
    pk_bulk_commit();
  }
  for (int j = lane; j < n_state_chunks; j += 32) {
    const int e0 = hc + 4 * j;
    float v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = (e0 + i < ns) ? sv[(e0 + i) / A] : 0.0f;
    g4[j] = make_float4(v[0], v[1], v[2], v[3]);
  }
  if (zero_tile == nullptr) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int j = n_state_chunks + lane; j < nchunks; j += 32) g4[j] = z;
  }
  if (lane < hc) g[lane] = sv[lane / A];
  {
    const int e = hc + 4 * nchunks + lane;  // tail floats (item rows: zero unless poked)
    if (e < L) g[e] = e < ns ? sv[e / A] : 0.0f;
  }
}

// choice rows: ones in the first n_act columns only (feature_extractor.py:49).  Must run after the fill
// of the same addresses is ordered before it (__syncwarp for the register stores, wait_group for TMA).
template <int A>
__device__ __forceinline__ void pk_poke_feature_block(float* __restrict__ g, const int* idx, int zero_row, int lane) {
#pragma unroll
  for (int it = 0; it < 3; ++it) {
    const int e = lane + 32 * it;
    if (e < 8 * A) {
      const int row = idx[e];
      if (row != zero_row) g[row * A + (e >> 3)] = 1.0f;
    }
  }
}

// optional phase trace (debugging / DESIGN.md timeline): globaltimer of lane 0 at up to 16 points of the FIRST
// battle of every warp; sel.trace == nullptr in normal operation
__device__ __forceinline__ void pk_trace(unsigned long long* tr, int slot, int phase) {
  if (tr != nullptr && (threadIdx.x & 31) == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tr[(size_t)slot * 16 + phase] = t;
  }
}

template <int CPL, int A, int MAXT, bool TC = false>
__global__ void __launch_bounds__(MAXT, 1)
pk_decide_kernel(const uint4* __restrict__ packed, int B, const float* __restrict__ prep, PkLayout lay,
                 float* __restrict__ feat, int* __restrict__ mask, float* __restrict__ q,
                 int* __restrict__ action, PkSelect sel, int zero_by_tma) {
  constexpr int HP = 32 * CPL;
  constexpr int AT = A;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const bool with_mlp = prep != nullptr;
  const int L = lay.C * A;  // floats per battle
  const bool use_tma = feat != nullptr && (zero_by_tma & 1) != 0;
  static_assert(!TC || CPL == 2, "the tensor-path hidden layer is written for 64 channels");
  constexpr bool tc = TC;  // hidden layers on the warp-level tensor path (compile-time: keeps the default kernel lean)
  const int HS = tc ? PK_TC_HS : HP;                                // row stride of hbuf
  const PkSmemPlan plan = pk_smem_plan(lay, AT, nwarps, with_mlp, use_tma, tc);

  pk_grid_launch_dependents();  // the next launch of the stream may begin its own read-only prologue
  unsigned long long* tr = sel.trace;
  const int tslot = blockIdx.x * nwarps + warp;
  pk_trace(tr, tslot, 0);

  unsigned char* wa = smem + plan.warp_area + warp * plan.warp_stride;
  uint4* rec = reinterpret_cast<uint4*>(wa);
  const uint8_t* r = reinterpret_cast<const uint8_t*>(rec);
  float* sv = reinterpret_cast<float*>(wa + PK_WARP_REC_BYTES);
  int* idx = reinterpret_cast<int*>(wa + PK_WARP_REC_BYTES + PK_WARP_SV_BYTES);
  float* hbuf = reinterpret_cast<float*>(wa + PK_WARP_FIXED_BYTES);
  const float* s_tail = reinterpret_cast<const float*>(smem + plan.tail);
  const float* s_w1s = reinterpret_cast<const float*>(smem + plan.w1s);
  const float* w1t = prep + lay.w1t;
  const int zero_row = lay.C;

  // battles are dealt round-robin over every warp of the grid.  The first record is requested before the
  // prologue and parsed while the weight staging loads are in flight (nothing in the parse needs them).
  const long long stride = (long long)gridDim.x * nwarps;
  long long b = (long long)warp * gridDim.x + blockIdx.x;
  uint4 cur = make_uint4(0u, 0u, 0u, 0u);
  if (b < B && lane < 9) cur = __ldg(packed + b * 9 + lane);

  uint64_t* wbar = reinterpret_cast<uint64_t*>(smem + plan.mbar);
  if (with_mlp && threadIdx.x == 0) {
    pk_mbar_init(wbar, 1);
    pk_fence_mbar_init();
  }
  const uint4* wfrag = reinterpret_cast<const uint4*>(smem + plan.wfrag);
  const float* zero_tile = nullptr;
  if (use_tma) {
    float4* z = reinterpret_cast<float4*>(smem + plan.zero_tile);
    for (int i = threadIdx.x; i < (int)(plan.zero_bytes >> 4); i += blockDim.x) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pk_fence_proxy_async_smem();  // the async proxy (TMA) reads what the generic proxy just wrote
    zero_tile = reinterpret_cast<const float*>(z);
  }

  // record -> smem, next record requested, state values and gather rows derived (the warp's own scratch only)
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
  pk_trace(tr, tslot, 1);
  if (b < B) take_and_parse();
  pk_trace(tr, tslot, 2);
  // Programmatic dependent launch: everything above reads only the packed records and fills shared memory, so it
  // may run while the previous launch of this stream is still finishing.  Weights (which a preceding pk_clamp_adam may
  // have written) are read and every global WRITE happens below griddepcontrol.wait, which returns once the preceding
  // kernel has completed and flushed.  Both griddepcontrol instructions are no-ops for a normal launch.
  // The network's small tensors (biases, hidden matrices, head; the 50 state rows of W0^T: 30 KB for 2 x 64) are staged
  // by two TMA bulk loads issued by one thread BEFORE the prologue barrier, i.e. before any warp's feature fill: the
  // SM's TMA unit serves requests in order, and loads queued behind sixteen 20 KB fills land microseconds later
  // (measured on the tensor-core variant).  They land while the first battle's features are stored;
  // the mbarrier is waited for just before the first use of the weights.
  if (with_mlp && warp == 0) {
    pk_grid_dependency_wait();
    if (lane == 0) {
      const uint32_t tail_bytes = (lay.total - lay.b[0]) * 4u, w1s_bytes = (uint32_t)PK_N_STATE * lay.HP * 4u;
      pk_mbar_expect_tx(wbar, tail_bytes + w1s_bytes);
      pk_bulk_load(smem + plan.tail, prep + lay.b[0], tail_bytes, wbar);
      pk_bulk_load(smem + plan.w1s, prep + lay.w1t, w1s_bytes, wbar);
    }
    __syncwarp();
  }
  if (with_mlp || use_tma) __syncthreads();
  pk_grid_dependency_wait();
  if constexpr (TC) {
    if (with_mlp) {
      for (int l = 1; l < lay.n_layers; ++l)
        pk_tc_build_wfrag(prep + lay.wl[l], reinterpret_cast<uint4*>(smem + plan.wfrag) + (l - 1) * 2048);
      __syncthreads();
    }
  }
  pk_trace(tr, tslot, 3);
  float b1r[CPL];

  for (bool first = true; b < B; b += stride, first = false) {
    if (!first) {
      __syncwarp();  // the previous battle's readers of rec / sv / idx are done
      take_and_parse();
    }
    const int n_act = min((int)r[PK_OFF_NACT], A);
    if (mask != nullptr && lane < A) mask[b * A + lane] = lane < n_act ? 1 : 0;  // feature_extractor.py:50-51

    float* gfeat = feat != nullptr ? feat + (size_t)b * L : nullptr;
    if (gfeat != nullptr) {
      pk_store_feature_block<A>(gfeat, L, sv, zero_tile, lane);
      if (!use_tma) {
        __syncwarp();  // orders the fill before the pokes of other lanes (same addresses)
        pk_poke_feature_block<A>(gfeat, idx, zero_row, lane);
      } else if (!with_mlp) {
        if (lane == 0) pk_bulk_wait_all();
        __syncwarp();
        pk_poke_feature_block<A>(gfeat, idx, zero_row, lane);
      }
    }
    pk_trace(tr, tslot, 4);
    if (!with_mlp) {
      tr = nullptr;
      continue;
    }

    if (first) {
      pk_mbar_wait(wbar, 0);  // the staged weights have landed (async proxy -> visible after the wait)
      pk_lds_vec<CPL>(s_tail, lane, b1r);  // lay.b[0] is the start of the tail
    }
    constexpr int GD = CPL <= 2 ? (AT < 4 ? AT : 4) : (AT < 2 ? AT : 2);  // gather ring depth (columns in flight)
    float gr[GD][8][CPL];
#pragma unroll
    for (int c = 0; c < GD; ++c) pk_gather_column<CPL>(w1t, idx + 8 * c, zero_row, HP, lane, gr[c]);
    // ---- layer 0, state part: shared by every column of this battle (two chains for ILP) ----------
    float base[CPL], base2[CPL];
#pragma unroll
    for (int i = 0; i < CPL; ++i) {
      base[i] = b1r[i];
      base2[i] = 0.0f;
    }
#pragma unroll 3
    for (int k4 = 0; k4 < 48; k4 += 4) {  // dense and branch-free: 0 * w adds nothing, and no compare/branch per row
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
      const float2 s2 = *reinterpret_cast<const float2*>(sv + 48);  // rows 48, 49
      float w[CPL];
      pk_lds_vec<CPL>(s_w1s + 48 * HP, lane, w);
      pk_fma_vec<CPL>(base, w, s2.x);
      pk_lds_vec<CPL>(s_w1s + 49 * HP, lane, w);
      pk_fma_vec<CPL>(base2, w, s2.y);
    }
#pragma unroll
    for (int i = 0; i < CPL; ++i) base[i] += base2[i];
    pk_trace(tr, tslot, 5);
    // ---- layer 0, choice part: 8 row slots per column gathered from W1^T (absent slots and illegal columns
    //      point at the all-zero row, so there is no branch).  The gathers miss L1 about half the time, so a ring of
    //      GD columns (8 GD row loads) is kept in flight: the first GD columns were requested BEFORE the state part
    //      above, column c + GD is requested when column c is summed; relu(h) goes straight to the warp's hbuf -------
#pragma unroll
    for (int c = 0; c < AT; ++c) {
      float hc[CPL];
#pragma unroll
      for (int i = 0; i < CPL; ++i) {
        const float s01 = gr[c % GD][0][i] + gr[c % GD][1][i], s23 = gr[c % GD][2][i] + gr[c % GD][3][i];
        const float s45 = gr[c % GD][4][i] + gr[c % GD][5][i], s67 = gr[c % GD][6][i] + gr[c % GD][7][i];
        hc[i] = fmaxf(base[i] + ((s01 + s23) + (s45 + s67)), 0.0f);
      }
      pk_sts_vec<CPL>(hbuf + c * HS, lane, hc);
      if (c + GD < AT) pk_gather_column<CPL>(w1t, idx + 8 * (c + GD), zero_row, HP, lane, gr[c % GD]);
    }
    __syncwarp();
    pk_trace(tr, tslot, 6);
    // ---- hidden layers (input in hbuf) + head ----------------------------------------------------------
    float qv[AT];
    bool q_done = false;
    if constexpr (TC) {
      {
        constexpr int NT = (A + 7) / 8;
        const unsigned tb = lay.b[0];
        const int g4 = lane >> 2, t4 = lane & 3;
        float d[4][NT][4];
        for (int l = 1; l < lay.n_layers; ++l) {
          pk_hidden_tc<A>(wfrag + (l - 1) * 2048, s_tail + (lay.b[l] - tb), hbuf, d, lane);
          if (l + 1 < lay.n_layers) {  // relu -> hbuf[column][channel] for the next layer
            __syncwarp();
#pragma unroll
            for (int mt = 0; mt < 4; ++mt)
#pragma unroll
              for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  const int col = 8 * nt + 2 * t4 + (i & 1), ch = 16 * mt + g4 + ((i & 2) ? 8 : 0);
                  if (col < A) hbuf[col * PK_TC_HS + ch] = fmaxf(d[mt][nt][i], 0.0f);
                }
            __syncwarp();
          }
        }
        const float* wo = s_tail + (lay.wout - tb);
        const float bo = s_tail[lay.bout - tb];
        float p[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) p[nt][0] = p[nt][1] = 0.0f;
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          const float w0 = wo[16 * mt + g4], w1 = wo[16 * mt + g4 + 8];
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            p[nt][0] = fmaf(w0, fmaxf(d[mt][nt][0], 0.0f), p[nt][0]);
            p[nt][1] = fmaf(w0, fmaxf(d[mt][nt][1], 0.0f), p[nt][1]);
            p[nt][0] = fmaf(w1, fmaxf(d[mt][nt][2], 0.0f), p[nt][0]);
            p[nt][1] = fmaf(w1, fmaxf(d[mt][nt][3], 0.0f), p[nt][1]);
          }
        }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) p[nt][i] += __shfl_xor_sync(0xffffffffu, p[nt][i], o);
#pragma unroll
        for (int c = 0; c < AT; ++c) qv[c] = __shfl_sync(0xffffffffu, p[c >> 3][c & 1], (c & 7) >> 1) + bo;
        q_done = true;
      }
    }
    if (!q_done) {
      const unsigned tb = lay.b[0];
      float acc[AT][CPL];
      for (int l = 1; l < lay.n_layers; ++l) {
        pk_hidden_from_smem<CPL, AT>(s_tail + (lay.wl[l] - tb), s_tail + (lay.b[l] - tb), hbuf, acc, lane);
        if (l + 1 < lay.n_layers) {  // feed the next hidden layer through hbuf
          __syncwarp();
#pragma unroll
          for (int c = 0; c < AT; ++c) {
            float hc[CPL];
#pragma unroll
            for (int i = 0; i < CPL; ++i) hc[i] = fmaxf(acc[c][i], 0.0f);
            pk_sts_vec<CPL>(hbuf + c * HP, lane, hc);
          }
          __syncwarp();
        }
      }
      if (lay.n_layers == 1) {  // single layer: the head reads relu(layer 0) back from hbuf
#pragma unroll
        for (int c = 0; c < AT; ++c) pk_lds_vec<CPL>(hbuf + c * HP, lane, acc[c]);
      }
      float wo[CPL];
      pk_lds_vec<CPL>(s_tail + (lay.wout - tb), lane, wo);
      const float bo = s_tail[lay.bout - tb];
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        float p = 0.0f;
#pragma unroll
        for (int i = 0; i < CPL; ++i) p = fmaf(wo[i], fmaxf(acc[c][i], 0.0f), p);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
        qv[c] = p + bo;
      }
    }
    __syncwarp();  // hbuf is rewritten by the next battle
    pk_trace(tr, tslot, 7);
    float myq = 0.0f;
#pragma unroll
    for (int c = 0; c < AT; ++c)
      if (lane == c) myq = qv[c];
    if (q != nullptr && lane < A) q[b * A + lane] = myq;
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
    tr = nullptr;  // first battle of the warp only
  }
}

// ---- standalone selection: q (B, A), mask (B, A) -> action (B,) ------------------------------------
__global__ void __launch_bounds__(256) pk_select_kernel(const float* __restrict__ q, const int* __restrict__ mask,
                                                        int B, int A, PkSelect sel, int* __restrict__ action) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float qv[PK_MAX_ACTIONS];
  uint32_t legal = 0;
#pragma unroll
  for (int c = 0; c < PK_MAX_ACTIONS; ++c) {
    qv[c] = 0.0f;
    if (c < A) {
      qv[c] = q[b * A + c];
      if (mask[b * A + c] != 0) legal |= 1u << c;
    }
  }
  action[b] = pk_select_dev<PK_MAX_ACTIONS>(qv, legal, A, sel, b, true);
}

// ---- dense-input forward: feat (B, C, A) -> q (B, A) ------------------------------------------------
// For callers that hand in dense features (Agent._calc_q_vector[_batch], agent.py:23-30).  HBM-bound by the
// 4*C*A bytes it has to read per battle.  One warp per battle, 16 warps per CTA, persistent grid.
//  * The (C, A) block is STREAMED, not staged: chunks of 32 rows (32*A floats, contiguous) go global -> shared
//    memory with 4-byte cp.async (battle blocks are only 4-byte aligned: 4*C*A is not a multiple of 16) through a
//    per-warp ring of PK_DENSE_DEPTH chunks that runs ahead across battle boundaries, so every warp always has
//    DEPTH-1 chunks (3.4 KB at A = 9) in flight and no registers are spent on the copy.  In shared memory a row is
//    padded to XS = 4, 8 or 12 floats, so a row is read with XS/4 16-byte loads (the kernel is bound by shared-memory
//    instructions: 9 scalar loads per row were the largest share of them).
//  * Rows that are entirely zero are skipped (exact: 0 * w adds nothing); one-hot features make ~80 % of the rows zero.
//  * The rows of W0^T a non-zero feature row multiplies come from shared memory for rows < cache_rows (as many as fit
//    beside the tail; the host sizes it) and from global memory, four loads in flight, above that.  With the carve-out
//    at its maximum the L1 is tiny, and a serial chain of L2-latency row loads was what bound the first version.
// Summation order: bias, then rows in ascending order -- the same as the first version (parity tests unchanged).
#define PK_DENSE_DEPTH 4
#define PK_DENSE_XS(AT) (((AT) + 3) & ~3)
template <int XS>
__device__ __forceinline__ void pk_dense_row(const float* row, float (&v)[XS]) {
#pragma unroll
  for (int i = 0; i < XS; i += 4) {
    const float4 t = *reinterpret_cast<const float4*>(row + i);
    v[i] = t.x;
    v[i + 1] = t.y;
    v[i + 2] = t.z;
    v[i + 3] = t.w;
  }
}
__device__ __forceinline__ void pk_cp_async4(uint32_t sdst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void pk_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void pk_cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int CPL, int AT>
__global__ void __launch_bounds__(512, 1)
pk_dense_forward_kernel(const float* __restrict__ feat, int B, int A, const float* __restrict__ prep, PkLayout lay,
                        int cache_rows, float* __restrict__ q) {
  constexpr int HP = 32 * CPL;
  constexpr int D = PK_DENSE_DEPTH;
  constexpr int XS = PK_DENSE_XS(AT);
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int L = lay.C * A;
  const int CH = 32 * A;                    // floats per chunk
  const int nch = (lay.C + 31) >> 5;        // chunks per battle
  // plan: tail | wcache[cache_rows][HP] | per warp { hbuf[AT*HP], ring[D][32][XS] }
  const unsigned tail_floats = lay.total - lay.b[0];
  float* s_tail = reinterpret_cast<float*>(smem);
  float* s_wc = s_tail + tail_floats;
  float* hbuf = s_wc + (size_t)cache_rows * HP + (size_t)warp * (AT * HP + D * 32 * XS);
  float* ring = hbuf + AT * HP;
  const float* w1t = prep + lay.w1t;

  // padding words of the ring are never written by the copies: zero them once so the row scan reads zeros there
  for (int i = lane; i < D * 32 * XS; i += 32) ring[i] = 0.0f;
  int doff[AT];  // where element e = j*32 + lane of a chunk lands: row e / A, column e % A
#pragma unroll
  for (int j = 0; j < AT; ++j) {
    const int e = j * 32 + lane;
    doff[j] = ((e / A) * XS + (e % A)) * 4;  // byte offset inside a slot
  }
  const uint32_t ring_u32 = pk_smem_u32(ring);
  __syncwarp();
  // the ring starts before the weights are staged: records first, weights under their latency
  const long long stride = (long long)gridDim.x * nwarps;
  long long ib = (long long)blockIdx.x * nwarps + warp;  // battle / chunk / slot the next issue refers to
  int ich = 0, islot = 0;
  auto issue = [&]() {
    if (ib < B) {
      const int off = ich * CH;
      const float* src = feat + (size_t)ib * L + off;
      const int n = min(CH, L - off);
      const uint32_t dst = ring_u32 + islot * (32 * XS * 4);
#pragma unroll
      for (int j = 0; j < AT; ++j) {
        const int e = j * 32 + lane;
        if (e < n) pk_cp_async4(dst + doff[j], src + e);
      }
    }
    pk_cp_async_commit();
    if (++ich == nch) {
      ich = 0;
      ib += stride;
    }
    islot = islot + 1 == D ? 0 : islot + 1;
  };
#pragma unroll
  for (int i = 0; i < D - 1; ++i) issue();

  {
    float4* dt = reinterpret_cast<float4*>(s_tail);
    const float4* st = reinterpret_cast<const float4*>(prep + lay.b[0]);
    const int nt = (int)(tail_floats >> 2);
    for (int i = threadIdx.x; i < nt; i += blockDim.x) dt[i] = __ldg(st + i);
    float4* dw = reinterpret_cast<float4*>(s_wc);
    const float4* sw = reinterpret_cast<const float4*>(w1t);
    const int nw = cache_rows * (HP / 4);
    for (int i = threadIdx.x; i < nw; i += blockDim.x) dw[i] = __ldg(sw + i);
  }
  __syncthreads();
  float b1r[CPL];
  pk_lds_vec<CPL>(s_tail, lane, b1r);

  int cslot = 0;
  for (long long b = (long long)blockIdx.x * nwarps + warp; b < B; b += stride) {
    float h[AT][CPL];
#pragma unroll
    for (int c = 0; c < AT; ++c)
#pragma unroll
      for (int i = 0; i < CPL; ++i) h[c][i] = b1r[i];
    for (int ch = 0; ch < nch; ++ch) {
      __syncwarp();  // every lane is done with the slot the next issue overwrites
      issue();
      pk_cp_async_wait<D - 1>();
      __syncwarp();
      const float* xt = ring + cslot * (32 * XS);
      cslot = cslot + 1 == D ? 0 : cslot + 1;
      const int k0 = ch << 5;
      bool nz = false;
      if (k0 + lane < lay.C) {
        float xv[XS];
        pk_dense_row<XS>(xt + lane * XS, xv);
#pragma unroll
        for (int c = 0; c < AT; ++c)
          if (c < A) nz |= xv[c] != 0.0f;
      }
      unsigned rows = __ballot_sync(0xffffffffu, nz);
      while (rows) {
        const int r0 = __ffs(rows) - 1;
        rows &= rows - 1u;
        if (k0 + r0 < cache_rows) {
          float w[CPL];
          pk_lds_vec<CPL>(s_wc + (k0 + r0) * HP, lane, w);
          float xv[XS];
          pk_dense_row<XS>(xt + r0 * XS, xv);
#pragma unroll
          for (int c = 0; c < AT; ++c)
            if (c < A) pk_fma_vec<CPL>(h[c], w, xv[c]);
        } else {  // rows ascend, so everything left in this chunk is uncached too: four row loads in flight
          int r[4];
          r[0] = r0;
#pragma unroll
          for (int t = 1; t < 4; ++t) {
            r[t] = rows ? __ffs(rows) - 1 : -1;
            rows &= rows - 1u;  // 0 & anything stays 0
          }
          float w[4][CPL];
#pragma unroll
          for (int t = 0; t < 4; ++t)
            if (r[t] >= 0) pk_ldg_vec<CPL>(w1t + (size_t)(k0 + r[t]) * HP, lane, w[t]);
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            if (r[t] >= 0) {
              float xv[XS];
              pk_dense_row<XS>(xt + r[t] * XS, xv);
#pragma unroll
              for (int c = 0; c < AT; ++c)
                if (c < A) pk_fma_vec<CPL>(h[c], w[t], xv[c]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < AT; ++c)
#pragma unroll
      for (int i = 0; i < CPL; ++i) h[c][i] = fmaxf(h[c][i], 0.0f);
    float qv[AT];
    pk_mlp_tail<CPL, AT>(s_tail, lay, hbuf, h, qv, lane);
    float myq = 0.0f;
#pragma unroll
    for (int c = 0; c < AT; ++c)
      if (lane == c) myq = qv[c];
    if (lane < A) q[b * A + lane] = myq;
  }
  pk_cp_async_wait<0>();
}

// ---- weight preparation: state_dict layout -> transposed / padded ---------------------------------------
// src (rows_out, cols_in) row-major -> dst [cols_in (+ zero rows)][HP], zero padded
__global__ void pk_transpose_pad_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows_out,
                                        int cols_in, int dst_rows, int HP) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)dst_rows * HP) return;
  const int k = (int)(i / HP), j = (int)(i - (long long)k * HP);
  dst[i] = (k < cols_in && j < rows_out) ? src[(size_t)j * cols_in + k] : 0.0f;
}
// src (rows, cols) row-major -> dst [HP][HP] zero padded, same orientation
__global__ void pk_pad2d_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int HP) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)HP * HP) return;
  const int j = (int)(i / HP), k = (int)(i - (long long)j * HP);
  dst[i] = (j < rows && k < cols) ? src[(size_t)j * cols + k] : 0.0f;
}
__global__ void pk_copy_pad_kernel(const float* __restrict__ src, float* __restrict__ dst, int n, int n_pad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pad) dst[i] = i < n ? src[i] : 0.0f;
}
// src (rows, cols) row-major hidden matrix -> tensor-core operand tiles (PkLayout::tc): hi tile then lo tile,
// 64 x 64 each, zero padded, K-major SWIZZLE_128B order
__global__ void pk_tc_tiles_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 64) return;
  const int j = i >> 6, k = i & 63;
  const float w = (j < rows && k < cols) ? src[(size_t)j * cols + k] : 0.0f;
  const float hi = pk_tf32_hi(w);
  const unsigned off = pk_tc_tile_off((unsigned)j, (unsigned)k, 2048u);
  dst[off] = hi;
  dst[off + PK_TC_TILE_FLOATS] = w - hi;
}
// W0 (rows, cols) row-major -> PkLayout::tc0: per k-block of 32 input rows a 128-row tile, rows 0-63 tf32(w), rows
// 64-127 the remainders, zero padded (pk_dense_tc_kernel's B operand)
__global__ void pk_tc0_tiles_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int nkb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 64 * 32 * nkb) return;
  const int k = i % (32 * nkb), j = i / (32 * nkb);
  const float w = (j < rows && k < cols) ? src[(size_t)j * cols + k] : 0.0f;
  const float hi = pk_tf32_hi(w);
  const unsigned off = pk_tc_tile_off((unsigned)j, (unsigned)k, PK_TC_TILE_FLOATS);
  dst[off] = hi;
  dst[off + PK_TC0_LO_DELTA] = w - hi;
}
