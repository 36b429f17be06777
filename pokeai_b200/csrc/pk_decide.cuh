// The decision kernel: packed observations -> [features] [mask] -> Q values -> action.
//
// One persistent CTA per SM, warps specialised by role, no block-level barrier after the
// prologue:
//
//   writer warps  materialise the (C, A) fp32 feature block of one battle at a time in a
//                 shared-memory staging tile that is all zeros except for ~450 state values and
//                 <= 72 one-hot pokes, hand it to the TMA engine as ONE bulk store
//                 (cp.async.bulk shared -> global, 22 KB), and once the engine has read the tile
//                 un-poke it instead of re-zeroing it.  HBM-write bound: 22 212 B per battle.
//   MLP warps     never read the dense features.  Layer 1 is evaluated in gather form
//                 (b1 + sum_k W1[:,k] s_k over the non-zero state values + the <= 8 one-hot rows of
//                 each action column, read from the transposed weights through L1), the hidden
//                 layers run out of shared-memory weights with 2 channels per lane and A columns
//                 per lane in registers, the head is a warp butterfly, and the masked
//                 argmax / epsilon-greedy / softmax draw is done in registers.
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

#define PK_WARP_REC_BYTES 576   // up to 4 records of 144 B (one per staging slot)
#define PK_WARP_SV_BYTES 256    // 64 floats: the 50 state values
#define PK_WARP_IDX_BYTES 320   // 80 ints: gather rows, 8 per action column
#define PK_MAX_NBUF 4

// ---- state features (state_feature_extractor.py:113-197) ----------------------------------------
// r points at one 144-byte record in shared memory.  Python computes a/b in double and numpy stores
// it as float32; IEEE fp32 division of two integers < 2^24 gives the same bits (double rounding is
// innocuous for division when 53 >= 2*24+2), so __fdiv_rn is exact parity -- never __fdividef.
__device__ __forceinline__ float pk_state_feature(const uint8_t* r, int k) {
  if (k < PK_ST_TYPE) {  // :113-123 remaining / total
    return __fdiv_rn((float)r[PK_OFF_REMAINING + k], (float)r[PK_OFF_TOTAL + k]);
  } else if (k < PK_ST_HP) {  // :125-138 opponent active's types
    uint32_t tm = c_tab.type_mask[r[PK_OFF_OPP]];
    return (float)((tm >> (k - PK_ST_TYPE)) & 1u);
  } else if (k < PK_ST_NV) {  // :140-151 hp ratio
    const uint16_t* hc = reinterpret_cast<const uint16_t*>(r + PK_OFF_HPCUR);
    const uint16_t* hm = reinterpret_cast<const uint16_t*>(r + PK_OFF_HPMAX);
    return __fdiv_rn((float)hc[k - PK_ST_HP], (float)hm[k - PK_ST_HP]);
  } else if (k < PK_ST_RANK) {  // :153-167 non-volatile condition one-hot
    int side = (k - PK_ST_NV) / 6;
    int cond = (k - PK_ST_NV) - 6 * side;
    return r[PK_OFF_STATUS + side] == cond + 1 ? 1.0f : 0.0f;
  } else if (k < PK_ST_WEATHER) {  // :169-182 (rank + 6) / 12
    int bst = (int)(int8_t)r[PK_OFF_BOOST + (k - PK_ST_RANK)];
    return __fdiv_rn((float)(bst + 6), 12.0f);
  } else if (k < PK_N_STATE) {  // :184-197 weather one-hot
    return r[PK_OFF_WEATHER] == (k - PK_ST_WEATHER) + 1 ? 1.0f : 0.0f;
  }
  return 0.0f;
}

// ---- choice rows (choice_to_vec.py:40-61) -----------------------------------------------------
// Slot t of action column c: 0 switch flag, 1 force_switch flag, 2 species, 3 item, 4..7 moves.
// Returns the absolute feature row, or -1 when the slot is empty / out of table range / a repeat
// of an earlier move of the same action (the reference sets the element to 1, it does not add).
__device__ __forceinline__ int pk_choice_row(const uint8_t* r, int c, int t) {
  const uint8_t* a = r + PK_OFF_ACT + 12 * c;
  const uint32_t flags = a[0];
  if (t == 0) return (flags & PK_FLAG_SWITCH) ? PK_N_STATE + PK_CH_SWITCH : -1;
  if (t == 1) return (flags & PK_FLAG_FORCE) ? PK_N_STATE + PK_CH_FORCE : -1;
  if (t == 2) {
    int sp = a[1];
    return sp < c_tab.n_species ? PK_N_STATE + PK_CH_SPECIES + sp : -1;
  }
  if (t == 3) {
    int it = a[2];
    return it < c_tab.n_items ? c_tab.row_item + it : -1;
  }
  const int mi = t - 4;
  if (mi >= (int)a[3] || mi >= PK_MAX_MOVES) return -1;
  const uint16_t* mv = reinterpret_cast<const uint16_t*>(a + 4);
  const int m = mv[mi];
  if (m >= c_tab.n_moves) return -1;
  for (int e = 0; e < mi; ++e)
    if (mv[e] == m) return -1;
  return c_tab.row_move + m;
}

// write `val` at every one-hot position of the choice block; tile points at element (row 0, col 0)
__device__ __forceinline__ void pk_poke_choice(float* tile, const uint8_t* r, int A, int n_act, int lane, float val) {
  for (int e = lane; e < 8 * n_act; e += 32) {
    const int c = e >> 3, t = e & 7;
    const int row = pk_choice_row(r, c, t);
    if (row >= 0) tile[row * A + c] = val;
  }
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

// ---- shared-memory plan ---------------------------------------------------------------------------
struct PkSmemPlan {
  unsigned tail;         // prepared floats from lay.b[0] to lay.total (biases, hidden weights, head)
  unsigned w1s;          // first PK_N_STATE rows of W1^T
  unsigned warp_area;    // per-warp scratch
  unsigned warp_stride;
  unsigned staging;      // writer staging tiles
  unsigned lbuf;         // floats per staging tile
  unsigned total;
};

__host__ __device__ inline PkSmemPlan pk_smem_plan(const PkLayout& lay, int AT, int nwarps, int n_writers, int nbuf,
                                                   int A, bool with_mlp) {
  PkSmemPlan p;
  unsigned o = 0;
  p.tail = o;
  if (with_mlp) o += (lay.total - lay.b[0]) * 4u;
  p.w1s = o;
  if (with_mlp) o += (unsigned)PK_N_STATE * lay.HP * 4u;
  p.warp_area = o;
  p.warp_stride = PK_WARP_REC_BYTES + PK_WARP_SV_BYTES + PK_WARP_IDX_BYTES + (with_mlp ? (unsigned)AT * lay.HP * 4u : 0u);
  o += p.warp_stride * nwarps;
  o = (o + 127u) & ~127u;
  p.staging = o;
  p.lbuf = ((unsigned)lay.C * A + 3u + 3u) & ~3u;
  o += p.lbuf * 4u * n_writers * nbuf;
  p.total = o;
  return p;
}

// ---- hidden layer: h (registers) -> relu(W h + b) (registers), weights in shared memory -------------------
template <int CPL, int AT>
__device__ __forceinline__ void pk_hidden_layer(const float* __restrict__ WT, const float* __restrict__ bias,
                                                float* hbuf, float (&h)[AT][CPL], int lane) {
  constexpr int HP = 32 * CPL;
#pragma unroll
  for (int c = 0; c < AT; ++c)
#pragma unroll
    for (int i = 0; i < CPL; ++i) hbuf[c * HP + lane + 32 * i] = h[c][i];
  __syncwarp();
  float acc[AT][CPL];
#pragma unroll
  for (int i = 0; i < CPL; ++i) {
    const float bv = bias[lane + 32 * i];
#pragma unroll
    for (int c = 0; c < AT; ++c) acc[c][i] = bv;
  }
#pragma unroll 2
  for (int k = 0; k < HP; k += 4) {
    float4 hv[AT];
#pragma unroll
    for (int c = 0; c < AT; ++c) hv[c] = *reinterpret_cast<const float4*>(hbuf + c * HP + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float w[CPL];
#pragma unroll
      for (int i = 0; i < CPL; ++i) w[i] = WT[(k + kk) * HP + lane + 32 * i];
#pragma unroll
      for (int c = 0; c < AT; ++c) {
        const float x = kk == 0 ? hv[c].x : kk == 1 ? hv[c].y : kk == 2 ? hv[c].z : hv[c].w;
#pragma unroll
        for (int i = 0; i < CPL; ++i) acc[c][i] = fmaf(w[i], x, acc[c][i]);
      }
    }
  }
  __syncwarp();  // every lane is done reading hbuf before the next layer overwrites it
#pragma unroll
  for (int c = 0; c < AT; ++c)
#pragma unroll
    for (int i = 0; i < CPL; ++i) h[c][i] = fmaxf(acc[c][i], 0.0f);
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
#pragma unroll
  for (int i = 0; i < CPL; ++i) wo[i] = s_tail[lay.wout - base + lane + 32 * i];
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

template <int CPL, int AT>
__global__ void __launch_bounds__(512, 1)
pk_decide_kernel(const uint4* __restrict__ packed, int B, int A, const float* __restrict__ prep, PkLayout lay,
                 float* __restrict__ feat, int* __restrict__ mask, float* __restrict__ q,
                 int* __restrict__ action, PkSelect sel, int n_writers, int nbuf) {
  constexpr int HP = 32 * CPL;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int n_mlp = nwarps - n_writers;
  const bool with_mlp = prep != nullptr && n_mlp > 0;
  const int L = lay.C * A;  // floats per battle
  const PkSmemPlan plan = pk_smem_plan(lay, AT, nwarps, n_writers, nbuf, A, with_mlp);

  // ---- prologue ---------------------------------------------------------------------------------
  if (with_mlp) pk_stage_weights(smem, plan, prep, lay);
  if (feat != nullptr) {
    float4* z = reinterpret_cast<float4*>(smem + plan.staging);
    const int nz = (int)((plan.lbuf * n_writers * nbuf) >> 2);
    for (int i = threadIdx.x; i < nz; i += blockDim.x) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();

  unsigned char* wa = smem + plan.warp_area + warp * plan.warp_stride;

  if (warp < n_writers) {
    // =========================== writer warp ====================================================
    if (feat == nullptr && (mask == nullptr || with_mlp)) return;
    uint4* recs = reinterpret_cast<uint4*>(wa);
    float* stg = reinterpret_cast<float*>(smem + plan.staging) + (size_t)warp * nbuf * plan.lbuf;
    const long long reuse_stride = (long long)nbuf * n_writers * gridDim.x;
    int use = 0;
    for (int n = warp;; n += n_writers, ++use) {
      const long long b = blockIdx.x + (long long)n * gridDim.x;
      if (b >= B) break;
      const int slot = use % nbuf;
      float* tile = stg + (size_t)slot * plan.lbuf;
      uint4* rec = recs + slot * 9;
      const uint8_t* r = reinterpret_cast<const uint8_t*>(rec);
      if (feat != nullptr && use >= nbuf) {
        // the bulk store that last used this tile must have finished READING it
        if (lane == 0) {
          switch (nbuf) {
            case 1: pk_bulk_wait_read<0>(); break;
            case 2: pk_bulk_wait_read<1>(); break;
            case 3: pk_bulk_wait_read<2>(); break;
            default: pk_bulk_wait_read<3>(); break;
          }
        }
        __syncwarp();
        const int m_old = (int)(((unsigned long long)(b - reuse_stride) * (unsigned long long)L) & 3ull);
        const int n_old = min((int)r[PK_OFF_NACT], A);
        pk_poke_choice(tile + m_old, r, A, n_old, lane, 0.0f);  // un-poke: the tile is zeros again (bar state rows)
        __syncwarp();
      }
      if (lane < 9) rec[lane] = __ldg(packed + b * 9 + lane);
      __syncwarp();
      const int n_act = min((int)r[PK_OFF_NACT], A);
      if (mask != nullptr && lane < A) mask[b * A + lane] = lane < n_act ? 1 : 0;  // feature_extractor.py:50-51
      if (feat != nullptr) {
        const int m = (int)(((unsigned long long)b * (unsigned long long)L) & 3ull);  // global 16 B phase of this block
        const int ns = PK_N_STATE * A;
        // a previous use of this tile with a different phase leaves <= 3 stale state values at either edge
        if (lane < 3) {
          if (lane < m) tile[lane] = 0.0f;
          if (m + lane < 3) tile[ns + m + lane] = 0.0f;
        }
        __syncwarp();
        // state rows fill every column, legal or not (feature_extractor.py:48)
        const float s0 = pk_state_feature(r, lane);
        const float s1 = pk_state_feature(r, lane + 32);
        for (int c = 0; c < A; ++c) tile[m + lane * A + c] = s0;
        if (lane + 32 < PK_N_STATE)
          for (int c = 0; c < A; ++c) tile[m + (lane + 32) * A + c] = s1;
        // choice rows only in the first n_act columns (feature_extractor.py:49)
        pk_poke_choice(tile + m, r, A, n_act, lane, 1.0f);
        pk_fence_proxy_async_smem();
        __syncwarp();
        const int hc = (4 - m) & 3;          // floats before the first 16 B boundary
        const int body = (L - hc) & ~3;      // floats moved by the TMA engine
        const int tc = L - hc - body;        // floats after the last 16 B boundary
        float* g = feat + (size_t)b * L;
        if (lane == 0) {
          pk_bulk_store(g + hc, tile + m + hc, (uint32_t)body * 4u);
          pk_bulk_commit();
        }
        if (lane < hc) g[lane] = tile[m + lane];
        if (lane >= 8 && lane < 8 + tc) {
          const int i = hc + body + (lane - 8);
          g[i] = tile[m + i];
        }
      }
    }
    if (feat != nullptr && lane == 0) pk_bulk_wait_all();
    return;
  }

  // =============================== MLP warp =======================================================
  if (!with_mlp) return;
  const bool mask_here = mask != nullptr && feat == nullptr;  // otherwise the writer warps emit it
  const int mw = warp - n_writers;
  uint4* rec = reinterpret_cast<uint4*>(wa);
  const uint8_t* r = reinterpret_cast<const uint8_t*>(rec);
  float* sv = reinterpret_cast<float*>(wa + PK_WARP_REC_BYTES);
  int* idx = reinterpret_cast<int*>(wa + PK_WARP_REC_BYTES + PK_WARP_SV_BYTES);
  float* hbuf = reinterpret_cast<float*>(wa + PK_WARP_REC_BYTES + PK_WARP_SV_BYTES + PK_WARP_IDX_BYTES);
  const float* s_tail = reinterpret_cast<const float*>(smem + plan.tail);
  const float* s_w1s = reinterpret_cast<const float*>(smem + plan.w1s);
  const float* w1t = prep + lay.w1t;
  const int zero_row = lay.C;
  float b1r[CPL];
#pragma unroll
  for (int i = 0; i < CPL; ++i) b1r[i] = s_tail[lane + 32 * i];  // lay.b[0] is the start of the tail

  for (int n = mw;; n += n_mlp) {
    const long long b = blockIdx.x + (long long)n * gridDim.x;
    if (b >= B) break;
    __syncwarp();
    if (lane < 9) rec[lane] = __ldg(packed + b * 9 + lane);
    __syncwarp();
    const int n_act = min((int)r[PK_OFF_NACT], A);
    if (mask_here && lane < A) mask[b * A + lane] = lane < n_act ? 1 : 0;
    sv[lane] = pk_state_feature(r, lane);
    sv[lane + 32] = pk_state_feature(r, lane + 32);
    for (int e = lane; e < 8 * AT; e += 32) {
      const int c = e >> 3, t = e & 7;
      int row = c < n_act ? pk_choice_row(r, c, t) : -1;
      idx[e] = row < 0 ? zero_row : row;
    }
    __syncwarp();

    // ---- layer 0, state part: shared by every column of this battle ------------------------------
    float base[CPL];
#pragma unroll
    for (int i = 0; i < CPL; ++i) base[i] = b1r[i];
#pragma unroll 1
    for (int k4 = 0; k4 < 52; k4 += 4) {
      const float4 s4 = *reinterpret_cast<const float4*>(sv + k4);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float x = kk == 0 ? s4.x : kk == 1 ? s4.y : kk == 2 ? s4.z : s4.w;
        if (x != 0.0f) {  // warp-uniform; k >= 50 reads zeros
#pragma unroll
          for (int i = 0; i < CPL; ++i) base[i] = fmaf(s_w1s[(k4 + kk) * HP + lane + 32 * i], x, base[i]);
        }
      }
    }
    // ---- layer 0, choice part: <= 8 one-hot rows per column, gathered from W1^T -------------------
    float h[AT][CPL];
#pragma unroll
    for (int c = 0; c < AT; ++c) {
#pragma unroll
      for (int i = 0; i < CPL; ++i) h[c][i] = base[i];
      if (c < n_act) {  // warp-uniform
        const int4 ia = *reinterpret_cast<const int4*>(idx + 8 * c);
        const int4 ib = *reinterpret_cast<const int4*>(idx + 8 * c + 4);
        float g0[CPL], g1[CPL], g2[CPL];
#pragma unroll
        for (int i = 0; i < CPL; ++i) {
          g0[i] = __ldg(w1t + (size_t)ia.z * HP + lane + 32 * i);  // species
          g1[i] = __ldg(w1t + (size_t)ia.w * HP + lane + 32 * i);  // item
          g2[i] = __ldg(w1t + (size_t)ib.x * HP + lane + 32 * i);  // first move
        }
        const uint8_t* a = r + PK_OFF_ACT + 12 * c;
        const bool more = a[0] != 0 || a[3] > 1;  // switch / force flags or a move list (switch target)
        if (more) {
          float g3[CPL], g4[CPL], g5[CPL], g6[CPL], g7[CPL];
#pragma unroll
          for (int i = 0; i < CPL; ++i) {
            g3[i] = __ldg(w1t + (size_t)ia.x * HP + lane + 32 * i);
            g4[i] = __ldg(w1t + (size_t)ia.y * HP + lane + 32 * i);
            g5[i] = __ldg(w1t + (size_t)ib.y * HP + lane + 32 * i);
            g6[i] = __ldg(w1t + (size_t)ib.z * HP + lane + 32 * i);
            g7[i] = __ldg(w1t + (size_t)ib.w * HP + lane + 32 * i);
          }
#pragma unroll
          for (int i = 0; i < CPL; ++i) h[c][i] = ((((h[c][i] + g3[i]) + g4[i]) + g5[i]) + g6[i]) + g7[i];
        }
#pragma unroll
        for (int i = 0; i < CPL; ++i) h[c][i] = ((h[c][i] + g0[i]) + g1[i]) + g2[i];
      }
#pragma unroll
      for (int i = 0; i < CPL; ++i) h[c][i] = fmaxf(h[c][i], 0.0f);
    }
    // ---- hidden layers + head -----------------------------------------------------------------------
    float qv[AT];
    pk_mlp_tail<CPL, AT>(s_tail, lay, hbuf, h, qv, lane);
    float myq = 0.0f;
#pragma unroll
    for (int c = 0; c < AT; ++c)
      if (lane == c) myq = qv[c];
    if (q != nullptr && lane < A) q[b * A + lane] = myq;
    if (action != nullptr) {
      const int act = pk_select_dev<AT>(qv, (1u << n_act) - 1u, A, sel, b, lane == 0);
      if (lane == 0) action[b] = act;
    }
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
// For callers that hand in dense features (Agent._calc_q_vector[_batch], agent.py:23-30).  One warp
// per battle: the (C, A) block is staged in shared memory with coalesced loads, rows that are entirely
// zero are skipped (exact: 0 * w adds nothing), the rest accumulate rows of W1^T; then the same tail.
template <int CPL, int AT>
__global__ void __launch_bounds__(256, 1)
pk_dense_forward_kernel(const float* __restrict__ feat, int B, int A, const float* __restrict__ prep, PkLayout lay,
                        float* __restrict__ q) {
  constexpr int HP = 32 * CPL;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int L = lay.C * A;
  // plan: tail | per warp { hbuf[AT*HP], xt[L rounded to 4] }
  const unsigned tail_bytes = (lay.total - lay.b[0]) * 4u;
  {
    float4* dt = reinterpret_cast<float4*>(smem);
    const float4* st = reinterpret_cast<const float4*>(prep + lay.b[0]);
    const int nt = (int)(tail_bytes >> 4);
    for (int i = threadIdx.x; i < nt; i += blockDim.x) dt[i] = __ldg(st + i);
  }
  __syncthreads();
  const float* s_tail = reinterpret_cast<const float*>(smem);
  const unsigned xt_floats = ((unsigned)L + 3u) & ~3u;
  float* hbuf = reinterpret_cast<float*>(smem + tail_bytes) + (size_t)warp * (AT * HP + xt_floats);
  float* xt = hbuf + AT * HP;
  const float* w1t = prep + lay.w1t;
  float b1r[CPL];
#pragma unroll
  for (int i = 0; i < CPL; ++i) b1r[i] = s_tail[lane + 32 * i];

  for (long long b = (long long)blockIdx.x * nwarps + warp; b < B; b += (long long)gridDim.x * nwarps) {
    const float* g = feat + (size_t)b * L;
    __syncwarp();
#pragma unroll 8
    for (int i = lane; i < L; i += 32) xt[i] = __ldg(g + i);
    __syncwarp();
    float h[AT][CPL];
#pragma unroll
    for (int c = 0; c < AT; ++c)
#pragma unroll
      for (int i = 0; i < CPL; ++i) h[c][i] = b1r[i];
    for (int k0 = 0; k0 < lay.C; k0 += 32) {
      const int k = k0 + lane;
      bool nz = false;
      if (k < lay.C)
        for (int c = 0; c < A; ++c) nz |= xt[k * A + c] != 0.0f;
      unsigned rows = __ballot_sync(0xffffffffu, nz);
      while (rows) {
        const int kk = k0 + __ffs(rows) - 1;
        rows &= rows - 1u;
        float w[CPL];
#pragma unroll
        for (int i = 0; i < CPL; ++i) w[i] = __ldg(w1t + (size_t)kk * HP + lane + 32 * i);
#pragma unroll
        for (int c = 0; c < AT; ++c) {
          if (c < A) {
            const float x = xt[kk * A + c];
#pragma unroll
            for (int i = 0; i < CPL; ++i) h[c][i] = fmaf(w[i], x, h[c][i]);
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
__global__ void pk_copy_pad_kernel(const float* __restrict__ src, float* __restrict__ dst, int n, int n_pad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_pad) dst[i] = i < n ? src[i] : 0.0f;
}
