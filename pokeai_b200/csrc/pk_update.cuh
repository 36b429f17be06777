// DQN minibatch update kernels (reference: pokeai/ai/generic_move_model/trainer.py:152-198).
//
// Transitions arrive packed (two 144-byte records each) and are re-encoded on the fly; the dense
// (617, A) feature tensors of the reference are never built.  One update is
//
//   pk_decide(next_state, online)  -> a* = argmax over legal slots            (trainer.py:171-174)
//   pk_decide(next_state, target)  -> Q_target(s', .)                         (trainer.py:175-182)
//   pk_dqn_backward_kernel         -> Q(s)[a] (only the chosen column carries loss: gather, :163),
//                                     y = gamma v + r (:184), Huber (:187), backward (:192) with
//                                     per-CTA partial gradients
//   pk_dqn_reduce_kernel / pk_dqn_flat_kernel -> fixed-order sum over CTAs, state_dict-ordered flat gradient
//   pk_clamp_adam_kernel           -> clamp to [-1, 1] (:193-194), Adam (:195), refresh of the prepared weights
//
// pk_dqn_backward_kernel: one persistent CTA per SM, 16 warps.  Per tile of 16 transitions:
//   phase 1  warp w runs transition w forward (layer 0 in gather form, hidden layers from shared memory) and
//            backward down to dz of every layer, leaving x (sparse), h_l and dz_l in the tile buffer;
//   phase 2  all 512 threads accumulate weight gradients into accumulators they OWN: registers for the dense
//            tensors (hidden matrices, biases, head, the 50 state rows of W0), shared memory for the one-hot
//            rows of W0 (each warp owns the rows r with r % 16 == warp).  No atomics, fixed order: the
//            gradient is bit-reproducible run to run.
// If the one-hot accumulator (567 x HP floats) does not fit beside the weights, that part alone falls back to
// red.global.add.f32 on the output buffer (SPARSE_SMEM = false).
#pragma once
#include "pk_common.cuh"

#define PK_UPD_TILE 16
#define PK_UPD_THREADS 512
#define PK_UPD_FIXED 112  // floats per tile record before h/dz: rec[36] sv[64] idx[8] g[4]
#define PK_UPD_MAX_LAYERS 4

struct PkUpdPlan {
  unsigned tail, wn, w1s, tile, tile_stride, gacc, total;  // byte offsets except tile_stride (floats)
};

__host__ __device__ inline PkUpdPlan pk_upd_plan(const PkLayout& lay, bool sparse_smem) {
  PkUpdPlan p;
  unsigned o = 0;
  p.tail = o;
  o += (lay.total - lay.b[0]) * 4u;
  p.wn = o;
  o += (lay.total_train - lay.total) * 4u;
  p.w1s = o;
  o += (unsigned)PK_N_STATE * lay.HP * 4u;
  p.tile = o;
  p.tile_stride = PK_UPD_FIXED + 2u * lay.n_layers * lay.HP;
  o += p.tile_stride * PK_UPD_TILE * 4u;
  o = (o + 15u) & ~15u;
  p.gacc = o;
  if (sparse_smem) o += (unsigned)(lay.C - PK_N_STATE) * lay.HP * 4u;
  p.total = o;
  return p;
}

// acc[0..CPL) += sum_k M[k][lane's channels] * x[k]   (M and x in shared memory, HP x HP)
template <int CPL>
__device__ __forceinline__ void pk_matvec_smem(const float* M, const float* x, float (&acc)[CPL], int lane) {
  constexpr int HP = 32 * CPL;
#pragma unroll 4
  for (int k = 0; k < HP; k += 4) {
    const float4 xv = *reinterpret_cast<const float4*>(x + k);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      float w[CPL];
      pk_lds_vec<CPL>(M + (k + kk) * HP, lane, w);
      pk_fma_vec<CPL>(acc, w, kk == 0 ? xv.x : kk == 1 ? xv.y : kk == 2 ? xv.z : xv.w);
    }
  }
}

struct PkUpdArgs {
  const uint4* state;
  const uint4* next_state;
  const int* action;
  const float* reward;
  const uint8_t* terminal;
  const int* a_star;      // [n] online argmax over legal next actions (double DQN), may be null otherwise
  const float* q_target;  // [n][A] target network on next_state
  const float* v_sel;     // [n] Q_target(s', a*) already selected (fused next-state pass), or null
  int n, A, double_dqn;
  float inv_n, gamma;
};

template <int CPL, int NL, bool SPARSE_SMEM>
__global__ void __launch_bounds__(PK_UPD_THREADS, 1)
pk_dqn_backward_kernel(PkUpdArgs a, const float* __restrict__ prep, PkLayout lay, float* __restrict__ partial,
                       float* __restrict__ gsum) {
  constexpr int HP = 32 * CPL;
  constexpr int EPT = HP * HP / PK_UPD_THREADS;                                   // hidden-matrix entries per thread
  constexpr int NSG = PK_UPD_THREADS / HP;                                        // row groups (threads per channel)
  constexpr int NS = ((PK_N_STATE + NSG - 1) / NSG + (NSG <= 8 ? 3 : 0)) & (NSG <= 8 ? ~3 : ~0);  // state rows per thread
  static_assert(EPT >= 1 && HP % EPT == 0, "hidden matrix must split evenly over the CTA");
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const PkUpdPlan plan = pk_upd_plan(lay, SPARSE_SMEM);
  float* s_tail = reinterpret_cast<float*>(smem + plan.tail);
  float* s_wn = reinterpret_cast<float*>(smem + plan.wn);
  float* s_w1s = reinterpret_cast<float*>(smem + plan.w1s);
  float* s_tile = reinterpret_cast<float*>(smem + plan.tile);
  float* s_gacc = reinterpret_cast<float*>(smem + plan.gacc);
  const unsigned tbase = lay.b[0];
  const int zero_row = lay.C;

  {  // stage the small tensors; zero the one-hot accumulator
    const float4* st = reinterpret_cast<const float4*>(prep + lay.b[0]);
    float4* dt = reinterpret_cast<float4*>(s_tail);
    for (int i = tid; i < (int)((lay.total - lay.b[0]) >> 2); i += PK_UPD_THREADS) dt[i] = __ldg(st + i);
    const float4* sn = reinterpret_cast<const float4*>(prep + lay.total);
    float4* dn = reinterpret_cast<float4*>(s_wn);
    for (int i = tid; i < (int)((lay.total_train - lay.total) >> 2); i += PK_UPD_THREADS) dn[i] = __ldg(sn + i);
    const float4* sw = reinterpret_cast<const float4*>(prep + lay.w1t);
    float4* dw = reinterpret_cast<float4*>(s_w1s);
    for (int i = tid; i < (PK_N_STATE * HP) >> 2; i += PK_UPD_THREADS) dw[i] = __ldg(sw + i);
    if (SPARSE_SMEM) {
      float4* z = reinterpret_cast<float4*>(s_gacc);
      for (int i = tid; i < ((lay.C - PK_N_STATE) * HP) >> 2; i += PK_UPD_THREADS) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  __syncthreads();

  // ---- accumulators this thread owns ------------------------------------------------------------------
  float acc_w[NL > 1 ? NL - 1 : 1][EPT];
  float acc_b[NL];
  float acc_s[NS];
  float acc_wout = 0.f, acc_bout = 0.f, acc_loss = 0.f;
#pragma unroll
  for (int l = 0; l < (NL > 1 ? NL - 1 : 1); ++l)
#pragma unroll
    for (int i = 0; i < EPT; ++i) acc_w[l][i] = 0.f;
#pragma unroll
  for (int l = 0; l < NL; ++l) acc_b[l] = 0.f;
#pragma unroll
  for (int i = 0; i < NS; ++i) acc_s[i] = 0.f;
  // hidden matrices: a thread owns EPT entries.  EPT == 8 (64 channels): a 2 (k) x 4 (j) block, so that one 8-byte load of
  // h and one 16-byte load of dz feed 8 FMAs (the phase is bound by shared-memory instructions); otherwise one k and EPT
  // adjacent j.
  constexpr bool BLK2 = EPT == 8;
  constexpr int JW = BLK2 ? 4 : EPT;  // adjacent channels per owned k
  const int own_k = BLK2 ? 2 * (tid / (HP / 4)) : (tid * EPT) / HP;
  const int own_j0 = BLK2 ? 4 * (tid % (HP / 4)) : (tid * EPT) % HP;

  float* T = s_tile + warp * plan.tile_stride;  // this warp's tile record
  uint4* rec = reinterpret_cast<uint4*>(T);
  const uint8_t* r = reinterpret_cast<const uint8_t*>(T);
  float* sv = T + 36;
  int* idx = reinterpret_cast<int*>(T + 100);
  float* gl = T + 108;
  float* hb = T + PK_UPD_FIXED;            // h[l] = output of layer l (post-ReLU), l = 0..NL-1
  float* dzb = hb + NL * HP;               // dz[l] = dLoss/d(pre-activation of layer l)
  const float* w1t = prep + lay.w1t;

  for (int t0 = blockIdx.x * PK_UPD_TILE; t0 < a.n; t0 += gridDim.x * PK_UPD_TILE) {
    const int nt = min(PK_UPD_TILE, a.n - t0);
    // ================= phase 1: one transition per warp =================
    if (warp < nt) {
      const long long i = t0 + warp;
      if (lane < 9) rec[lane] = __ldg(a.state + i * 9 + lane);
      __syncwarp();
      const int act = a.action[i];
      sv[lane] = pk_state_feature(r, lane);
      sv[lane + 32] = pk_state_feature(r, lane + 32);
      if (lane < 8) {
        const int row = (act >= 0 && act < PK_MAX_ACTIONS) ? pk_choice_row(r, act, lane) : -1;
        idx[lane] = row < 0 ? zero_row : row;
      }
      __syncwarp();
      // ---- forward, chosen column only (state_action_values = model(s).gather(1, a), trainer.py:163) ----
      float z[CPL];
      pk_lds_vec<CPL>(s_tail, lane, z);  // b0
#pragma unroll 1
      for (int k4 = 0; k4 < 52; k4 += 4) {
        const float4 s4 = *reinterpret_cast<const float4*>(sv + k4);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const float x = kk == 0 ? s4.x : kk == 1 ? s4.y : kk == 2 ? s4.z : s4.w;
          if (x != 0.0f) {
            float w[CPL];
            pk_lds_vec<CPL>(s_w1s + (k4 + kk) * HP, lane, w);
            pk_fma_vec<CPL>(z, w, x);
          }
        }
      }
      {
        float gr[8][CPL];
#pragma unroll
        for (int s = 0; s < 8; ++s) pk_ldg_vec<CPL>(w1t + (size_t)idx[s] * HP, lane, gr[s]);
#pragma unroll
        for (int s = 0; s < 8; ++s)
#pragma unroll
          for (int c = 0; c < CPL; ++c) z[c] += gr[s][c];
      }
      float h[CPL];
#pragma unroll
      for (int c = 0; c < CPL; ++c) h[c] = fmaxf(z[c], 0.f);
      pk_sts_vec<CPL>(hb, lane, h);
#pragma unroll
      for (int l = 1; l < NL; ++l) {
        __syncwarp();
        pk_lds_vec<CPL>(s_tail + (lay.b[l] - tbase), lane, z);
        pk_matvec_smem<CPL>(s_tail + (lay.wl[l] - tbase), hb + (l - 1) * HP, z, lane);
#pragma unroll
        for (int c = 0; c < CPL; ++c) h[c] = fmaxf(z[c], 0.f);
        pk_sts_vec<CPL>(hb + l * HP, lane, h);
      }
      float wo[CPL];
      pk_lds_vec<CPL>(s_tail + (lay.wout - tbase), lane, wo);
      float q = 0.f;
#pragma unroll
      for (int c = 0; c < CPL; ++c) q = fmaf(wo[c], h[c], q);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      q += s_tail[lay.bout - tbase];
      // ---- TD target (trainer.py:164-184) ----
      float v = 0.f;
      if (!a.terminal[i]) {
        if (a.v_sel != nullptr) {
          v = a.v_sel[i];
        } else if (a.double_dqn) {
          v = a.q_target[i * a.A + a.a_star[i]];
        } else {
          const uint4 w1 = __ldg(a.next_state + i * 9 + 1);
          const int n_next = min((int)((w1.w >> 16) & 0xffu), a.A);
          v = -CUDART_INF_F;
          for (int c = 0; c < n_next; ++c) v = fmaxf(v, a.q_target[i * a.A + c]);
        }
      }
      const float y = __fadd_rn(__fmul_rn(v, a.gamma), a.reward[i]);
      const float d = q - y;
      const float ad = fabsf(d);
      const float li = ad < 1.0f ? 0.5f * d * d : ad - 0.5f;        // smooth_l1, beta = 1 (trainer.py:187)
      const float g = fminf(fmaxf(d, -1.0f), 1.0f) * a.inv_n;        // d(mean Huber)/dq
      if (lane == 0) {
        gl[0] = g;
        gl[1] = li * a.inv_n;
      }
      // ---- backward to the pre-activations of every layer ----
      float dz[CPL];
#pragma unroll
      for (int c = 0; c < CPL; ++c) dz[c] = h[c] > 0.f ? g * wo[c] : 0.f;
      pk_sts_vec<CPL>(dzb + (NL - 1) * HP, lane, dz);
#pragma unroll
      for (int l = NL - 1; l >= 1; --l) {
        __syncwarp();
        float dh[CPL], hp[CPL];
#pragma unroll
        for (int c = 0; c < CPL; ++c) dh[c] = 0.f;
        pk_matvec_smem<CPL>(s_wn + (lay.wn[l] - lay.total), dzb + l * HP, dh, lane);  // dh = W_l^T dz_l
        pk_lds_vec<CPL>(hb + (l - 1) * HP, lane, hp);
#pragma unroll
        for (int c = 0; c < CPL; ++c) dz[c] = hp[c] > 0.f ? dh[c] : 0.f;
        pk_sts_vec<CPL>(dzb + (l - 1) * HP, lane, dz);
      }
    }
    __syncthreads();
    // ================= phase 2: owned accumulation over the tile =================
    for (int t = 0; t < nt; ++t) {
      const float* R = s_tile + t * plan.tile_stride;
      const float* Rh = R + PK_UPD_FIXED;
      const float* Rdz = Rh + NL * HP;
      if (tid < HP) {
        acc_wout = fmaf(R[108], Rh[(NL - 1) * HP + tid], acc_wout);
#pragma unroll
        for (int l = 0; l < NL; ++l) acc_b[l] += Rdz[l * HP + tid];
      }
      if (tid == 0) {
        acc_bout += R[108];
        acc_loss += R[109];
      }
#pragma unroll
      for (int l = 1; l < NL; ++l) {
        if constexpr (BLK2) {
          const float2 h2 = *reinterpret_cast<const float2*>(Rh + (l - 1) * HP + own_k);
          const float4 d4 = *reinterpret_cast<const float4*>(Rdz + l * HP + own_j0);
          acc_w[l - 1][0] = fmaf(h2.x, d4.x, acc_w[l - 1][0]);
          acc_w[l - 1][1] = fmaf(h2.x, d4.y, acc_w[l - 1][1]);
          acc_w[l - 1][2] = fmaf(h2.x, d4.z, acc_w[l - 1][2]);
          acc_w[l - 1][3] = fmaf(h2.x, d4.w, acc_w[l - 1][3]);
          acc_w[l - 1][4] = fmaf(h2.y, d4.x, acc_w[l - 1][4]);
          acc_w[l - 1][5] = fmaf(h2.y, d4.y, acc_w[l - 1][5]);
          acc_w[l - 1][6] = fmaf(h2.y, d4.z, acc_w[l - 1][6]);
          acc_w[l - 1][7] = fmaf(h2.y, d4.w, acc_w[l - 1][7]);
        } else {
          const float hk = Rh[(l - 1) * HP + own_k];
          const float* dzl = Rdz + l * HP + own_j0;  // EPT adjacent channels: 16-byte loads when EPT allows
          if constexpr (EPT % 4 == 0) {
#pragma unroll
            for (int i = 0; i < EPT; i += 4) {
              const float4 d4 = *reinterpret_cast<const float4*>(dzl + i);
              acc_w[l - 1][i] = fmaf(hk, d4.x, acc_w[l - 1][i]);
              acc_w[l - 1][i + 1] = fmaf(hk, d4.y, acc_w[l - 1][i + 1]);
              acc_w[l - 1][i + 2] = fmaf(hk, d4.z, acc_w[l - 1][i + 2]);
              acc_w[l - 1][i + 3] = fmaf(hk, d4.w, acc_w[l - 1][i + 3]);
            }
          } else {
#pragma unroll
            for (int i = 0; i < EPT; ++i) acc_w[l - 1][i] = fmaf(hk, dzl[i], acc_w[l - 1][i]);
          }
        }
      }
      {  // state rows of W0: a thread owns one channel (tid % HP) and NS ADJACENT rows starting at NS * (tid / HP), so the
         // state values arrive as 16-byte broadcast loads (rows >= 50 are the zero padding of sv and are dropped at the end)
        static_assert(PK_UPD_THREADS % HP == 0, "state-row ownership assumes HP divides the CTA size");
        static_assert(NS * (PK_UPD_THREADS / HP) <= 64, "sv is padded to 64 values");
        const float dzj = Rdz[tid % HP];
        const float* svk = R + 36 + NS * (tid / HP);
        if constexpr (NS % 4 == 0) {
#pragma unroll
          for (int i = 0; i < NS; i += 4) {
            const float4 s4 = *reinterpret_cast<const float4*>(svk + i);
            acc_s[i] = fmaf(s4.x, dzj, acc_s[i]);
            acc_s[i + 1] = fmaf(s4.y, dzj, acc_s[i + 1]);
            acc_s[i + 2] = fmaf(s4.z, dzj, acc_s[i + 2]);
            acc_s[i + 3] = fmaf(s4.w, dzj, acc_s[i + 3]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < NS; ++i) acc_s[i] = fmaf(svk[i], dzj, acc_s[i]);
        }
      }
    }
    // one-hot rows of W0: warp w owns rows with (row & 15) == w; pairs (t, slot) scanned 32 at a time
    for (int p0 = 0; p0 < nt * 8; p0 += 32) {
      const int p = p0 + lane;
      int row = zero_row;
      if (p < nt * 8) row = reinterpret_cast<const int*>(s_tile + (p >> 3) * plan.tile_stride + 100)[p & 7];
      unsigned mine = __ballot_sync(0xffffffffu, row != zero_row && (row & 15) == warp);
      while (mine) {
        const int src = __ffs(mine) - 1;
        mine &= mine - 1u;
        const int rr = __shfl_sync(0xffffffffu, row, src);
        const float* Rdz0 = s_tile + ((p0 + src) >> 3) * plan.tile_stride + PK_UPD_FIXED + NL * HP;
        float dz[CPL];
        pk_lds_vec<CPL>(Rdz0, lane, dz);
        if (SPARSE_SMEM) {
          float* dst = s_gacc + (rr - PK_N_STATE) * HP;
          float cur[CPL];
          pk_lds_vec<CPL>(dst, lane, cur);
#pragma unroll
          for (int c = 0; c < CPL; ++c) cur[c] += dz[c];
          pk_sts_vec<CPL>(dst, lane, cur);
        } else {
#pragma unroll
          for (int c = 0; c < CPL; ++c) atomicAdd(gsum + lay.w1t + (size_t)rr * HP + CPL * lane + c, dz[c]);
        }
      }
    }
    __syncthreads();  // the tile buffer is rewritten by the next tile's phase 1
  }

  // ---- this CTA's partial gradient, prepared layout (every position of [0, lay.total) written once) ----
  float* out = partial + (size_t)blockIdx.x * lay.total;
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    const int k = NS * (tid / HP) + i;
    if (k < PK_N_STATE) out[lay.w1t + k * HP + (tid % HP)] = acc_s[i];
  }
  for (int e = tid; e < (lay.C - PK_N_STATE) * HP; e += PK_UPD_THREADS)
    out[lay.w1t + PK_N_STATE * HP + e] = SPARSE_SMEM ? s_gacc[e] : 0.f;
  if (tid < HP) {
    out[lay.w1t + (size_t)lay.C * HP + tid] = 0.f;  // the zero row stays zero
    out[lay.wout + tid] = acc_wout;
#pragma unroll
    for (int l = 0; l < NL; ++l) out[lay.b[l] + tid] = acc_b[l];
  }
#pragma unroll
  for (int l = 1; l < NL; ++l)
#pragma unroll
    for (int i = 0; i < EPT; ++i) out[lay.wl[l] + (own_k + i / JW) * HP + own_j0 + i % JW] = acc_w[l - 1][i];
  if (tid == 0) {  // bout[0] = d/d(output bias), bout[1] = this CTA's share of the loss
    out[lay.bout + 0] = acc_bout;
    out[lay.bout + 1] = acc_loss;
    out[lay.bout + 2] = 0.f;
    out[lay.bout + 3] = 0.f;
  }
}

// gsum[p] (+)= sum over CTAs of partial[cta][p], in CTA order (deterministic)
__global__ void pk_dqn_reduce_kernel(const float* __restrict__ partial, int n_cta, unsigned total, float* __restrict__ gsum,
                                     int accumulate) {
  const unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= total) return;
  float s = accumulate ? gsum[p] : 0.f;
  for (int c = 0; c < n_cta; ++c) s += partial[(size_t)c * total + p];
  gsum[p] = s;
}

// the two kernels above in one launch.  One thread per PREPARED position p (so that the 148 partial vectors are read
// with coalesced loads: the flat order is the transpose of the prepared one): s = (gsum[p] if accumulate) + sum over
// CTAs of partial[cta][p], in CTA order (deterministic), written to the flat state_dict-ordered gradient at the
// parameter p belongs to (padding positions are dropped); the loss shares ride at bout + 1
// Launch: grid = ceil(lay.total / 32), 256 threads.  Thread (px, gy) adds the partials of CTAs gy, gy + 8, ... for
// position 32 blockIdx + px (a warp reads 128 contiguous bytes of one partial vector); the eight sums of a position are
// combined through shared memory in a fixed order.  (One thread per position with four loads in flight was latency
// bound: 15 us for 148 partials of 176 KB that sit in L2.)
#define PK_REDUCE_GROUPS 8
__global__ void __launch_bounds__(32 * PK_REDUCE_GROUPS)
pk_dqn_reduce_flat_kernel(const float* __restrict__ partial, int n_cta, PkParamLayout P, PkLayout lay,
                          const float* __restrict__ gsum, int accumulate, float* __restrict__ grad,
                          float* __restrict__ loss) {
  __shared__ float sh[PK_REDUCE_GROUPS][32];
  const int px = threadIdx.x & 31, gy = threadIdx.x >> 5;
  const unsigned p = blockIdx.x * 32u + (unsigned)px;
  float s0 = 0.f, s1 = 0.f;
  if (p < lay.total) {
    const float* src = partial + p;
    int c = gy;
    for (; c + PK_REDUCE_GROUPS < n_cta; c += 2 * PK_REDUCE_GROUPS) {
      const float v0 = src[(size_t)c * lay.total], v1 = src[(size_t)(c + PK_REDUCE_GROUPS) * lay.total];
      s0 += v0;
      s1 += v1;
    }
    if (c < n_cta) s0 += src[(size_t)c * lay.total];
  }
  sh[gy][px] = s0 + s1;
  __syncthreads();
  if (gy != 0 || p >= lay.total) return;
  const bool is_loss = p == lay.bout + 1;
  const int i = pk_prepared_to_flat(P, lay, p);
  if (i < 0 && !is_loss) return;
  float s = ((sh[0][px] + sh[1][px]) + (sh[2][px] + sh[3][px])) + ((sh[4][px] + sh[5][px]) + (sh[6][px] + sh[7][px]));
  if (accumulate && !is_loss) s += gsum[p];
  if (is_loss) {
    if (loss != nullptr) *loss = s;
  } else {
    grad[i] = s;
  }
}

// prepared-layout sums -> flat state_dict-ordered gradient (+ the loss)
__global__ void pk_dqn_flat_kernel(const float* __restrict__ gsum, PkParamLayout P, PkLayout lay, float* __restrict__ grad,
                                   float* __restrict__ loss) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0 && loss != nullptr) *loss = gsum[lay.bout + 1];
  if (i >= P.total) return;
  unsigned wn;
  grad[i] = gsum[pk_flat_to_prepared(P, lay, i, &wn)];
}

// clamp(grad) -> Adam (torch.optim.Adam defaults: no weight decay, no amsgrad) -> params, and the prepared copies
struct PkAdam {
  float one_minus_b1, b2, one_minus_b2, step_size, bc2_sqrt, eps;
};
// The same factors from the device-resident control block (pk_update_ctl, include/pokeai_b200.h) instead of launch
// arguments, so that a captured CUDA graph of the update step can be replayed: epoch and step advance on the device.
// torch/optim/adam.py computes the bias corrections in double on the host; thread 0 of every block does the same here.
__device__ __forceinline__ void pk_adam_from_ctl(const pk_update_ctl* ctl, PkAdam* h, unsigned* epoch) {
  const double b1 = (double)ctl->beta1, b2 = (double)ctl->beta2;
  const double bc1 = 1.0 - pow(b1, (double)ctl->step), bc2 = 1.0 - pow(b2, (double)ctl->step);
  h->one_minus_b1 = (float)(1.0 - b1);
  h->b2 = ctl->beta2;
  h->one_minus_b2 = (float)(1.0 - b2);
  h->step_size = (float)((double)ctl->lr / bc1);
  h->bc2_sqrt = (float)sqrt(bc2);
  h->eps = ctl->eps;
  *epoch = ctl->epoch;
}
__global__ void pk_update_ctl_advance_kernel(pk_update_ctl* ctl) {
  ctl->epoch += 1u;
  ctl->step += 1;
}
// ---- gradient exchange over NVLink peer memory, fused with clamp + Adam ----------------------------------------------
// The data-parallel update has one exchange: the sum over ranks of the flat gradient, followed by the
// clamp and the Adam step on every rank.  NCCL does it as allreduce -> separate kernel; for the default network the
// vector is 175 KB, so the step is pure latency.  Here ONE kernel does both: every rank's gradient lives in a buffer
// that all peers have mapped (torch symmetric memory = cuMem handles exchanged once; NVSwitch gives every GPU a direct
// path to every peer), a flag per rank says "my gradient of epoch e is complete", and each thread loads its element
// from all W peers IN RANK ORDER, sums, clamps and applies Adam.  The fixed order makes the reduced gradient and the
// updated weights bit-identical on every rank (NCCL's ring / tree order is not specified).
// Protocol: gradients are double-buffered by epoch parity.  Publishing epoch e+1 happens after this rank's epoch-e
// kernel finished (stream order), i.e. after it stopped reading the peers' parity-e buffers, so a peer that has seen
// every flag reach e+1 may overwrite its parity-e buffer for epoch e+2: one flag round per step, no second barrier.
// A rank that never arrives would spin the others for ever: the wait gives up after ~2 s and raises status[0].
__device__ __forceinline__ unsigned long long pk_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
struct PkPeers {
  const float* grad[PK_MAX_PEERS];  // every rank's gradient buffer of this epoch's parity, (n + 1) floats: gradient, loss share
  unsigned* flags[PK_MAX_PEERS];    // every rank's flag words; flags[d][r] = last epoch rank r published to rank d
  int rank, world;
  unsigned epoch;
  unsigned long long timeout_ns;    // how long a CTA waits for the peers' flags before it gives up (PK_PEER_TIMEOUT_MS)
};
__device__ __forceinline__ unsigned pk_ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float pk_ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__global__ void pk_peer_clamp_adam_kernel(float* __restrict__ params, PkPeers pe, float* __restrict__ m,
                                          float* __restrict__ v, PkParamLayout P, PkLayout lay, PkAdam h,
                                          float* __restrict__ prepared, float* __restrict__ grad_sum,
                                          float* __restrict__ loss_status, const pk_update_ctl* __restrict__ ctl) {
  if (ctl != nullptr) {  // graph-replayable form: epoch and Adam factors come from device memory
    __shared__ PkAdam s_h;
    __shared__ unsigned s_epoch;
    if (threadIdx.x == 0) pk_adam_from_ctl(ctl, &s_h, &s_epoch);
    __syncthreads();
    h = s_h;
    pe.epoch = s_epoch;
  }
  if (blockIdx.x == 0 && threadIdx.x < pe.world) {  // publish: this rank's gradient was written by earlier kernels
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pe.flags[threadIdx.x] + pe.rank), "r"(pe.epoch) : "memory");
  }
  // Every CTA waits on this rank's own flag words (local memory).  A rank that never arrives must not hang the GPU
  // for ever: after `timeout_ns` the CTA gives up -- it raises loss_status[1] and returns WITHOUT touching params, m,
  // v or prepared (the decision is shared through shared memory, so a CTA is all-or-nothing).  The host checks the
  // flag after every step (Trainer raises); a step with loss_status[1] != 0 must be treated as not taken.
  __shared__ int s_timed_out;
  if (threadIdx.x == 0) s_timed_out = 0;
  __syncthreads();
  if (threadIdx.x < pe.world) {
    const unsigned* f = pe.flags[pe.rank] + threadIdx.x;
    const unsigned long long t0 = pk_globaltimer();
    while ((int)(pk_ld_acquire_sys(f) - pe.epoch) < 0) {
      if (pk_globaltimer() - t0 > pe.timeout_ns) {
        s_timed_out = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (s_timed_out) {
    if (threadIdx.x == 0 && loss_status != nullptr) loss_status[1] = 1.0f;
    return;
  }
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > P.total) return;
  float gs = 0.0f;
  for (int r = 0; r < pe.world; ++r) gs += pk_ld_relaxed_sys(pe.grad[r] + i);  // rank order: same bits everywhere
  if (i == P.total) {  // the loss shares ride in the last slot
    if (loss_status != nullptr) loss_status[0] = gs;
    return;
  }
  if (grad_sum != nullptr) grad_sum[i] = gs;
  const float g = fminf(fmaxf(gs, -1.0f), 1.0f);                  // trainer.py:193-194, after the reduce
  const float mi = m[i] + (g - m[i]) * h.one_minus_b1;
  const float vi = v[i] * h.b2 + (h.one_minus_b2 * g) * g;
  const float denom = sqrtf(vi) / h.bc2_sqrt + h.eps;
  const float p = params[i] + (-h.step_size * mi) / denom;
  m[i] = mi;
  v[i] = vi;
  params[i] = p;
  if (prepared != nullptr) {
    unsigned wn, tc;
    prepared[pk_flat_to_prepared(P, lay, i, &wn, &tc)] = p;
    if (wn) prepared[wn] = p;
    if (tc) {  // tensor-core operand tiles (pk_decide_tc_kernel, pk_dense_tc_kernel): hi = tf32(p), lo = p - hi
      const float hi = pk_tf32_hi(p);
      prepared[tc] = hi;
      prepared[tc + pk_tc_lo_delta(lay, tc)] = p - hi;
    }
  }
}

// ---- sum of a LARGE gradient over the ranks of one NVSwitch box, through peer memory (wide network: 55 MB) --------
// pk_peer_clamp_adam_kernel reads every element from every peer on every rank: right for the 175 KB gradient of the
// default network, W times too much traffic for the wide one.  Here the vector is cut into W slices:
//   phase A  wait for every rank's "gradient of epoch e written" flag; rank r sums slice r over the W buffers IN RANK
//            ORDER (16-byte system-scope loads over NVLink) and writes the sums back into ITS buffer's slice r
//            (nobody else reads that slice of that buffer in this phase);
//   flag     the last CTA of the grid to finish phase A (a counter that grows by gridDim.x per call) publishes
//            "slice r of epoch e reduced" to every peer;
//   phase B  wait for every rank's "reduced" flag and copy slice p from rank p's buffer for every p != r.
// Afterwards every rank's buffer holds the same bits: each element was summed by exactly one rank, in rank order.
// Buffers alternate by epoch parity exactly as for pk_peer_clamp_adam (a buffer is rewritten at epoch e + 2, when
// every rank has left the kernel of epoch e: they all published "written e + 1" after it).  Traffic per rank:
// 2 (W - 1) / W of the vector over NVLink instead of NCCL's ring / tree passes; one launch, no second barrier.
// Flag words: flags[d][r] = "written", flags[d][8 + r] = "reduced", flags[rank][16] = the phase-A counter.
// Time-out as above: a CTA that gives up raises status[1] and returns; the exchange is then unusable (the counter is
// off) and the host must stop (Trainer raises).
__device__ __forceinline__ float4 pk_ld_relaxed_sys_v4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool pk_peer_wait_flags(const unsigned* f, int world, unsigned epoch, unsigned long long timeout_ns,
                                                   int* s_timed_out) {
  if (threadIdx.x == 0) *s_timed_out = 0;
  __syncthreads();
  if (threadIdx.x < (unsigned)world) {
    const unsigned long long t0 = pk_globaltimer();
    while ((int)(pk_ld_acquire_sys(f + threadIdx.x) - epoch) < 0) {
      if (pk_globaltimer() - t0 > timeout_ns) {
        *s_timed_out = 1;
        break;
      }
    }
  }
  __syncthreads();
  return *s_timed_out == 0;
}
__global__ void __launch_bounds__(512) pk_peer_allreduce_kernel(PkPeers pe, unsigned n4, float* __restrict__ status) {
  __shared__ int s_timed_out;
  unsigned* my_flags = pe.flags[pe.rank];
  if (blockIdx.x == 0 && threadIdx.x < pe.world) {  // publish: this rank's gradient was written by earlier kernels
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pe.flags[threadIdx.x] + pe.rank), "r"(pe.epoch) : "memory");
  }
  if (!pk_peer_wait_flags(my_flags, pe.world, pe.epoch, pe.timeout_ns, &s_timed_out)) {
    if (threadIdx.x == 0 && status != nullptr) status[1] = 1.0f;
    return;
  }
  const unsigned per = (n4 + pe.world - 1) / pe.world;
  float4* mine = reinterpret_cast<float4*>(const_cast<float*>(pe.grad[pe.rank]));
  {  // phase A: my slice, summed over the ranks in rank order
    const unsigned lo = per * pe.rank, hi = min(n4, lo + per);
    for (unsigned i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) {
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r = 0; r < pe.world; ++r) {
        const float4 v = pk_ld_relaxed_sys_v4(reinterpret_cast<const float4*>(pe.grad[r]) + i);
        s.x += v.x;
        s.y += v.y;
        s.z += v.z;
        s.w += v.w;
      }
      mine[i] = s;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned old = atomicAdd(my_flags + 16, 1u);
    if (old + 1u == pe.epoch * gridDim.x) {  // the last CTA of this call: the whole slice is in memory
      __threadfence_system();
      for (int d = 0; d < pe.world; ++d)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pe.flags[d] + 8 + pe.rank), "r"(pe.epoch) : "memory");
    }
  }
  if (!pk_peer_wait_flags(my_flags + 8, pe.world, pe.epoch, pe.timeout_ns, &s_timed_out)) {
    if (threadIdx.x == 0 && status != nullptr) status[1] = 1.0f;
    return;
  }
  for (int k = 1; k < pe.world; ++k) {  // phase B: the other ranks' reduced slices, starting with the next rank
    const int p = (pe.rank + k) % pe.world;
    const unsigned lo = per * p, hi = min(n4, lo + per);
    const float4* src = reinterpret_cast<const float4*>(pe.grad[p]);
    for (unsigned i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x)
      mine[i] = pk_ld_relaxed_sys_v4(src + i);
  }
}

__global__ void pk_clamp_adam_kernel(float* __restrict__ params, const float* __restrict__ grad, float* __restrict__ m,
                                     float* __restrict__ v, PkParamLayout P, PkLayout lay, PkAdam h,
                                     float* __restrict__ prepared, const pk_update_ctl* __restrict__ ctl) {
  if (ctl != nullptr) {
    __shared__ PkAdam s_h;
    __shared__ unsigned s_epoch;
    if (threadIdx.x == 0) pk_adam_from_ctl(ctl, &s_h, &s_epoch);
    __syncthreads();
    h = s_h;
  }
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.total) return;
  const float g = fminf(fmaxf(grad[i], -1.0f), 1.0f);            // trainer.py:193-194
  const float mi = m[i] + (g - m[i]) * h.one_minus_b1;            // exp_avg.lerp_(grad, 1 - beta1)
  const float vi = v[i] * h.b2 + (h.one_minus_b2 * g) * g;        // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
  const float denom = sqrtf(vi) / h.bc2_sqrt + h.eps;
  const float p = params[i] + (-h.step_size * mi) / denom;        // param.addcdiv_(exp_avg, denom, value=-step_size)
  m[i] = mi;
  v[i] = vi;
  params[i] = p;
  if (prepared != nullptr) {
    unsigned wn, tc;
    prepared[pk_flat_to_prepared(P, lay, i, &wn, &tc)] = p;
    if (wn) prepared[wn] = p;
    if (tc) {  // tensor-core operand tiles (pk_decide_tc_kernel, pk_dense_tc_kernel): hi = tf32(p), lo = p - hi
      const float hi = pk_tf32_hi(p);
      prepared[tc] = hi;
      prepared[tc + pk_tc_lo_delta(lay, tc)] = p - hi;
    }
  }
}
