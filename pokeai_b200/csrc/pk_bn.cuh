// Train-mode BatchNorm update (MLPModel(bn=True), pokeai/ai/generic_move_model/mlp_model.py:12-15,24-26 under
// Trainer._update, trainer.py:152-195).  The reference never calls .eval() on Trainer.model / target_model, so with
// bn=True all three forwards of the update normalise with the statistics of the batch they are given (over batch x action
// columns, per channel), move the running statistics, and the gradient flows through the batch statistics.
//
// Every recorded run of the reference uses bn=False: this path is built for coverage and exactness, not for speed.  It is
// a chain of plain FP32 kernels over activations kept in HBM as [column][channel] (column = (transition, action)):
//   pk_bn_gemm_kernel      strided FP32 GEMM, 64x64 tiles, optional split-K with atomics   (conv1d k=1, dW, dH)
//   pk_bn_stats_kernel     per-channel sum and sum of squares in DOUBLE (atomicAdd double)
//   pk_bn_finalize_kernel  mean, 1/sqrt(var + eps), running statistics (momentum 0.1, unbiased variance)
//   pk_bn_apply_kernel     y = relu(gamma * (z - mean) * rstd + beta)
//   pk_bn_head_kernel      q = w_out . y + b_out
//   pk_bn_dq_kernel        Double-DQN target, Huber loss, dq (one CTA, fixed summation order)
//   pk_bn_bwd_reduce_kernel / pk_bn_bwd_apply_kernel   BatchNorm backward (dgamma, dbeta, dz)
//   pk_clamp_adam_flat_kernel   clamp + Adam over a plain vector (same arithmetic as pk_clamp_adam_kernel)
#pragma once
#include "pk_common.cuh"
#include "pk_decide.cuh"  // pk_state_feature, pk_choice_row
#include "pk_update.cuh"  // PkAdam

// flat parameter order of a bn=True MLPModel (nn.Module.parameters(): layers.*.weight, then bn_layers.*.weight / .bias
// interleaved per layer, then output.weight, output.bias); conv layers carry no bias (mlp_model.py:12)
struct PkBnLayout {
  unsigned w[PK_MAX_LAYERS], gamma[PK_MAX_LAYERS], beta[PK_MAX_LAYERS];
  unsigned wout, bout, total;
};
static inline PkBnLayout pk_bn_make_layout(int nl, int H, int C) {
  PkBnLayout L;
  unsigned off = 0;
  for (int l = 0; l < nl; ++l) {
    L.w[l] = off;
    off += (unsigned)H * (unsigned)(l == 0 ? C : H);
  }
  for (int l = 0; l < nl; ++l) {
    L.gamma[l] = off;
    off += (unsigned)H;
    L.beta[l] = off;
    off += (unsigned)H;
  }
  L.wout = off;
  off += (unsigned)H;
  L.bout = off;
  off += 1u;
  L.total = off;
  return L;
}

// workspace of one forward (floats): [acc: nl * 2H doubles][mean, rstd: nl * 2H floats][z_l, y_l: nl * 2 * N * H]
struct PkBnWs {
  size_t acc, stat, act, total;
};
static inline PkBnWs pk_bn_ws_layout(int nl, int H, size_t N) {
  PkBnWs w;
  w.acc = 0;
  w.stat = (size_t)nl * 4 * H;
  w.act = w.stat + (size_t)nl * 2 * H;
  w.act = (w.act + 3) & ~(size_t)3;
  w.total = w.act + (size_t)nl * 2 * N * H;
  return w;
}

// operand of the strided GEMM: element(r, k) = p[(r / rdiv) * r_hi + (r % rdiv) * r_lo + (k / kdiv) * k_hi + (k % kdiv) * k_lo]
struct PkBnOp {
  const float* p;
  long long r_hi, r_lo, k_hi, k_lo;
  int rdiv, kdiv, k_fast;
};
__device__ __forceinline__ float pk_bn_op_at(const PkBnOp& o, int r, int k) {
  if (o.rdiv == 1 && o.kdiv == 1) return o.p[(long long)r * o.r_hi + (long long)k * o.k_hi];  // plain strided matrix
  return o.p[(long long)(r / o.rdiv) * o.r_hi + (long long)(r % o.rdiv) * o.r_lo + (long long)(k / o.kdiv) * o.k_hi +
             (long long)(k % o.kdiv) * o.k_lo];
}

#define PK_BN_TILE 64
#define PK_BN_BK 16
// out[m * ldc + n] (=|+=) sum_k A(m, k) * B(n, k); gridDim.z > 1 splits K and accumulates with atomicAdd (out zeroed)
__global__ void __launch_bounds__(256) pk_bn_gemm_kernel(PkBnOp a, PkBnOp b, float* __restrict__ out, int M, int Nn, int K,
                                                         int ldc, int kchunk) {
  __shared__ float As[PK_BN_BK][PK_BN_TILE + 1], Bs[PK_BN_BK][PK_BN_TILE + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * PK_BN_TILE, n0 = blockIdx.x * PK_BN_TILE;
  const int k_lo = blockIdx.z * kchunk, k_hi = min(K, k_lo + kchunk);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
  for (int k0 = k_lo; k0 < k_hi; k0 += PK_BN_BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      {
        const int r = a.k_fast ? e / PK_BN_BK : e % PK_BN_TILE, k = a.k_fast ? e % PK_BN_BK : e / PK_BN_TILE;
        As[k][r] = (m0 + r < M && k0 + k < k_hi) ? pk_bn_op_at(a, m0 + r, k0 + k) : 0.0f;
      }
      {
        const int r = b.k_fast ? e / PK_BN_BK : e % PK_BN_TILE, k = b.k_fast ? e % PK_BN_BK : e / PK_BN_TILE;
        Bs[k][r] = (n0 + r < Nn && k0 + k < k_hi) ? pk_bn_op_at(b, n0 + r, k0 + k) : 0.0f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PK_BN_BK; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[k][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (m < M && n < Nn) {
        if (gridDim.z > 1)
          atomicAdd(out + (size_t)m * ldc + n, acc[i][j]);
        else
          out[(size_t)m * ldc + n] = acc[i][j];
      }
    }
}

// per-channel sum and sum of squares of z[N][H] in double: block (32 channels, 8 row lanes), grid (row chunks, channel chunks)
#define PK_BN_ROWS_PER_BLOCK 256
__global__ void __launch_bounds__(256) pk_bn_stats_kernel(const float* __restrict__ z, int N, int H, double* __restrict__ acc) {
  __shared__ double s1[8][32], s2[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int o = blockIdx.y * 32 + tx;
  const int r0 = blockIdx.x * PK_BN_ROWS_PER_BLOCK, r1 = min(N, r0 + PK_BN_ROWS_PER_BLOCK);
  double a = 0.0, b = 0.0;
  if (o < H)
    for (int r = r0 + ty; r < r1; r += 8) {
      const double v = (double)z[(size_t)r * H + o];
      a += v;
      b += v * v;
    }
  s1[ty][tx] = a;
  s2[ty][tx] = b;
  __syncthreads();
  if (ty == 0 && o < H) {
    for (int i = 1; i < 8; ++i) {
      a += s1[i][tx];
      b += s2[i][tx];
    }
    atomicAdd(acc + o, a);
    atomicAdd(acc + H + o, b);
  }
}

// mean / rstd of the batch and the running statistics (torch.nn.BatchNorm1d defaults: eps 1e-5, momentum 0.1,
// running_var moves towards the UNBIASED batch variance); running == nullptr leaves them alone
__global__ void pk_bn_finalize_kernel(const double* __restrict__ acc, int N, int H, float eps, float momentum,
                                      float* __restrict__ stat, float* __restrict__ running) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= H) return;
  const double mean = acc[o] / (double)N;
  double var = acc[H + o] / (double)N - mean * mean;
  if (var < 0.0) var = 0.0;
  stat[o] = (float)mean;
  stat[H + o] = (float)(1.0 / sqrt(var + (double)eps));
  if (running != nullptr) {
    const double unbiased = N > 1 ? var * ((double)N / (double)(N - 1)) : var;
    running[o] = (float)((1.0 - (double)momentum) * (double)running[o] + (double)momentum * mean);
    running[H + o] = (float)((1.0 - (double)momentum) * (double)running[H + o] + (double)momentum * unbiased);
  }
}

__global__ void pk_bn_apply_kernel(const float* __restrict__ z, const float* __restrict__ stat,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, size_t total, int H,
                                   float* __restrict__ y) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(i % (size_t)H);
    const float xhat = (z[i] - stat[o]) * stat[H + o];
    y[i] = fmaxf(fmaf(xhat, gamma[o], beta[o]), 0.0f);
  }
}

// q[n] = b_out + sum_o w_out[o] * y[n][o]: one warp per column
__global__ void pk_bn_head_kernel(const float* __restrict__ y, const float* __restrict__ wout, const float* __restrict__ bout,
                                  int N, int H, float* __restrict__ q) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  float s = 0.0f;
  for (int o = lane; o < H; o += 32) s = fmaf(wout[o], y[(size_t)warp * H + o], s);
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if (lane == 0) q[warp] = s + bout[0];
}

// Double-DQN target, Huber loss and its derivative (trainer.py:163-187).  slot[i] = row of transition i's next state in
// the compacted non-final arrays (q_no / q_nt / mask_next), -1 when terminal.  dq has one entry per (transition, action)
// column, zero except at the chosen action.  One CTA: the loss is summed in a fixed order.
__global__ void __launch_bounds__(1024)
pk_bn_dq_kernel(const float* __restrict__ q, const int* __restrict__ action, const float* __restrict__ reward,
                const int* __restrict__ slot, const float* __restrict__ q_no, const float* __restrict__ q_nt,
                const int* __restrict__ mask_next, int n, int A, int n_global, float gamma, int double_dqn,
                float* __restrict__ dq, float* __restrict__ loss) {
  __shared__ float sh[1024];
  float acc = 0.0f;
  for (int i = threadIdx.x; i < n; i += 1024) {
    float v = 0.0f;
    const int s = slot[i];
    if (s >= 0) {
      float best = -CUDART_INF_F;
      int bi = 0;
      const float* sel = double_dqn ? q_no : q_nt;
      for (int c = 0; c < A; ++c) {
        const float x = mask_next[(size_t)s * A + c] != 0 ? sel[(size_t)s * A + c] : -CUDART_INF_F;
        if (x > best) {
          best = x;
          bi = c;
        }
      }
      v = double_dqn ? q_nt[(size_t)s * A + bi] : best;
    }
    const float y = fmaf(v, gamma, reward[i]);
    const float d = q[(size_t)i * A + action[i]] - y;
    const float ad = fabsf(d);
    acc += ad < 1.0f ? 0.5f * d * d : ad - 0.5f;
    for (int c = 0; c < A; ++c) dq[(size_t)i * A + c] = 0.0f;
    dq[(size_t)i * A + action[i]] = fminf(fmaxf(d, -1.0f), 1.0f) / (float)n_global;
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = sh[0] / (float)n_global;
}

// ---- backward -----------------------------------------------------------------------------------------------------------
// dy of a layer's output: the head's (dq[n] * w_out[o]) for the last layer, the dH buffer of the layer above otherwise
__device__ __forceinline__ float pk_bn_dy(const float* __restrict__ dq, const float* __restrict__ wout,
                                          const float* __restrict__ dh, size_t r, int o, int H) {
  return dq != nullptr ? dq[r] * wout[o] : dh[r * H + o];
}

// head: acc[o] += dq[n] * y[n][o] (d w_out), acc[H] += dq[n] (d b_out)
__global__ void __launch_bounds__(256) pk_bn_head_bwd_kernel(const float* __restrict__ dq, const float* __restrict__ y, int N,
                                                             int H, double* __restrict__ acc) {
  __shared__ double s1[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int o = blockIdx.y * 32 + tx;  // o == H is the bias column
  const int r0 = blockIdx.x * PK_BN_ROWS_PER_BLOCK, r1 = min(N, r0 + PK_BN_ROWS_PER_BLOCK);
  double a = 0.0;
  if (o <= H)
    for (int r = r0 + ty; r < r1; r += 8) {
      const float d = dq[r];
      if (d != 0.0f) a += (double)d * (o < H ? (double)y[(size_t)r * H + o] : 1.0);
    }
  s1[ty][tx] = a;
  __syncthreads();
  if (ty == 0 && o <= H) {
    for (int i = 1; i < 8; ++i) a += s1[i][tx];
    atomicAdd(acc + o, a);
  }
}

// BatchNorm backward, reductions: acc[o] += dyr (d beta), acc[H + o] += dyr * xhat (d gamma); dyr = dy where y > 0
__global__ void __launch_bounds__(256)
pk_bn_bwd_reduce_kernel(const float* __restrict__ dq, const float* __restrict__ wout, const float* __restrict__ dh,
                        const float* __restrict__ y, const float* __restrict__ z, const float* __restrict__ stat, int N, int H,
                        double* __restrict__ acc) {
  __shared__ double s1[8][32], s2[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int o = blockIdx.y * 32 + tx;
  const int r0 = blockIdx.x * PK_BN_ROWS_PER_BLOCK, r1 = min(N, r0 + PK_BN_ROWS_PER_BLOCK);
  double a = 0.0, b = 0.0;
  if (o < H) {
    const float mean = stat[o], rstd = stat[H + o];
    for (int r = r0 + ty; r < r1; r += 8) {
      const size_t i = (size_t)r * H + o;
      if (y[i] > 0.0f) {
        const float dy = pk_bn_dy(dq, wout, dh, (size_t)r, o, H);
        a += (double)dy;
        b += (double)dy * (double)((z[i] - mean) * rstd);
      }
    }
  }
  s1[ty][tx] = a;
  s2[ty][tx] = b;
  __syncthreads();
  if (ty == 0 && o < H) {
    for (int i = 1; i < 8; ++i) {
      a += s1[i][tx];
      b += s2[i][tx];
    }
    atomicAdd(acc + o, a);
    atomicAdd(acc + H + o, b);
  }
}

// dz = gamma * rstd * (dyr - mean(dyr) - xhat * mean(dyr * xhat))
__global__ void pk_bn_bwd_apply_kernel(const float* __restrict__ dq, const float* __restrict__ wout,
                                       const float* __restrict__ dh, const float* __restrict__ y, const float* __restrict__ z,
                                       const float* __restrict__ stat, const float* __restrict__ gamma,
                                       const double* __restrict__ acc, int N, int H, float* __restrict__ dz) {
  const size_t total = (size_t)N * H;
  const float inv_n = 1.0f / (float)N;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int o = (int)(i % (size_t)H);
    const size_t r = i / (size_t)H;
    const float rstd = stat[H + o];
    const float xhat = (z[i] - stat[o]) * rstd;
    const float dyr = y[i] > 0.0f ? pk_bn_dy(dq, wout, dh, r, o, H) : 0.0f;
    const float db = (float)acc[o], dg = (float)acc[H + o];
    dz[i] = gamma[o] * rstd * (dyr - db * inv_n - xhat * dg * inv_n);
  }
}

// double accumulators -> the flat gradient: bn weight / bias of every layer, output.weight, output.bias
__global__ void pk_bn_bwd_params_kernel(const double* __restrict__ acc, PkBnLayout L, int nl, int H, float* __restrict__ grad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nl * H) {
    const int l = i / H, o = i % H;
    grad[L.beta[l] + o] = (float)acc[(size_t)l * 2 * H + o];
    grad[L.gamma[l] + o] = (float)acc[(size_t)l * 2 * H + H + o];
  } else if (i < nl * H + H + 1) {
    const int o = i - nl * H;
    grad[L.wout + o] = (float)acc[(size_t)nl * 2 * H + o];  // o == H lands on L.bout (= L.wout + H)
  }
}

// clamp(grad) -> Adam over a plain vector; the arithmetic of pk_clamp_adam_kernel (pk_update.cuh) without a layout
__global__ void pk_clamp_adam_flat_kernel(float* __restrict__ params, const float* __restrict__ grad, float* __restrict__ m,
                                          float* __restrict__ v, unsigned n, PkAdam h) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float g = fminf(fmaxf(grad[i], -1.0f), 1.0f);
  const float mi = m[i] + (g - m[i]) * h.one_minus_b1;
  const float vi = v[i] * h.b2 + (h.one_minus_b2 * g) * g;
  const float denom = sqrtf(vi) / h.bc2_sqrt + h.eps;
  params[i] = params[i] + (-h.step_size * mi) / denom;
  m[i] = mi;
  v[i] = vi;
}

// ---- layer 0 from the PACKED records ---------------------------------------------------------------------------------
// A feature column is the battle's 50 state values (shared by its action columns, state_feature_extractor.py:92-197) plus at
// most eight one-hot choice rows (choice_to_vec.py:40-61): z0 = sum_k sv[k] W0[:, k] + sum_t W0[:, row_t], 58 rows of
// W0^T instead of 617, and no dense (B, 617, A) block in HBM.  One warp per battle, lanes over output channels.
#define PK_BN_L0_WARPS 8
struct PkBnL0Smem {
  uint4 rec[PK_BN_L0_WARPS][9];
  float sv[PK_BN_L0_WARPS][64];
  int idx[PK_BN_L0_WARPS][8 * PK_MAX_ACTIONS];
};
// parses battle b into this warp's slots; returns n_act (choice rows of columns >= n_act are -1: feature_extractor.py:49)
__device__ __forceinline__ int pk_bn_l0_parse(PkBnL0Smem& sm, int warp, int lane, const uint4* __restrict__ packed, int b,
                                              int A) {
  if (lane < 9) sm.rec[warp][lane] = __ldg(packed + (size_t)b * 9 + lane);
  __syncwarp();
  const uint8_t* r = reinterpret_cast<const uint8_t*>(sm.rec[warp]);
  const int na = min((int)r[PK_OFF_NACT], A);
  sm.sv[warp][lane] = pk_state_feature(r, lane);
  sm.sv[warp][lane + 32] = pk_state_feature(r, lane + 32);
  for (int e = lane; e < 8 * A; e += 32) {
    const int c = e >> 3;
    sm.idx[warp][e] = c < na ? pk_choice_row(r, c, e & 7) : -1;
  }
  __syncwarp();
  return na;
}

// in [R][Cc] -> out [Cc][R] (out[c * R + r] = in[r * Cc + c]) for the rows r >= r_from of `in`; block (32, 8)
__global__ void pk_bn_transpose_kernel(const float* __restrict__ in, int R, int Cc, int r_from, float* __restrict__ out) {
  __shared__ float t[32][33];
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    t[i][threadIdx.x] = (r < R && c < Cc) ? in[(size_t)r * Cc + c] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < Cc && r >= r_from) out[(size_t)c * R + r] = t[threadIdx.x][i];
  }
}

__global__ void __launch_bounds__(32 * PK_BN_L0_WARPS)
pk_bn_l0_fwd_kernel(const uint4* __restrict__ packed, int B, int A, const float* __restrict__ w0t, int H,
                    float* __restrict__ z, int* __restrict__ mask) {
  __shared__ PkBnL0Smem sm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * PK_BN_L0_WARPS + warp;
  if (b >= B) return;
  const int na = pk_bn_l0_parse(sm, warp, lane, packed, b, A);
  if (mask != nullptr && lane < A) mask[(size_t)b * A + lane] = lane < na ? 1 : 0;
  for (int o = lane; o < H; o += 32) {
    float base = 0.0f;
#pragma unroll 10
    for (int k = 0; k < PK_N_STATE; ++k) base = fmaf(sm.sv[warp][k], w0t[(size_t)k * H + o], base);
    for (int a = 0; a < A; ++a) {
      float v = base;
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int row = sm.idx[warp][8 * a + t];
        if (row >= 0) v += w0t[(size_t)row * H + o];
      }
      z[((size_t)b * A + a) * H + o] = v;
    }
  }
}

// backward of layer 0: the one-hot part of dW0 as coalesced atomics into a TRANSPOSED buffer gw0t [C][H] (zeroed by the
// caller); the state part leaves as two compact matrices, SV [B][64] and DS [B][H] = sum over the battle's columns of
// dz, which one small GEMM (pk_bn_gemm_kernel) contracts over the batch
__global__ void __launch_bounds__(32 * PK_BN_L0_WARPS)
pk_bn_l0_bwd_kernel(const uint4* __restrict__ packed, int B, int A, const float* __restrict__ dz, int H,
                    float* __restrict__ gw0t, float* __restrict__ svm, float* __restrict__ dsm) {
  __shared__ PkBnL0Smem sm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * PK_BN_L0_WARPS + warp;
  if (b >= B) return;
  pk_bn_l0_parse(sm, warp, lane, packed, b, A);
  svm[(size_t)b * 64 + lane] = sm.sv[warp][lane];
  svm[(size_t)b * 64 + lane + 32] = sm.sv[warp][lane + 32];
  for (int o = lane; o < H; o += 32) {
    float ds = 0.0f;
    for (int a = 0; a < A; ++a) {
      const float d = dz[((size_t)b * A + a) * H + o];
      ds += d;
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int row = sm.idx[warp][8 * a + t];
        if (row >= 0) atomicAdd(gw0t + (size_t)row * H + o, d);
      }
    }
    dsm[(size_t)b * H + o] = ds;
  }
}
