// Shared device/host definitions of the pokeai_b200 kernels (sm_100a only).
//
// Feature row order follows the reference bit for bit:
//   state rows  0..49   pokeai/ai/state_feature_extractor.py:74-90,113-197
//   choice rows 50..    pokeai/ai/generic_move_model/choice_to_vec.py:16-28,40-61
//   assembly            pokeai/ai/generic_move_model/feature_extractor.py:40-52
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pokeai_b200.h"

static_assert(sizeof(pk_action) == 12, "pk_action must be 12 bytes");
static_assert(sizeof(pk_obs) == PK_OBS_BYTES, "pk_obs must be 144 bytes");

// ---- state block offsets (state_feature_extractor.py:42-48 order) -------------------------
#define PK_ST_REMAINING 0
#define PK_ST_TYPE 2
#define PK_ST_HP 19
#define PK_ST_NV 21
#define PK_ST_RANK 33
#define PK_ST_WEATHER 47
// ---- choice block offsets relative to row PK_N_STATE (choice_to_vec.py:17-26) --------------
#define PK_CH_SWITCH 0
#define PK_CH_FORCE 1
#define PK_CH_SPECIES 2

// byte offsets inside pk_obs
#define PK_OFF_REMAINING 0
#define PK_OFF_TOTAL 2
#define PK_OFF_HPCUR 4
#define PK_OFF_HPMAX 8
#define PK_OFF_STATUS 12
#define PK_OFF_BOOST 14
#define PK_OFF_OPP 28
#define PK_OFF_WEATHER 29
#define PK_OFF_NACT 30
#define PK_OFF_ACT 32

struct PkDeviceTables {
  int n_species, n_moves, n_items;
  int C;         // feature channels
  int row_move;  // first move row   (absolute channel index)
  int row_item;  // first item row
  int pad0, pad1;
  uint32_t type_mask[256];
};

// Layout of the prepared (transposed, padded) network, in floats.  HP = H rounded up to 32.
//   w1t  [C+1][HP]  row k = input channel k, row C = zeros (target of "no row" gathers)
//   b[0] [HP]
//   wl[l] [HP][HP] (l = 1..n_layers-1), b[l] [HP]
//   wout [HP], bout [4]                                   <- `total` ends here (what inference reads)
//   wn[l] [HP][HP] (l = 1..n_layers-1): the same hidden matrices NOT transposed, row j = output channel,
//                  read by the backward pass (dh = W^T dz)           <- `total_train` ends here
//   tc[l] (HP == 64 only, l = 1..n_layers-1): the same hidden matrix as the B operand of tcgen05.mma.kind::tf32 --
//                  two 64 x 64 fp32 tiles, hi = tf32(w) and lo = w - hi, each in the K-major SWIZZLE_128B shared-memory
//                  layout (pk_tc_tile_off), so pk_decide_tc_kernel stages them with one TMA bulk load
//   tc0 (HP == 64 only): W0 for pk_dense_tc_kernel -- per k-block of 32 input rows one 128-row K-major SWIZZLE_128B tile,
//                  rows 0-63 = tf32(W0[j][k]), rows 64-127 = the remainders (hi element + PK_TC0_LO_DELTA), zero padded
//                  beyond C and H; 2 ceil(C / 64) tiles of PK_TC_TILE_FLOATS                            <- `total_all`
#define PK_TC_TILE_FLOATS 4096u
#define PK_TC0_LO_DELTA 2048u  // 64 rows of 32 floats
struct PkLayout {
  int C, H, HP, n_layers;
  unsigned w1t, wout, bout, total, total_train, total_all;
  unsigned b[PK_MAX_LAYERS];
  unsigned wl[PK_MAX_LAYERS];
  unsigned wn[PK_MAX_LAYERS];
  unsigned tc[PK_MAX_LAYERS];
  unsigned tc0;  // layer 0 as tensor-core B operand, see below (0 when HP != 64)
};

// float offset of element (row, k) of a K-major SWIZZLE_128B fp32 operand tile with 64 rows x 64 k: k-blocks of 32
// floats (128 B) are whole sub-tiles of `kb_floats`; inside one, 8-row groups are 1024 B apart and the 16-byte chunk
// index is XORed with the row's index in its group (cute Swizzle<3,4,3>)
__host__ __device__ inline unsigned pk_tc_tile_off(unsigned row, unsigned k, unsigned kb_floats) {
  const unsigned kb = k >> 5, chunk = (k >> 2) & 7u, r8 = row & 7u;
  return kb * kb_floats + (row >> 3) * 256u + r8 * 32u + ((chunk ^ r8) << 2) + (k & 3u);
}
__device__ __forceinline__ float pk_tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__host__ __device__ inline PkLayout pk_make_layout(int n_layers, int H, int C) {
  PkLayout L;
  L.C = C;
  L.H = H;
  L.HP = (H + 31) & ~31;
  L.n_layers = n_layers;
  unsigned o = 0;
  L.w1t = o;
  o += (unsigned)(C + 1) * L.HP;
  L.b[0] = o;
  o += L.HP;
  L.wl[0] = 0;
  for (int l = 1; l < PK_MAX_LAYERS; ++l) {
    if (l < n_layers) {
      L.wl[l] = o;
      o += (unsigned)L.HP * L.HP;
      L.b[l] = o;
      o += L.HP;
    } else {
      L.wl[l] = 0;
      L.b[l] = 0;
    }
  }
  L.wout = o;
  o += L.HP;
  L.bout = o;
  o += 4;
  L.total = o;
  for (int l = 0; l < PK_MAX_LAYERS; ++l) {
    if (l >= 1 && l < n_layers) {
      L.wn[l] = o;
      o += (unsigned)L.HP * L.HP;
    } else {
      L.wn[l] = 0;
    }
  }
  L.total_train = o;
  for (int l = 0; l < PK_MAX_LAYERS; ++l) {
    if (l >= 1 && l < n_layers && L.HP == 64) {
      L.tc[l] = o;
      o += 2u * PK_TC_TILE_FLOATS;
    } else {
      L.tc[l] = 0;
    }
  }
  L.tc0 = 0;
  if (L.HP == 64) {
    L.tc0 = o;
    o += (unsigned)((C + 63) >> 6) * 2u * PK_TC_TILE_FLOATS;  // whole 64-row k-blocks
  }
  L.total_all = o;
  return L;
}

// flat parameter index (state_dict order) -> offset in the prepared layout; *wn_off = offset of the
// untransposed copy for hidden-layer weights, 0 otherwise
struct PkParamLayout;
__host__ __device__ inline unsigned pk_flat_to_prepared(const PkParamLayout& P, const PkLayout& L, unsigned i,
                                                        unsigned* wn_off, unsigned* tc_off = nullptr);

// Flat parameter vector in state_dict order (layers.i.weight, layers.i.bias, ..., output.weight, output.bias)
struct PkParamLayout {
  int C, H, n_layers;
  unsigned w[PK_MAX_LAYERS], b[PK_MAX_LAYERS], wout, bout, total;
};

__host__ __device__ inline PkParamLayout pk_make_param_layout(int n_layers, int H, int C) {
  PkParamLayout P;
  P.C = C;
  P.H = H;
  P.n_layers = n_layers;
  unsigned o = 0;
  for (int l = 0; l < PK_MAX_LAYERS; ++l) {
    if (l < n_layers) {
      P.w[l] = o;
      o += (unsigned)H * (l == 0 ? C : H);
      P.b[l] = o;
      o += H;
    } else {
      P.w[l] = 0;
      P.b[l] = 0;
    }
  }
  P.wout = o;
  o += H;
  P.bout = o;
  o += 1;
  P.total = o;
  return P;
}

// *tc_off (optional) = offset of the hi element in the tensor-core tile of a weight (lo at + pk_tc_lo_delta), 0 when
// the layout has none
__host__ __device__ inline unsigned pk_flat_to_prepared(const PkParamLayout& P, const PkLayout& L, unsigned i,
                                                        unsigned* wn_off, unsigned* tc_off) {
  *wn_off = 0;
  if (tc_off) *tc_off = 0;
  for (int l = 0; l < P.n_layers; ++l) {
    if (i < P.b[l]) {
      const unsigned e = i - P.w[l], cin = l == 0 ? (unsigned)P.C : (unsigned)P.H;
      const unsigned j = e / cin, k = e - j * cin;
      if (l > 0) *wn_off = L.wn[l] + j * L.HP + k;
      if (l > 0 && tc_off && L.tc[l]) *tc_off = L.tc[l] + pk_tc_tile_off(j, k, 2048u);
      if (l == 0 && tc_off && L.tc0) *tc_off = L.tc0 + pk_tc_tile_off(j, k, PK_TC_TILE_FLOATS);
      return (l == 0 ? L.w1t : L.wl[l]) + k * L.HP + j;
    }
    if (i < P.b[l] + (unsigned)P.H) return L.b[l] + (i - P.b[l]);
  }
  if (i < P.bout) return L.wout + (i - P.wout);
  return L.bout;
}

// distance from the hi element of a tensor-core tile to its lo element
__host__ __device__ inline unsigned pk_tc_lo_delta(const PkLayout& L, unsigned tc_off) {
  return (L.tc0 && tc_off >= L.tc0) ? PK_TC0_LO_DELTA : PK_TC_TILE_FLOATS;
}

struct PkSelect {
  int mode;
  float eps;
  float temperature;
  const float* u_explore;
  const float* u_pick;
  float* probs;
  unsigned long long* trace;  // debug phase trace, normally null
};

// ---- small PTX helpers -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t pk_smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
// generic-proxy writes to shared memory -> visible to the async proxy (TMA) that reads them next
__device__ __forceinline__ void pk_fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// TMA 1-D bulk store shared -> global; size and both addresses multiples of 16
__device__ __forceinline__ void pk_bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(pk_smem_u32(ssrc)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void pk_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void pk_bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void pk_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// TMA 1-D bulk load global -> shared, completion on an mbarrier
__device__ __forceinline__ void pk_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pk_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void pk_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pk_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pk_bulk_load(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   pk_smem_u32(sdst)),
               "l"(gsrc), "r"(bytes), "r"(pk_smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void pk_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(pk_smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void pk_fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ float pk_ldg_f(const float* p) { return __ldg(p); }
// programmatic dependent launch (cudaLaunchAttributeProgrammaticStreamSerialization)
__device__ __forceinline__ void pk_grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pk_grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// inverse of pk_flat_to_prepared for the inference part of the prepared layout: position p in [0, L.total) -> flat
// parameter index, or -1 for padding / the zero row
__host__ __device__ inline int pk_prepared_to_flat(const PkParamLayout& P, const PkLayout& L, unsigned p) {
  const unsigned HP = (unsigned)L.HP, H = (unsigned)P.H;
  if (p < L.b[0]) {  // W0^T [C + 1][HP]
    const unsigned k = p / HP, j = p - k * HP;
    return (k < (unsigned)P.C && j < H) ? (int)(P.w[0] + j * (unsigned)P.C + k) : -1;
  }
  for (int l = 0; l < P.n_layers; ++l) {
    if (p >= L.b[l] && p < L.b[l] + HP) return (p - L.b[l] < H) ? (int)(P.b[l] + (p - L.b[l])) : -1;
    if (l > 0 && p >= L.wl[l] && p < L.wl[l] + HP * HP) {
      const unsigned e = p - L.wl[l], k = e / HP, j = e - k * HP;
      return (k < H && j < H) ? (int)(P.w[l] + j * H + k) : -1;
    }
  }
  if (p >= L.wout && p < L.wout + HP) return (p - L.wout < H) ? (int)(P.wout + (p - L.wout)) : -1;
  if (p == L.bout) return (int)P.bout;
  return -1;
}
