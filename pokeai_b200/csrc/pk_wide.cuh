// Wide policy network (BASELINE.json configs[4]: 4 x 2048 hidden) -- the tensor-core path.
//
// The reference evaluates the same MLPModel (mlp_model.py:5-30) whatever its width; at 2048 channels the
// per-column MLP is a chain of dense contractions (36 864 rows x 2048 x 2048 at 4 096 battles), 11 000 FLOP/B,
// so it belongs on the 5th-generation tensor cores:
//
//   pk_wide_encode_kernel   packed records -> X (B*A, 640) bf16, K-major rows (one row per action column: the
//                           50 state values, the <= 8 ones, zero padding 617 -> 640); the dense feature tensor of
//                           the reference is never built
//   pk_wide_gemm_kernel     Y = act(X W^T + b): TMA (cp.async.bulk.tensor, SWIZZLE_128B) fills a 4-stage shared-memory
//                           ring, ONE elected thread issues tcgen05.mma (cta_group::1, kind::f16, bf16 x bf16 ->
//                           fp32) into a double-buffered TMEM accumulator (2 x 256 columns), eight epilogue warps
//                           read it back with tcgen05.ld, add the bias, apply ReLU and store bf16 (or fp32) rows.
//                           Persistent: one CTA per SM walks the (m, n) tiles; the 8 n-tiles of one m-block are
//                           adjacent in the walk so the activation block is read from HBM once.
//   pk_wide_head_kernel     q = h . w_out + b_out per row, then the masked selection of pk_decide.cuh
//
// Numerics: operands are rounded to bf16, accumulation is fp32 -- the bf16 tolerance of SURVEY.md 8c
// (|dq| <= 2e-2 on O(1) logits) applies, not the 1e-5 of the fp32 small-net path.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>

#include "pk_decide.cuh"

#define PKW_BM 128
#define PKW_BN 256
#define PKW_BK 64
#define PKW_STAGES 4
#define PKW_KPAD0 640  // 617 input channels padded to a multiple of PKW_BK
#define PKW_EPI_WARPS 8   // two per TMEM lane quarter, each takes half of the tile's 256 columns
#define PKW_THREADS (64 + 32 * PKW_EPI_WARPS)
#define PKW_A_BYTES (PKW_BM * PKW_BK * 2)
#define PKW_B_BYTES (PKW_BN * PKW_BK * 2)
#define PKW_STAGE_BYTES (PKW_A_BYTES + PKW_B_BYTES)
#define PKW_OUT_STAGE_BYTES 2048  // one epilogue warp's 32 x 32 bf16 block on its way to a TMA store
#define PKW_OUT_BYTES (PKW_EPI_WARPS * 2 * PKW_OUT_STAGE_BYTES)  // double-buffered per warp
#define PKW_SMEM_BYTES (PKW_STAGES * PKW_STAGE_BYTES + PKW_OUT_BYTES + 1024 + 256)

// ---- PTX wrappers --------------------------------------------------------------------------------------
__device__ __forceinline__ void pkw_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(pk_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void pkw_tma_load_2d(void* sdst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          pk_smem_u32(sdst)),
      "l"(map), "r"(pk_smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// TMA 2-D store shared -> global (bulk async-group completion); out-of-bounds rows / columns of the box are clipped
__device__ __forceinline__ void pkw_tma_store_2d(const CUtensorMap* map, const void* ssrc, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map),
               "r"(pk_smem_u32(ssrc)), "r"(c0), "r"(c1)
               : "memory");
}
// ---- CTA-pair (cta_group::2) forms: two CTAs of a cluster on adjacent SMs run ONE MMA over 256 rows -----------------
#define PKW_PEER_MASK 0xFEFFFFFFu  // clears the CTA-rank bit of a shared::cluster address: the pair's even (leader) CTA
__device__ __forceinline__ uint32_t pkw_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void pkw_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");  // not .aligned: lanes 0 of the producer / MMA warps arrive apart from their warps
}
// TMA load whose byte count is credited to the LEADER CTA's mbarrier (same offset in the leader's shared memory)
__device__ __forceinline__ void pkw_tma_load_2d_pair(void* sdst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          pk_smem_u32(sdst)),
      "l"(map), "r"(pk_smem_u32(bar) & PKW_PEER_MASK), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void pkw_mma_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// arrives on the barrier at this offset in BOTH CTAs of the pair once every MMA issued so far has completed
__device__ __forceinline__ void pkw_mma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   pk_smem_u32(bar)),
               "h"((unsigned short)3)
               : "memory");
}
// arrive on the LEADER CTA's barrier at this offset (from either CTA of the pair)
__device__ __forceinline__ void pkw_mbar_arrive_leader(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\tmapa.shared::cluster.u32 ra, %0, 0;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(pk_smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void pkw_tmem_alloc_pair(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(pk_smem_u32(slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void pkw_tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
#define PKW_STAGES2 5  // CTA-pair form: a stage is A 16 KB + half of B 16 KB
#define PKW_SMEM_BYTES2 (PKW_STAGES2 * (PKW_A_BYTES + PKW_B_BYTES / 2) + PKW_OUT_BYTES + 1024 + 256)

__device__ __forceinline__ void pkw_prefetch_map(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void pkw_tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pkw_tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pkw_tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(pk_smem_u32(slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void pkw_tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; issued by one thread for the whole CTA
__device__ __forceinline__ void pkw_mma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
// arrives on the mbarrier once every MMA issued so far by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void pkw_mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(pk_smem_u32(bar))
               : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets TMEM lane (base lane + t)
__device__ __forceinline__ void pkw_tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void pkw_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row groups 1024 B apart (cute::UMMA::SmemDescriptor)
__device__ __forceinline__ uint64_t pkw_smem_desc_k128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFFu);  // start address, 16-byte units
  d |= (uint64_t)1u << 16;                  // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024u >> 4) << 32;        // stride byte offset: 8 rows x 128 B
  d |= (uint64_t)1u << 46;                  // descriptor version (Blackwell)
  d |= (uint64_t)2u << 61;                  // SWIZZLE_128B
  return d;
}
// MN-major operand (the contraction index is the ROW of the row-major source, e.g. dW = dZ^T H): the tile is a
// row of 64-element-wide TMA boxes, each box 64 k-rows x 128 B; 8-row groups 1024 B apart (stride byte offset),
// successive 64-element chunks of the M/N extent one box = 8192 B apart (leading byte offset)
__device__ __forceinline__ uint64_t pkw_smem_desc_mn128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
  d |= (uint64_t)(8192u >> 4) << 16;
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1u << 46;
  d |= (uint64_t)2u << 61;
  return d;
}
// instruction descriptor, kind::f16: fp32 accumulate, bf16 x bf16, M x N tile; mn_major != 0: both operands MN-major
__host__ __device__ constexpr uint32_t pkw_instr_desc(int M, int N, int mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (mn_major ? (3u << 15) : 0u) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

struct PkwGemmArgs {
  const float* bias;      // (N,) fp32 or null
  __nv_bfloat16* out_bf;  // (M, N) bf16 row-major or null
  float* out_f32;         // (M, N) fp32 row-major or null
  const __nv_bfloat16* relu_mask;  // (M, N) bf16 or null: out = mask > 0 ? acc : 0 (backward through a ReLU)
  int M, N, K;            // N % 32 == 0; tails in M, N (to 256) and K (to 64) are zero-filled by TMA
  int relu;
  int tma_store;          // out_bf (and only out_bf) is written by TMA stores through map_o
  int ksplit;             // >= 1: the contraction is cut into ksplit ranges; range s writes out_f32 + s * M * N
  // fused 1-channel head (mlp_model.py:26-28): head_part[m][2 nb + half] = sum over the half tile's 128 columns of
  // bf16(act(acc + bias))[n] * head_w[n]; pk_wide_head_sum_kernel adds the 2 N / 256 partials of a row in a fixed order.
  // With out_bf == nullptr the activations themselves are never written (inference: 151 MB at 4 096 states x 2048).
  const float* head_w;    // (N,) fp32 or null
  float* head_part;       // (M, 2 * ceil(N / 256)) fp32
};

// MN == 0:  Y (M, N) = act(A (M, K) . W (N, K)^T + bias): both operands bf16 row-major with K contiguous.
// MN == 1:  Y (M, N) = A (K, M)^T . W (K, N): both operands bf16 row-major with the CONTRACTION index as the row
//           (weight gradient dW = dZ^T H); the tensor maps then have 64 x 64 boxes.
// CG == 2: launched as clusters of two CTAs.  A pair works on a 256-row x 256-column tile: each CTA loads
// its 128 rows of A and HALF of the weight tile (128 of its 256 rows), the leader CTA issues one cta_group::2 MMA per
// K step over both halves, and each CTA's epilogue drains its own 128 accumulator rows.  Per CTA the weight tile costs
// half the shared-memory traffic and half the TMA bytes of the CG == 1 form.
template <int MN, int CG = 1>
__global__ void __launch_bounds__(PKW_THREADS, 1)
pk_wide_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                    const __grid_constant__ CUtensorMap map_o, PkwGemmArgs g) {
  constexpr int STAGES = CG == 2 ? PKW_STAGES2 : PKW_STAGES;
  constexpr uint32_t B_BYTES = PKW_B_BYTES / CG;            // this CTA's part of the weight tile
  constexpr uint32_t STAGE_BYTES = PKW_A_BYTES + B_BYTES;
  constexpr int TILE_M = PKW_BM * CG;                       // rows of an output tile (per CTA: PKW_BM)
  const int crank = CG == 2 ? (int)pkw_cluster_ctarank() : 0;
  const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, n_units = CG == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  // SWIZZLE_128B tiles need 1024-byte alignment: asked for in the declaration, so that the accesses stay in the shared
  // address space (STS / LDS instead of generic stores and loads)
  extern __shared__ __align__(1024) unsigned char pkw_smem[];
  unsigned char* smem = pkw_smem;
  if ((pk_smem_u32(smem) & 1023u) != 0u) __trap();
  unsigned char* tiles = smem;
  unsigned char* out_stage = smem + STAGES * STAGE_BYTES;  // 1024-byte aligned: SWIZZLE_64B staging
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + PKW_OUT_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_mblk = (g.M + TILE_M - 1) / TILE_M, n_nblk = (g.N + PKW_BN - 1) / PKW_BN,
            n_kblk = (g.K + PKW_BK - 1) / PKW_BK;
  const int n_mn = n_mblk * n_nblk;
  const int n_tiles = n_mn * g.ksplit;                      // tile = split * n_mn + mb * n_nblk + nb
  const int kps = (n_kblk + g.ksplit - 1) / g.ksplit;       // k-blocks per split

  if (threadIdx.x == 0) {
    pkw_prefetch_map(&map_a);
    pkw_prefetch_map(&map_w);
    if (g.tma_store) pkw_prefetch_map(&map_o);
    for (int i = 0; i < STAGES; ++i) {
      pk_mbar_init(full + i, 1);
      pk_mbar_init(empty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      pk_mbar_init(tfull + i, 1);
      pk_mbar_init(tempty + i, PKW_EPI_WARPS * CG);  // pair form: both CTAs' epilogue warps arrive on the leader's
    }
    pk_fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (CG == 2) pkw_tmem_alloc_pair(tmem_slot, 512);
    else pkw_tmem_alloc(tmem_slot, 512);
  }
  pkw_tc_fence_before();
  if constexpr (CG == 2) pkw_cluster_sync();  // both CTAs' barriers exist before anything arrives on them remotely
  else __syncthreads();
  pkw_tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    int s = 0;
    uint32_t ph = 0;
    for (int tile = unit; tile < n_tiles; tile += n_units) {
      const int sp = tile / n_mn, mn = tile - sp * n_mn;
      const int mb = mn / n_nblk, nb = mn - mb * n_nblk;
      const int kb0 = sp * kps, kb1 = min(n_kblk, kb0 + kps);
      for (int kb = kb0; kb < kb1; ++kb) {
        pk_mbar_wait(empty + s, ph ^ 1u);
        unsigned char* sa = tiles + s * STAGE_BYTES;
        if constexpr (CG == 2) {
          // the leader's barrier counts both CTAs' bytes; each CTA brings its rows of A and its half of the weights
          if (crank == 0) pk_mbar_expect_tx(full + s, 2u * STAGE_BYTES);
          if constexpr (MN == 0) {
            pkw_tma_load_2d_pair(sa, &map_a, kb * PKW_BK, mb * TILE_M + crank * PKW_BM, full + s);
            pkw_tma_load_2d_pair(sa + PKW_A_BYTES, &map_w, kb * PKW_BK, nb * PKW_BN + crank * (PKW_BN / 2), full + s);
          } else {
#pragma unroll
            for (int j = 0; j < PKW_BM / 64; ++j)
              pkw_tma_load_2d_pair(sa + j * 8192, &map_a, mb * TILE_M + crank * PKW_BM + 64 * j, kb * PKW_BK, full + s);
#pragma unroll
            for (int j = 0; j < PKW_BN / 2 / 64; ++j)
              pkw_tma_load_2d_pair(sa + PKW_A_BYTES + j * 8192, &map_w, nb * PKW_BN + crank * (PKW_BN / 2) + 64 * j,
                                   kb * PKW_BK, full + s);
          }
          if (++s == STAGES) {
            s = 0;
            ph ^= 1u;
          }
          continue;
        }
        pk_mbar_expect_tx(full + s, STAGE_BYTES);
        if constexpr (MN == 0) {
          pkw_tma_load_2d(sa, &map_a, kb * PKW_BK, mb * PKW_BM, full + s);
          pkw_tma_load_2d(sa + PKW_A_BYTES, &map_w, kb * PKW_BK, nb * PKW_BN, full + s);
        } else {
#pragma unroll
          for (int j = 0; j < PKW_BM / 64; ++j)
            pkw_tma_load_2d(sa + j * 8192, &map_a, mb * PKW_BM + 64 * j, kb * PKW_BK, full + s);
#pragma unroll
          for (int j = 0; j < PKW_BN / 64; ++j)
            pkw_tma_load_2d(sa + PKW_A_BYTES + j * 8192, &map_w, nb * PKW_BN + 64 * j, kb * PKW_BK, full + s);
        }
        if (++s == STAGES) {
          s = 0;
          ph ^= 1u;
        }
      }
    }
  } else if (warp == 1 && lane == 0 && crank == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = pkw_instr_desc(TILE_M, PKW_BN, MN);
    int s = 0;
    uint32_t ph = 0;
    int it = 0;
    for (int tile = unit; tile < n_tiles; tile += n_units, ++it) {
      const int as = it & 1;
      const uint32_t aph = (uint32_t)(it >> 1) & 1u;
      pk_mbar_wait(tempty + as, aph ^ 1u);  // the epilogue has drained this accumulator
      pkw_tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(as * PKW_BN);
      const int sp = tile / n_mn;
      const int kb0 = sp * kps, kb1 = min(n_kblk, kb0 + kps);
      for (int kb = kb0; kb < kb1; ++kb) {
        pk_mbar_wait(full + s, ph);
        pkw_tc_fence_after();
        const uint32_t sa = pk_smem_u32(tiles + s * STAGE_BYTES);
        const uint64_t adesc = MN ? pkw_smem_desc_mn128(sa) : pkw_smem_desc_k128(sa);
        const uint64_t bdesc = MN ? pkw_smem_desc_mn128(sa + PKW_A_BYTES) : pkw_smem_desc_k128(sa + PKW_A_BYTES);
        // UMMA_K = 16: K-major advances 16 bf16 = 32 B inside the swizzle atom (+2 in the 16-byte address field),
        // MN-major advances 16 k-rows of 128 B = 2048 B (+128)
        constexpr uint64_t kstep = MN ? 128u : 2u;
#pragma unroll
        for (int k = 0; k < PKW_BK / 16; ++k) {
          if constexpr (CG == 2)
            pkw_mma_pair(d_tmem, adesc + kstep * (uint64_t)k, bdesc + kstep * (uint64_t)k, idesc, (kb > kb0 || k != 0) ? 1u : 0u);
          else
            pkw_mma(d_tmem, adesc + kstep * (uint64_t)k, bdesc + kstep * (uint64_t)k, idesc, (kb > kb0 || k != 0) ? 1u : 0u);
        }
        // frees the stage (in both CTAs of a pair) once these MMAs have read it
        if constexpr (CG == 2) pkw_mma_commit_pair(empty + s);
        else pkw_mma_commit(empty + s);
        if (++s == STAGES) {
          s = 0;
          ph ^= 1u;
        }
      }
      // accumulator complete (each CTA of a pair holds its 128 rows)
      if constexpr (CG == 2) pkw_mma_commit_pair(tfull + as);
      else pkw_mma_commit(tfull + as);
    }
  } else if (warp >= 2) {
    // ===== epilogue: TMEM -> registers -> bias, ReLU -> global =====
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes 32 * (warp % 4) .. + 31
    // eight epilogue warps: warps 2..5 take the first half of the tile's column chunks, warps 6..9 the second (round 2:
    // with four warps the epilogue of a K = 640 tile took as long as its MMAs and held layer 0 at 44 % tensor-active)
    constexpr int CHW = (PKW_BN / 32) / (PKW_EPI_WARPS / 4);
    const int ch0 = ((warp - 2) >> 2) * CHW;
    // bf16 outputs leave through shared memory and TMA: the block is written in the SWIZZLE_64B order the store expects
    // (16-byte piece j of row r at piece j ^ ((r >> 1) & 3): conflict-free for a warp's 16-byte stores) and ONE thread
    // hands 32 rows x 64 B to the copy engine.  With direct stores every warp-level store instruction touched 32
    // different lines (one 16-byte piece of each of 32 rows 4 KB apart): 4 096 line accesses per 128 x 256 tile.
    unsigned char* my_stage = out_stage + (warp - 2) * (2 * PKW_OUT_STAGE_BYTES);
    int sbuf = 0;
    int it = 0;
    for (int tile = unit; tile < n_tiles; tile += n_units, ++it) {
      const int sp = tile / n_mn, mn = tile - sp * n_mn;
      const int mb = mn / n_nblk, nb = mn - mb * n_nblk;
      float* const out_f32 = g.out_f32 != nullptr ? g.out_f32 + (size_t)sp * g.M * g.N : nullptr;
      const int as = it & 1;
      const uint32_t aph = (uint32_t)(it >> 1) & 1u;
      pk_mbar_wait(tfull + as, aph);
      pkw_tc_fence_after();
      const int row0 = mb * TILE_M + crank * PKW_BM + quarter * 32;  // first row of this warp's TMEM lane quarter
      const int row = row0 + lane;
      const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * PKW_BN);
      float hs0 = 0.0f, hs1 = 0.0f, hs2 = 0.0f, hs3 = 0.0f;  // fused head: this row's partial over the warp's columns
#pragma unroll 1
      for (int ch = ch0; ch < ch0 + CHW; ++ch) {
        uint32_t v[32];
        pkw_tmem_ld32(t0 + (uint32_t)(ch * 32), v);
        pkw_tmem_ld_wait();
        const int n0 = nb * PKW_BN + ch * 32;
        if (n0 >= g.N) break;  // N tail (warp-uniform)
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
        if (g.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 bv = __ldg(reinterpret_cast<const float4*>(g.bias + n0 + j));
            f[j] += bv.x;
            f[j + 1] += bv.y;
            f[j + 2] += bv.z;
            f[j + 3] += bv.w;
          }
        }
        if (g.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
        }
        if (g.relu_mask != nullptr && row < g.M) {
          const uint4* mp = reinterpret_cast<const uint4*>(g.relu_mask + (size_t)row * g.N + n0);
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const uint4 mv = __ldg(mp + (j >> 3));
            const __nv_bfloat162* mh = reinterpret_cast<const __nv_bfloat162*>(&mv);
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float2 mf = __bfloat1622float2(mh[t]);
              if (!(mf.x > 0.0f)) f[j + 2 * t] = 0.0f;
              if (!(mf.y > 0.0f)) f[j + 2 * t + 1] = 0.0f;
            }
          }
        }
        if (g.head_w != nullptr) {  // the head reads the activations as the next layer would: rounded to bf16
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 wv = __ldg(reinterpret_cast<const float4*>(g.head_w + n0 + j));
            hs0 = fmaf(__bfloat162float(__float2bfloat16_rn(f[j])), wv.x, hs0);
            hs1 = fmaf(__bfloat162float(__float2bfloat16_rn(f[j + 1])), wv.y, hs1);
            hs2 = fmaf(__bfloat162float(__float2bfloat16_rn(f[j + 2])), wv.z, hs2);
            hs3 = fmaf(__bfloat162float(__float2bfloat16_rn(f[j + 3])), wv.w, hs3);
          }
        }
        if (g.tma_store) {
          if (lane == 0) pk_bulk_wait_read<1>();  // the store that read this staging buffer two chunks ago is done with it
          __syncwarp();
          unsigned char* st = my_stage + sbuf * PKW_OUT_STAGE_BYTES;
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 o;
            __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]);
            __nv_bfloat162 p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
            __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]);
            __nv_bfloat162 p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
            o.x = *reinterpret_cast<uint32_t*>(&p0);
            o.y = *reinterpret_cast<uint32_t*>(&p1);
            o.z = *reinterpret_cast<uint32_t*>(&p2);
            o.w = *reinterpret_cast<uint32_t*>(&p3);
            *reinterpret_cast<uint4*>(st + lane * 64 + ((((j >> 3) ^ (lane >> 1)) & 3) << 4)) = o;
          }
          pk_fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            pkw_tma_store_2d(&map_o, st, n0, row0);
            pk_bulk_commit();
          }
          sbuf ^= 1;
        } else if (row < g.M) {
          if (g.out_bf != nullptr) {
            uint4* dst = reinterpret_cast<uint4*>(g.out_bf + (size_t)row * g.N + n0);
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 o;
              __nv_bfloat162 p0 = __floats2bfloat162_rn(f[j], f[j + 1]);
              __nv_bfloat162 p1 = __floats2bfloat162_rn(f[j + 2], f[j + 3]);
              __nv_bfloat162 p2 = __floats2bfloat162_rn(f[j + 4], f[j + 5]);
              __nv_bfloat162 p3 = __floats2bfloat162_rn(f[j + 6], f[j + 7]);
              o.x = *reinterpret_cast<uint32_t*>(&p0);
              o.y = *reinterpret_cast<uint32_t*>(&p1);
              o.z = *reinterpret_cast<uint32_t*>(&p2);
              o.w = *reinterpret_cast<uint32_t*>(&p3);
              dst[j >> 3] = o;
            }
          }
          if (out_f32 != nullptr) {
            float4* dst = reinterpret_cast<float4*>(out_f32 + (size_t)row * g.N + n0);
#pragma unroll
            for (int j = 0; j < 32; j += 4) dst[j >> 2] = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
          }
        }
      }
      if (g.head_w != nullptr && row < g.M)
        g.head_part[(size_t)row * (2 * n_nblk) + 2 * nb + ((warp - 2) >> 2)] = (hs0 + hs1) + (hs2 + hs3);
      pkw_tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) pkw_mbar_arrive_leader(tempty + as);
        else pkw_mbar_arrive(tempty + as);
      }
    }
  }
  if (warp >= 2 && lane == 0) pk_bulk_wait_all();  // TMA stores of the last tiles (no-op when none were issued)
  pkw_tc_fence_before();
  if constexpr (CG == 2) pkw_cluster_sync();  // the peer's shared memory and TMEM stay alive until both CTAs are done
  else __syncthreads();
  if (warp == 1) {
    __syncwarp();
    if constexpr (CG == 2) pkw_tmem_dealloc_pair(tmem_base, 512);
    else pkw_tmem_dealloc(tmem_base, 512);
  }
}

// ---- packed records -> X (B*A, KPAD) bf16 (feature_extractor.py:40-52 transposed to one row per column) ----
#define PKW_ENC_WARPS 8
__host__ __device__ inline unsigned pkw_enc_warp_bytes(int A, int kpad) {
  return PK_WARP_FIXED_BYTES + 16u + (unsigned)A * (unsigned)kpad * 2u;
}
__global__ void __launch_bounds__(PKW_ENC_WARPS * 32)
pk_wide_encode_kernel(const uint4* __restrict__ packed, int B, int A, int C, int kpad, __nv_bfloat16* __restrict__ X,
                      int* __restrict__ mask, const int* __restrict__ col_select) {
  // col_select == null: A rows per battle (one per action column).  Otherwise ONE row per battle, the column
  // col_select[b] (the chosen action of a transition: trainer.py:163 gathers only that column).
  extern __shared__ __align__(16) unsigned char enc_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wa = enc_smem + warp * pkw_enc_warp_bytes(A, kpad);
  uint4* rec = reinterpret_cast<uint4*>(wa);
  const uint8_t* r = reinterpret_cast<const uint8_t*>(rec);
  float* sv = reinterpret_cast<float*>(wa + PK_WARP_REC_BYTES);
  int* idx = reinterpret_cast<int*>(wa + PK_WARP_REC_BYTES + PK_WARP_SV_BYTES);
  __nv_bfloat16* rows = reinterpret_cast<__nv_bfloat16*>(wa + PK_WARP_FIXED_BYTES + 16u);  // [A][kpad]
  const int RA = col_select != nullptr ? 1 : A;  // rows emitted per battle
  const int n16 = RA * kpad / 8;                 // uint4 chunks of the battle's rows
  const __nv_bfloat16 one = __float2bfloat16_rn(1.0f);
  for (long long b = (long long)blockIdx.x * PKW_ENC_WARPS + warp; b < B; b += (long long)gridDim.x * PKW_ENC_WARPS) {
    __syncwarp();
    if (lane < 9) rec[lane] = __ldg(packed + b * 9 + lane);
    uint4* r16 = reinterpret_cast<uint4*>(rows);
    for (int i = lane; i < n16; i += 32) r16[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    const int n_act = min((int)r[PK_OFF_NACT], A);
    if (mask != nullptr && lane < A) mask[b * A + lane] = lane < n_act ? 1 : 0;
    sv[lane] = pk_state_feature(r, lane);
    sv[lane + 32] = pk_state_feature(r, lane + 32);
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const int e = lane + 32 * it;
      if (e < 8 * A) {
        const int c = e >> 3;
        const int row = c < n_act ? pk_choice_row(r, c, e & 7) : -1;
        idx[e] = row < 0 ? C : row;
      }
    }
    __syncwarp();
    const int sel = col_select != nullptr ? __ldg(col_select + b) : -1;
    for (int k = lane; k < PK_N_STATE; k += 32) {  // state rows fill every column (feature_extractor.py:48)
      const __nv_bfloat16 v = __float2bfloat16_rn(sv[k]);
      for (int c = 0; c < RA; ++c) rows[c * kpad + k] = v;
    }
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const int e = lane + 32 * it;
      if (e < 8 * A) {
        const int row = idx[e];
        const int c = e >> 3;
        if (row != C && (sel < 0 || c == sel)) rows[(sel < 0 ? c : 0) * kpad + row] = one;
      }
    }
    __syncwarp();
    uint4* dst = reinterpret_cast<uint4*>(X + (size_t)b * RA * kpad);
    for (int i = lane; i < n16; i += 32) dst[i] = r16[i];
  }
}

// ---- head: q[m] = h[m] . w_out + b_out, one warp per row (mlp_model.py:26-28, the 1-channel output conv) ----
__global__ void __launch_bounds__(256)
pk_wide_head_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ wout, const float* __restrict__ bout,
                    int M, int N, float* __restrict__ q) {
  const int lane = threadIdx.x & 31;
  const long long m = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (m >= M) return;
  const uint4* row = reinterpret_cast<const uint4*>(h + (size_t)m * N);
  float s = 0.0f;
  for (int i = lane; i < N / 8; i += 32) {
    const uint4 v = __ldg(row + i);
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(wout + 8 * i));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(wout + 8 * i + 4));
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
    const float2 a = __bfloat1622float2(p[0]), b = __bfloat1622float2(p[1]);
    const float2 c = __bfloat1622float2(p[2]), d = __bfloat1622float2(p[3]);
    s = fmaf(a.x, w0.x, s);
    s = fmaf(a.y, w0.y, s);
    s = fmaf(b.x, w0.z, s);
    s = fmaf(b.y, w0.w, s);
    s = fmaf(c.x, w1.x, s);
    s = fmaf(c.y, w1.y, s);
    s = fmaf(d.x, w1.z, s);
    s = fmaf(d.y, w1.w, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) q[m] = s + __ldg(bout);
}

// the fused head's partials (PkwGemmArgs::head_part) -> q[m] = sum of the row's nparts partials, in order, + b_out
__global__ void __launch_bounds__(256)
pk_wide_head_sum_kernel(const float* __restrict__ part, int nparts, const float* __restrict__ bout, int M,
                        float* __restrict__ q) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  const float* p = part + (size_t)m * nparts;
  float s = 0.0f;
  for (int i = 0; i < nparts; ++i) s += p[i];
  q[m] = s + __ldg(bout);
}

// fp32 (rows, cols) row-major -> bf16 (rows, cols_pad), zero padded columns
__global__ void pk_wide_cast_pad_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols,
                                        int cols_pad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)rows * cols_pad) return;
  const int r = (int)(i / cols_pad), c = (int)(i - (long long)r * cols_pad);
  dst[i] = __float2bfloat16_rn(c < cols ? src[(size_t)r * cols + c] : 0.0f);
}

// ---- backward helpers (Trainer._update's loss.backward(), trainer.py:192, for the wide network) -------------
// fp32 (rows, cols) -> bf16 (cols, rows): the transposed weight copy read by the dH GEMM
__global__ void pk_wide_transpose_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows,
                                              int cols) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? src[(size_t)r * cols + c] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < cols && r < rows) dst[(size_t)c * rows + r] = __float2bfloat16_rn(tile[threadIdx.x][i]);
  }
}

// dZ_L[m][n] = dq[m] * w_out[n] * (h[m][n] > 0)   (backward of the 1-channel head through the last ReLU)
__global__ void __launch_bounds__(256)
pk_wide_dz_head_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ dq, const float* __restrict__ wout,
                       int M, int N, __nv_bfloat16* __restrict__ dz) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per 8 elements
  const long long total = (long long)M * N / 8;
  if (i >= total) return;
  const long long e = i * 8;
  const int m = (int)(e / N), n = (int)(e - (long long)m * N);
  const uint4 hv = __ldg(reinterpret_cast<const uint4*>(h + e));
  const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&hv);
  const float d = __ldg(dq + m);
  const float4 w0 = __ldg(reinterpret_cast<const float4*>(wout + n));
  const float4 w1 = __ldg(reinterpret_cast<const float4*>(wout + n + 4));
  const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
  uint4 o;
  __nv_bfloat162* op = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float2 hf = __bfloat1622float2(hp[t]);
    op[t] = __floats2bfloat162_rn(hf.x > 0.0f ? d * w[2 * t] : 0.0f, hf.y > 0.0f ? d * w[2 * t + 1] : 0.0f);
  }
  *reinterpret_cast<uint4*>(dz + e) = o;
}

// partial[s][n] = sum over the rows of split s of scale[m] * z[m][n]  (scale == null: 1).  Fixed order: each
// thread owns two columns and walks its rows in order; the splits are summed in order by the finish kernel.
#define PKW_COLSUM_SPLITS 256
__global__ void __launch_bounds__(128)
pk_wide_colsum_kernel(const __nv_bfloat16* __restrict__ z, const float* __restrict__ scale, int M, int N,
                      float* __restrict__ partial) {
  const int n = (blockIdx.x * 128 + threadIdx.x) * 8;  // eight columns per thread: one 16-byte load per row
  if (n >= N) return;
  const int s = blockIdx.y;
  const int rows_per = (M + PKW_COLSUM_SPLITS - 1) / PKW_COLSUM_SPLITS;
  const int m0 = s * rows_per, m1 = min(M, m0 + rows_per);
  float a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = 0.0f;
  int m = m0;
  for (; m + 4 <= m1; m += 4) {  // four independent loads in flight, accumulated in row order
    uint4 v[4];
    float sc[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      v[u] = __ldg(reinterpret_cast<const uint4*>(z + (size_t)(m + u) * N + n));
      sc[u] = scale != nullptr ? __ldg(scale + m + u) : 1.0f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float2 f = __bfloat1622float2(p[t]);
        a[2 * t] = fmaf(sc[u], f.x, a[2 * t]);
        a[2 * t + 1] = fmaf(sc[u], f.y, a[2 * t + 1]);
      }
    }
  }
  for (; m < m1; ++m) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(z + (size_t)m * N + n));
    const float sc = scale != nullptr ? __ldg(scale + m) : 1.0f;
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float2 f = __bfloat1622float2(p[t]);
      a[2 * t] = fmaf(sc, f.x, a[2 * t]);
      a[2 * t + 1] = fmaf(sc, f.y, a[2 * t + 1]);
    }
  }
  float4* o = reinterpret_cast<float4*>(partial + (size_t)s * N + n);
  o[0] = make_float4(a[0], a[1], a[2], a[3]);
  o[1] = make_float4(a[4], a[5], a[6], a[7]);
}
__global__ void pk_wide_colsum_finish_kernel(const float* __restrict__ partial, int N, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float a = 0.0f;
  for (int s = 0; s < PKW_COLSUM_SPLITS; ++s) a += partial[(size_t)s * N + n];
  out[n] = a;
}
// out[0] = sum_m dq[m] (gradient of output.bias), one block, fixed order
__global__ void __launch_bounds__(256) pk_wide_sum_kernel(const float* __restrict__ x, int M, float* __restrict__ out) {
  __shared__ float sh[256];
  float a = 0.0f;
  for (int m = threadIdx.x; m < M; m += 256) a += x[m];
  sh[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0];
}
// (rows, cols_pad) fp32 -> (rows, cols) fp32: drops the K padding of layer 0's weight gradient
__global__ void pk_wide_compact_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols,
                                       int cols_pad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)rows * cols) return;
  const int r = (int)(i / cols), c = (int)(i - (long long)r * cols);
  dst[i] = src[(size_t)r * cols_pad + c];
}
// out (rows, cols) = sum over s (in order) of partial[s] (rows, cols_pad): split-K reduction + drop of the K padding
__global__ void pk_wide_splitk_reduce_kernel(const float* __restrict__ partial, int nsplit, float* __restrict__ dst,
                                             int rows, int cols, int cols_pad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)rows * cols) return;
  const int r = (int)(i / cols), c = (int)(i - (long long)r * cols);
  float a = 0.0f;
  for (int sp = 0; sp < nsplit; ++sp) a += partial[((size_t)sp * rows + r) * cols_pad + c];
  dst[i] = a;
}

// ---- Double-DQN target, Huber loss and d loss / d q for the wide network (trainer.py:163-187) -----------------
// q_sa (n,): Q of the chosen column; q_no / q_nt (n, A): online / target Q of next_state; mask_next (n, A).
// dq[i] = clamp(q_sa - y, -1, 1) / n_global (derivative of smooth_l1 with beta = 1, mean over the GLOBAL batch);
// loss[0] = sum_i huber_i / n_global, summed in a fixed order by one block.
__global__ void __launch_bounds__(1024)
pk_wide_dqn_dq_kernel(const float* __restrict__ q_sa, const float* __restrict__ q_no, const float* __restrict__ q_nt,
                      const int* __restrict__ mask_next, const float* __restrict__ reward,
                      const unsigned char* __restrict__ terminal, int n, int A, int n_global, float gamma, int double_dqn,
                      float* __restrict__ dq, float* __restrict__ loss) {
  __shared__ float sh[1024];
  float acc = 0.0f;
  for (int i = threadIdx.x; i < n; i += 1024) {
    float v = 0.0f;
    if (!terminal[i]) {
      float best = -CUDART_INF_F;
      int bi = 0;
      const float* sel = double_dqn ? q_no : q_nt;  // :171-182
      for (int c = 0; c < A; ++c) {
        const float x = mask_next[(size_t)i * A + c] != 0 ? sel[(size_t)i * A + c] : -CUDART_INF_F;
        if (x > best) {
          best = x;
          bi = c;
        }
      }
      v = double_dqn ? q_nt[(size_t)i * A + bi] : best;
    }
    const float y = fmaf(v, gamma, reward[i]);  // :184
    const float d = q_sa[i] - y;
    const float ad = fabsf(d);
    acc += ad < 1.0f ? 0.5f * d * d : ad - 0.5f;  // F.smooth_l1_loss, :187
    dq[i] = fminf(fmaxf(d, -1.0f), 1.0f) / (float)n_global;
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = sh[0] / (float)n_global;
}
