// Microbenchmark: throughput of small 1-D TMA bulk loads (cp.async.bulk global -> shared) as pk_dense_tc_kernel
// issues them: `bytes` per copy, source stride `stride` bytes between copies, `nw` issuing warps per CTA (lane 0 each),
// `depth` copies in flight per warp.  One CTA per SM.  Prints aggregate GB/s.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_load_bw bulk_load_bw.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(512, 1)
k(const unsigned char* src, size_t total, uint32_t bytes, size_t stride, int depth, long long per_warp, int mode) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // [nw][32]
  unsigned char* buf = smem + 8192;
  const uint32_t slot = (bytes + 127u) & ~127u;
  if (lane == 0)
    for (int d = 0; d < depth; ++d)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bars + warp * 32 + d)));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const long long w = (long long)blockIdx.x * nw + warp, W = (long long)gridDim.x * nw;
  if (mode == 0) {
    if (lane != 0) return;
    for (long long i = 0; i < per_warp; ++i) {
      const int d = (int)(i % depth);
      uint64_t* bar = bars + warp * 32 + d;
      if (i >= depth) {
        const uint32_t par = (uint32_t)((i / depth - 1) & 1);
        asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W;\n}" ::"r"(s32(bar)), "r"(par) : "memory");
      }
      const size_t off = ((size_t)((i * W + w)) * stride) % (total - 65536);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       s32(buf + ((size_t)warp * depth + d) * slot)), "l"(src + (off & ~(size_t)15)), "r"(bytes), "r"(s32(bar)) : "memory");
    }
    for (int d = 0; d < depth && d < per_warp; ++d) {
      const long long uses = (per_warp - d + depth - 1) / depth;
      const uint32_t par = (uint32_t)((uses - 1) & 1);
      asm volatile("{\n.reg .pred p;\nW2: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@!p bra W2;\n}" ::"r"(s32(bars + warp * 32 + d)), "r"(par) : "memory");
    }
  } else {
    // mode 1: the whole warp loads the copy with LDG.128 (lane l: chunks l, l + 32, ...), `depth` copies unrolled
    float4 acc = make_float4(0, 0, 0, 0);
    const uint32_t nch = bytes >> 4;
    for (long long i = 0; i < per_warp; i += 4) {
      float4 v[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const size_t off = (((size_t)(((i + u) * W + w)) * stride) % (total - 65536)) & ~(size_t)15;
#pragma unroll
        for (int c = 0; c < 3; ++c)
          v[u][c] = (lane + 32 * c < nch) ? __ldg(reinterpret_cast<const float4*>(src + off) + lane + 32 * c) : make_float4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int c = 0; c < 3; ++c) { acc.x += v[u][c].x; acc.y += v[u][c].y; acc.z += v[u][c].z; acc.w += v[u][c].w; }
    }
    if (acc.x + acc.y + acc.z + acc.w == 1.2345f) buf[threadIdx.x] = 1;
  }
}
int main(int argc, char** argv) {
  const size_t total = (size_t)(argc > 1 ? atoi(argv[1]) : 1536) << 20;  // MiB touched: <= ~100 stays in L2
  unsigned char* d;
  cudaMalloc(&d, total);
  cudaMemset(d, 0, total);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  struct Cfg { uint32_t bytes; size_t stride; int nw, depth, mode; } cfgs[] = {
      {1184, 22212, 1, 16, 0}, {1184, 22212, 4, 16, 0}, {1184, 22212, 16, 8, 0}, {1184, 22272, 16, 8, 0},
      {1152, 22272, 16, 8, 0}, {2368, 22212, 16, 4, 0},  {4736, 22212, 8, 4, 0},  {16384, 22272, 4, 2, 0},
      {2336, 22212, 1, 16, 0}, {2336, 22212, 2, 16, 0}, {2336, 22212, 14, 4, 0}, {4640, 22212, 14, 2, 0}, {592, 22212, 16, 8, 0},
      {1184, 22212, 16, 4, 1}, {1184, 22272, 16, 4, 1}};
  for (auto& c : cfgs) {
    const long long per_warp = (long long)((size_t)1 << 30) / c.bytes / (148 * c.nw) / 4 * 4;
    const size_t smem = 8192 + (size_t)c.nw * c.depth * ((c.bytes + 127) & ~127u);
    if (smem > 200 * 1024) { printf("skip\n"); continue; }
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      k<<<148, c.nw * 32, smem>>>(d, total, c.bytes, c.stride, c.depth, per_warp, c.mode);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double gb = (double)per_warp * 148 * c.nw * c.bytes / 1e9;
    printf("mode=%d bytes=%u stride=%zu warps=%d depth=%d: %.1f GB/s (%.3f ms) %s\n", c.mode, c.bytes, c.stride, c.nw, c.depth,
           gb / (ms * 1e-3), ms, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
