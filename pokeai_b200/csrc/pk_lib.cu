// C ABI of pokeai_b200 (see include/pokeai_b200.h).  Single translation unit: the kernels live in the
// .cuh files included below so that the __constant__ tables are shared without -rdc.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "pk_common.cuh"
#include "pk_decide2.cuh"
#include "pk_update.cuh"
#include "pk_bn.cuh"
#include "pk_wide.cuh"
#include "pk_decide_tc.cuh"
#include "pk_dense_tc.cuh"

// ---- error handling -----------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int pk_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define PK_CUDA(call)                                                                            \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) return pk_fail(PK_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
  } while (0)
#define PK_REQUIRE(cond, ...) \
  do {                        \
    if (!(cond)) return pk_fail(PK_ERR_INVALID, __VA_ARGS__); \
  } while (0)

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- global state (one device per process, as in one-process-per-GPU deployments) -----------------
struct PkGlobal {
  int device = -1;
  int sm_count = 0;
  int max_smem = 0;
  PkDeviceTables tab;
  // tuning knobs (environment: PK_DECIDE_WARPS / PK_ENCODE_WARPS / PK_CTAS_PER_SM / PK_ZERO_TMA)
  int decide_warps = 16;
  int encode_warps = 16;
  int decide_pair = 1;
  int decide_tc = 0;  // 64-channel nets: hidden layers on tcgen05 (PK_DECIDE_TC: 0 never (measured best), 1 when features are written, 2 always)
  int hidden_tc = 0;
  // programmatic dependent launch (the record-only prologue of a pk_decide / pk_encode launch overlaps the tail of the
  // previous kernel of the stream): OFF for the public device-pointer entry points unless the caller opts in
  // (pk_set_overlap_launches / PK_PDL=1: the records must not be produced by the preceding kernel), ON for the host-buffer
  // sessions, whose records always arrive from host memory (PK_PDL_SESSION=0 turns that off)
  int pdl = 0;
  int pdl_session = 1;
  int dense_tc = 1;  // 64-channel nets: the dense-input forward on tcgen05 (PK_DENSE_TC; 0 = FP32-pipe kernel)
  int dense_warps = 16;  // warps per CTA of the dense-input forward (PK_DENSE_WARPS; kernel iteration only)
  int zero_copy = 1;  // host-buffer calls: kernels read / write pinned host buffers directly (PK_ZERO_COPY)
  int peer_timeout_ms = 30000;  // pk_peer_clamp_adam: how long a rank waits for its peers' gradients (PK_PEER_TIMEOUT_MS)
  int upd_fused = 1;  // pk_dqn_grad: online + target next-state pass in one kernel (PK_UPDATE_FUSED)
  int ctas_per_sm = 1;
  int zero_by_tma = 1;
  unsigned long long* trace = nullptr;  // debug phase trace buffer (pk_debug_set_trace)
  // side stream of the wide backward: the HBM-bound column sums run beside the tensor-bound GEMMs (PK_WIDE_SIDE)
  int wide_tma_store = 1;  // wide GEMM: bf16 outputs through shared memory + TMA stores (PK_WIDE_TMA_STORE)
  int wide_side = 1;
  int wide_2cta = 1;  // wide GEMM: CTA-pair (cta_group::2) form for the K-major GEMMs (PK_WIDE_2CTA)
  int wide_fused_head = 1;  // pk_wide_forward: head folded into the last GEMM's epilogue (PK_WIDE_FUSED_HEAD)
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // pk_dqn_grad_static: device address of pk_static_scratch on `device` (cudaGetSymbolAddress in pk_init) and the
  // return code of that query
  float* static_scratch = nullptr;
  int static_scratch_device = -1;
  cudaError_t static_scratch_rc = cudaErrorNotReady;
};
static PkGlobal g;

// Scratch of pk_dqn_grad_static, the entry point that allocates nothing: file-scope device memory, held by every GPU
// that loads the library until the process exits.  In HOST code this name is only a stand-in for the symbol, not a
// device address: kernels get the pointer cudaGetSymbolAddress returned for the current device (pk_init).
static __device__ __align__(256) float pk_static_scratch[PK_DQN_STATIC_SCRATCH_FLOATS];

static int env_int(const char* name, int dflt, int lo, int hi) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  int x = atoi(v);
  if (x < lo) x = lo;
  if (x > hi) x = hi;
  return x;
}

extern "C" int pk_abi_version(void) { return PK_ABI_VERSION; }
extern "C" const char* pk_last_error(void) { return g_err; }

extern "C" int pk_init(const pk_tables* t, int device) {
  PK_REQUIRE(t != nullptr && t->species_type_mask != nullptr, "pk_init: null tables");
  PK_REQUIRE(t->n_species > 0 && t->n_species <= 256, "pk_init: n_species %d not in 1..256 (u8 index)", t->n_species);
  PK_REQUIRE(t->n_items > 0 && t->n_items <= 256, "pk_init: n_items %d not in 1..256 (u8 index)", t->n_items);
  PK_REQUIRE(t->n_moves > 0 && t->n_moves <= 65536, "pk_init: n_moves %d not in 1..65536 (u16 index)", t->n_moves);
  PK_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  PK_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return pk_fail(PK_ERR_UNSUPPORTED, "pk_init: device %d is sm_%d%d; this library is built for sm_100a only", device,
                   prop.major, prop.minor);
  memset(&g.tab, 0, sizeof(g.tab));
  g.tab.n_species = t->n_species;
  g.tab.n_moves = t->n_moves;
  g.tab.n_items = t->n_items;
  g.tab.row_move = PK_N_STATE + PK_CH_SPECIES + t->n_species;
  g.tab.row_item = g.tab.row_move + t->n_moves;
  g.tab.C = g.tab.row_item + t->n_items;
  for (int i = 0; i < t->n_species; ++i) g.tab.type_mask[i] = t->species_type_mask[i] & ((1u << PK_N_TYPES) - 1u);
  PK_CUDA(cudaMemcpyToSymbol(c_tab, &g.tab, sizeof(g.tab)));
  g.device = device;
  {  // once per device, with that device current; a failure only disables pk_dqn_grad_static, which reports it
    void* p = nullptr;
    g.static_scratch_rc = cudaGetSymbolAddress(&p, pk_static_scratch);
    if (g.static_scratch_rc != cudaSuccess) {
      (void)cudaGetLastError();
      p = nullptr;
    }
    g.static_scratch = reinterpret_cast<float*>(p);
    g.static_scratch_device = device;
  }
  g.sm_count = prop.multiProcessorCount;
  g.max_smem = (int)prop.sharedMemPerBlockOptin;
  g.decide_warps = env_int("PK_DECIDE_WARPS", 16, 1, 32);
  g.encode_warps = env_int("PK_ENCODE_WARPS", 16, 1, 16);
  g.ctas_per_sm = env_int("PK_CTAS_PER_SM", 1, 1, 4);
  g.zero_by_tma = env_int("PK_ZERO_TMA", 1, 0, 1);
  g.hidden_tc = env_int("PK_HIDDEN_TC", 0, 0, 1);  // 64-channel hidden layers on the 3xTF32 mma.sync path (measured: no gain)
  g.pdl = env_int("PK_PDL", 0, 0, 1);
  g.upd_fused = env_int("PK_UPDATE_FUSED", 1, 0, 1);
  g.pdl_session = env_int("PK_PDL_SESSION", 1, 0, 1);
  g.peer_timeout_ms = env_int("PK_PEER_TIMEOUT_MS", 30000, 1, 3600000);
  g.zero_copy = env_int("PK_ZERO_COPY", 1, 0, 2);
  g.dense_warps = env_int("PK_DENSE_WARPS", 16, 1, 16);
  g.wide_side = env_int("PK_WIDE_SIDE", 1, 0, 1);
  g.wide_fused_head = env_int("PK_WIDE_FUSED_HEAD", 1, 0, 1);
  g.wide_2cta = env_int("PK_WIDE_2CTA", 1, 0, 1);
  g.wide_tma_store = env_int("PK_WIDE_TMA_STORE", 1, 0, 1);
  g.decide_pair = env_int("PK_DECIDE_PAIR", 1, 0, 2);  // 0 never, 1 when features are not written, 2 always
  g.decide_tc = env_int("PK_DECIDE_TC", 0, 0, 2);
  g.dense_tc = env_int("PK_DENSE_TC", 1, 0, 1);
  return PK_OK;
}

#define PK_CHECK_INIT() \
  do {                  \
    if (g.device < 0) return pk_fail(PK_ERR_NOT_INIT, "pk_init has not been called"); \
  } while (0)

extern "C" int pk_debug_set_trace(unsigned long long* device_buffer) {
  g.trace = device_buffer;
  return PK_OK;
}
extern "C" int pk_set_overlap_launches(int on) {
  g.pdl = on ? 1 : 0;
  return PK_OK;
}
extern "C" int pk_debug_set_hidden_tc(int on) {
  g.hidden_tc = on ? 1 : 0;
  return PK_OK;
}
extern "C" int pk_debug_set_dense_tc(int on) {
  g.dense_tc = on ? 1 : 0;
  return PK_OK;
}
extern "C" int pk_debug_set_decide_tc(int mode) {
  if (mode < 0 || mode > 2) return pk_fail(PK_ERR_INVALID, "pk_debug_set_decide_tc: mode %d not in 0..2", mode);
  g.decide_tc = mode;
  return PK_OK;
}
extern "C" int pk_debug_set_pairing(int mode) {
  if (mode < 0 || mode > 2) return pk_fail(PK_ERR_INVALID, "pk_debug_set_pairing: mode %d not in 0..2", mode);
  g.decide_pair = mode;
  return PK_OK;
}
extern "C" int pk_n_channels(void) { return g.device < 0 ? 0 : g.tab.C; }
extern "C" int pk_sm_count(void) { return g.device < 0 ? 0 : g.sm_count; }

// ---- weights --------------------------------------------------------------------------------------
extern "C" size_t pk_mlp_prepared_floats(int n_layers, int n_channels, int in_channels) {
  if (n_layers < 1 || n_layers > PK_MAX_LAYERS || n_channels < 1 || in_channels < 1) return 0;
  return pk_make_layout(n_layers, n_channels, in_channels).total_all;
}
extern "C" size_t pk_mlp_param_floats(int n_layers, int n_channels, int in_channels) {
  if (n_layers < 1 || n_layers > PK_MAX_LAYERS || n_channels < 1 || in_channels < 1) return 0;
  return pk_make_param_layout(n_layers, n_channels, in_channels).total;
}

static int pk_prepare_impl(const pk_mlp* m, float* prepared, cudaStream_t st) {
  const PkLayout lay = pk_make_layout(m->n_layers, m->n_channels, m->in_channels);
  const int H = m->n_channels, HP = lay.HP, T = 256;
  for (int l = 0; l < m->n_layers; ++l) {
    PK_REQUIRE(m->w[l] != nullptr && m->b[l] != nullptr, "pk_mlp_prepare: null tensor in layer %d", l);
    const int cin = l == 0 ? m->in_channels : H;
    const int rows = l == 0 ? cin + 1 : HP;
    float* dst = prepared + (l == 0 ? lay.w1t : lay.wl[l]);
    const long long n = (long long)rows * HP;
    pk_transpose_pad_kernel<<<(unsigned)((n + T - 1) / T), T, 0, st>>>(m->w[l], dst, H, cin, rows, HP);
    pk_copy_pad_kernel<<<(HP + T - 1) / T, T, 0, st>>>(m->b[l], prepared + lay.b[l], H, HP);
    if (l > 0)  // untransposed padded copy for the backward pass
      pk_pad2d_kernel<<<(unsigned)(((long long)HP * HP + T - 1) / T), T, 0, st>>>(m->w[l], prepared + lay.wn[l], H, H, HP);
    if (l > 0 && lay.tc[l])  // tcgen05 operand tiles (hi / lo, swizzled) for pk_decide_tc_kernel
      pk_tc_tiles_kernel<<<(64 * 64 + T - 1) / T, T, 0, st>>>(m->w[l], prepared + lay.tc[l], H, H);
    if (l == 0 && lay.tc0) {  // layer 0 as tcgen05 operand tiles for pk_dense_tc_kernel
      const int nkb = ((cin + 63) >> 6) * 2;
      pk_tc0_tiles_kernel<<<(64 * 32 * nkb + T - 1) / T, T, 0, st>>>(m->w[0], prepared + lay.tc0, H, cin, nkb);
    }
  }
  PK_REQUIRE(m->w_out != nullptr && m->b_out != nullptr, "pk_mlp_prepare: null output tensor");
  pk_copy_pad_kernel<<<(HP + T - 1) / T, T, 0, st>>>(m->w_out, prepared + lay.wout, H, HP);
  pk_copy_pad_kernel<<<1, 32, 0, st>>>(m->b_out, prepared + lay.bout, 1, 4);
  PK_CUDA(cudaGetLastError());
  return PK_OK;
}

extern "C" int pk_mlp_prepare(const pk_mlp* m, float* prepared, void* stream) {
  PK_CHECK_INIT();
  PK_REQUIRE(m != nullptr && prepared != nullptr, "pk_mlp_prepare: null argument");
  PK_REQUIRE(m->n_layers >= 1 && m->n_layers <= PK_MAX_LAYERS, "pk_mlp_prepare: n_layers %d not in 1..%d", m->n_layers,
             PK_MAX_LAYERS);
  PK_REQUIRE(m->n_channels >= 1 && m->in_channels >= 1, "pk_mlp_prepare: bad channel counts");
  PK_REQUIRE(aligned16(prepared), "pk_mlp_prepare: prepared buffer must be 16-byte aligned");
  return pk_prepare_impl(m, prepared, (cudaStream_t)stream);
}

// ---- launch helpers ---------------------------------------------------------------------------------
template <typename K>
static int pk_set_smem(K kernel, unsigned bytes) {
  PK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return PK_OK;
}

// Launch with (optionally) programmatic dependent launch: the kernel's read-only prologue may overlap the tail of the
// previous kernel of the stream; it does griddepcontrol.wait before its first global write.
template <typename... KArgs, typename... Args>
static cudaError_t pk_launch_pdl(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, bool pdl,
                                 Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define PK_DISPATCH_CPL_AT(CPL_, AT_, MACRO)                                 \
  do {                                                                       \
    if (CPL_ == 1 && AT_ == 4) { MACRO(1, 4); }                              \
    else if (CPL_ == 1 && AT_ == 6) { MACRO(1, 6); }                         \
    else if (CPL_ == 1 && AT_ == 9) { MACRO(1, 9); }                         \
    else if (CPL_ == 2 && AT_ == 4) { MACRO(2, 4); }                         \
    else if (CPL_ == 2 && AT_ == 6) { MACRO(2, 6); }                         \
    else if (CPL_ == 2 && AT_ == 9) { MACRO(2, 9); }                         \
    else if (CPL_ == 4 && AT_ == 4) { MACRO(4, 4); }                         \
    else if (CPL_ == 4 && AT_ == 6) { MACRO(4, 6); }                         \
    else if (CPL_ == 4 && AT_ == 9) { MACRO(4, 9); }                         \
    else return pk_fail(PK_ERR_UNSUPPORTED, "no kernel for %d channels per lane / %d action slots", CPL_, AT_); \
  } while (0)

static inline int pk_at_for(int A) { return A <= 4 ? 4 : (A <= 6 ? 6 : 9); }

static int pk_decide_impl(const pk_obs* packed, int B, int A, const float* prepared, int n_layers, int n_channels,
                          float* feat, int32_t* mask, float* q, int32_t* action, const pk_select_params* sel,
                          cudaStream_t st, bool pdl) {
  PK_CHECK_INIT();
  PK_REQUIRE(B >= 0, "negative batch");
  PK_REQUIRE(A >= 1 && A <= PK_MAX_ACTIONS, "A=%d not in 1..%d", A, PK_MAX_ACTIONS);
  if (B == 0) return PK_OK;
  PK_REQUIRE(packed != nullptr && aligned16(packed), "packed observations must be non-null and 16-byte aligned");
  PK_REQUIRE(feat == nullptr || (reinterpret_cast<uintptr_t>(feat) & 3u) == 0, "feature buffer must be 4-byte aligned");
  const bool with_mlp = prepared != nullptr;
  PkLayout lay;
  PkSelect ps;
  memset(&ps, 0, sizeof(ps));
  ps.temperature = 1.0f;
  ps.trace = g.trace;
  if (with_mlp) {
    PK_REQUIRE(n_layers >= 1 && n_layers <= PK_MAX_LAYERS && n_channels >= 1, "bad network shape %d x %d", n_layers,
               n_channels);
    PK_REQUIRE(aligned16(prepared), "prepared weights must be 16-byte aligned");
    PK_REQUIRE(q != nullptr || action != nullptr, "pk_decide: nothing to compute (q and action both null)");
    lay = pk_make_layout(n_layers, n_channels, g.tab.C);
    if (sel != nullptr) {
      ps.mode = sel->mode;
      ps.eps = sel->eps;
      ps.temperature = sel->temperature;
      ps.u_explore = sel->u_explore;
      ps.u_pick = sel->u_pick;
      ps.probs = sel->probs;
      PK_REQUIRE(ps.mode >= PK_SELECT_ARGMAX && ps.mode <= PK_SELECT_SOFTMAX, "unknown selection mode %d", ps.mode);
      if (ps.mode == PK_SELECT_EPSGREEDY)
        PK_REQUIRE(ps.u_explore != nullptr && ps.u_pick != nullptr, "epsilon-greedy needs u_explore and u_pick");
      if (ps.mode == PK_SELECT_SOFTMAX)
        PK_REQUIRE(ps.u_pick != nullptr && ps.temperature > 0.0f, "softmax sampling needs u_pick and temperature > 0");
    }
  } else {
    PK_REQUIRE(feat != nullptr || mask != nullptr, "pk_encode: nothing to compute (feat and mask both null)");
    lay = pk_make_layout(1, 32, g.tab.C);
  }
  const int CPL = lay.HP / 32;
  const bool use_tma = feat != nullptr && g.zero_by_tma != 0;
  const uint4* pk = reinterpret_cast<const uint4*>(packed);
  // 64-channel networks with one or two hidden layers: the hidden layers of the four battles of a warp group run as
  // one tcgen05 contraction (pk_decide_tc.cuh)
  if (with_mlp && lay.HP == 64 && n_layers >= 2 && (g.decide_tc == 2 || (g.decide_tc == 1 && feat != nullptr))) {
    const PktSmemPlan tplan = pkt_smem_plan(lay, A, use_tma, (unsigned)g.max_smem);
    if (tplan.total <= (unsigned)g.max_smem) {
      int tgrid = (B + PKT_WARPS - 1) / PKT_WARPS;
      if (tgrid > g.sm_count) tgrid = g.sm_count;
#define PK_LAUNCH_TC(A_)                                                                                          \
  {                                                                                                               \
    int rc_ = pk_set_smem(pk_decide_tc_kernel<A_>, tplan.total);                                                  \
    if (rc_) return rc_;                                                                                          \
    PK_CUDA(pk_launch_pdl(pk_decide_tc_kernel<A_>, tgrid, PKT_WARPS * 32, tplan.total, st, pdl, pk, B,            \
                          prepared, lay, feat, mask, q, action, ps, use_tma ? 1 : 0, (unsigned)g.max_smem));      \
  }
      switch (A) {
        case 1: PK_LAUNCH_TC(1) break;
        case 2: PK_LAUNCH_TC(2) break;
        case 3: PK_LAUNCH_TC(3) break;
        case 4: PK_LAUNCH_TC(4) break;
        case 5: PK_LAUNCH_TC(5) break;
        case 6: PK_LAUNCH_TC(6) break;
        case 7: PK_LAUNCH_TC(7) break;
        case 8: PK_LAUNCH_TC(8) break;
        default: PK_LAUNCH_TC(9) break;
      }
#undef PK_LAUNCH_TC
      PK_CUDA(cudaGetLastError());
      return PK_OK;
    }
  }
  // measured (profiles/r01_pair_vs_single.txt): pairing wins when the features are not materialised (305 vs 275 M
  // decisions/s at 65 536 states) and loses when they are (two 22 KB blocks per warp back-pressure the store path)
  if (with_mlp && lay.HP <= 64 && (g.decide_pair == 2 || (g.decide_pair == 1 && feat == nullptr))) {
    // paired kernel (pk_decide2.cuh): a warp takes two battles, one per half-warp in the MLP
    int nw = g.decide_warps > 16 ? 16 : g.decide_warps;
    PkSmemPlan plan2 = pk2_smem_plan(lay, A, nw, use_tma);
    while (plan2.total > (unsigned)g.max_smem && nw > 1) plan2 = pk2_smem_plan(lay, A, --nw, use_tma);
    if (plan2.total <= (unsigned)g.max_smem) {
      const long long NP = ((long long)B + 1) / 2;
      int grid2 = (int)((NP + nw - 1) / nw);
      if (grid2 > g.sm_count * g.ctas_per_sm) grid2 = g.sm_count * g.ctas_per_sm;
#define PK_LAUNCH_PAIR(CH_, A_)                                                                           \
  {                                                                                                       \
    int rc_ = pk_set_smem(pk_decide2_kernel<CH_, A_>, plan2.total);                                       \
    if (rc_) return rc_;                                                                                  \
    pk_decide2_kernel<CH_, A_><<<grid2, nw * 32, plan2.total, st>>>(pk, B, prepared, lay, feat, mask, q,  \
                                                                    action, ps, use_tma ? 1 : 0);         \
  }
#define PK_LAUNCH_PAIR_A(CH_)                                       \
  switch (A) {                                                      \
    case 1: PK_LAUNCH_PAIR(CH_, 1) break;                           \
    case 2: PK_LAUNCH_PAIR(CH_, 2) break;                           \
    case 3: PK_LAUNCH_PAIR(CH_, 3) break;                           \
    case 4: PK_LAUNCH_PAIR(CH_, 4) break;                           \
    case 5: PK_LAUNCH_PAIR(CH_, 5) break;                           \
    case 6: PK_LAUNCH_PAIR(CH_, 6) break;                           \
    case 7: PK_LAUNCH_PAIR(CH_, 7) break;                           \
    case 8: PK_LAUNCH_PAIR(CH_, 8) break;                           \
    default: PK_LAUNCH_PAIR(CH_, 9) break;                          \
  }
      if (lay.HP == 32) { PK_LAUNCH_PAIR_A(2) }
      else { PK_LAUNCH_PAIR_A(4) }
#undef PK_LAUNCH_PAIR_A
#undef PK_LAUNCH_PAIR
      PK_CUDA(cudaGetLastError());
      return PK_OK;
    }
  }
  int nwarps = with_mlp ? g.decide_warps : g.encode_warps;
  if (nwarps > 16) nwarps = 16;  // the 128-channel instantiation needs > 64 registers per thread
  bool tc = with_mlp && lay.HP == 64 && n_layers >= 2 && g.hidden_tc != 0;
  if (tc && pk_smem_plan(lay, A, nwarps, with_mlp, use_tma, true).total > (unsigned)g.max_smem) tc = false;
  PkSmemPlan plan = pk_smem_plan(lay, A, nwarps, with_mlp, use_tma, tc);
  while (plan.total > (unsigned)g.max_smem && nwarps > 1) {  // drop warps until the plan fits
    --nwarps;
    plan = pk_smem_plan(lay, A, nwarps, with_mlp, use_tma, tc);
  }
  if (plan.total > (unsigned)g.max_smem)
    return pk_fail(PK_ERR_UNSUPPORTED,
                   "network %d x %d does not fit the fused small-net kernel (%u B shared memory needed, %d available)",
                   n_layers, n_channels, plan.total, g.max_smem);
  int grid = (B + nwarps - 1) / nwarps;
  if (grid > g.sm_count * g.ctas_per_sm) grid = g.sm_count * g.ctas_per_sm;
#define PK_LAUNCH_DECIDE(C_, A_)                                                                              \
  {                                                                                                           \
    if constexpr (C_ == 2) {                                                                                  \
      if (tc) {                                                                                               \
        int rc_ = pk_set_smem(pk_decide_kernel<2, A_, 512, true>, plan.total);                                \
        if (rc_) return rc_;                                                                                  \
        PK_CUDA(pk_launch_pdl(pk_decide_kernel<2, A_, 512, true>, grid, nwarps * 32, plan.total, st,          \
                              pdl, pk, B, prepared, lay, feat, mask, q, action, ps, use_tma ? 1 : 0));        \
        break;                                                                                                \
      }                                                                                                       \
    }                                                                                                         \
    int rc_ = pk_set_smem(pk_decide_kernel<C_, A_, 512>, plan.total);                                         \
    if (rc_) return rc_;                                                                                      \
    PK_CUDA(pk_launch_pdl(pk_decide_kernel<C_, A_, 512>, grid, nwarps * 32, plan.total, st, pdl, pk,          \
                          B, prepared, lay, feat, mask, q, action, ps, use_tma ? 1 : 0));                     \
  }
#define PK_LAUNCH_DECIDE_A(C_)                                        \
  switch (A) {                                                        \
    case 1: PK_LAUNCH_DECIDE(C_, 1) break;                            \
    case 2: PK_LAUNCH_DECIDE(C_, 2) break;                            \
    case 3: PK_LAUNCH_DECIDE(C_, 3) break;                            \
    case 4: PK_LAUNCH_DECIDE(C_, 4) break;                            \
    case 5: PK_LAUNCH_DECIDE(C_, 5) break;                            \
    case 6: PK_LAUNCH_DECIDE(C_, 6) break;                            \
    case 7: PK_LAUNCH_DECIDE(C_, 7) break;                            \
    case 8: PK_LAUNCH_DECIDE(C_, 8) break;                            \
    default: PK_LAUNCH_DECIDE(C_, 9) break;                           \
  }
  if (CPL == 1) { PK_LAUNCH_DECIDE_A(1) }
  else if (CPL == 2) { PK_LAUNCH_DECIDE_A(2) }
  else if (CPL == 4) { PK_LAUNCH_DECIDE_A(4) }
  else return pk_fail(PK_ERR_UNSUPPORTED, "the fused small-net kernel supports up to 128 channels (got %d)", n_channels);
#undef PK_LAUNCH_DECIDE_A
#undef PK_LAUNCH_DECIDE
  PK_CUDA(cudaGetLastError());
  return PK_OK;
}

extern "C" int pk_encode(const pk_obs* packed, int B, int A, float* feat, int32_t* mask, void* stream) {
  return pk_decide_impl(packed, B, A, nullptr, 0, 0, feat, mask, nullptr, nullptr, nullptr, (cudaStream_t)stream,
                        g.pdl != 0);
}

extern "C" int pk_decide(const pk_obs* packed, int B, int A, const float* prepared, int n_layers, int n_channels,
                         float* feat, int32_t* mask, float* q, int32_t* action, const pk_select_params* sel,
                         void* stream) {
  PK_REQUIRE(prepared != nullptr, "pk_decide: null prepared weights");
  return pk_decide_impl(packed, B, A, prepared, n_layers, n_channels, feat, mask, q, action, sel, (cudaStream_t)stream,
                        g.pdl != 0);
}

extern "C" int pk_mlp_forward(const float* prepared, int n_layers, int n_channels, int in_channels, const float* feat,
                              int B, int A, float* q, void* stream) {
  PK_CHECK_INIT();
  PK_REQUIRE(B >= 0, "negative batch");
  PK_REQUIRE(A >= 1 && A <= PK_MAX_ACTIONS, "A=%d not in 1..%d", A, PK_MAX_ACTIONS);
  if (B == 0) return PK_OK;
  PK_REQUIRE(prepared != nullptr && feat != nullptr && q != nullptr, "pk_mlp_forward: null argument");
  PK_REQUIRE(aligned16(prepared), "prepared weights must be 16-byte aligned");
  PK_REQUIRE(n_layers >= 1 && n_layers <= PK_MAX_LAYERS && n_channels >= 1 && in_channels >= 1, "bad network shape");
  const PkLayout lay = pk_make_layout(n_layers, n_channels, in_channels);
  cudaStream_t st = (cudaStream_t)stream;
  if (g.dense_tc && lay.tc0 && n_layers >= 2) {
    // 64-channel networks: every layer on the tensor cores (pk_dense_tc.cuh)
    const PkdSmemPlan plan = pkd_smem_plan(lay, A, (unsigned)g.max_smem);
    if (plan.x_stages >= 2) {
      int rc = pk_set_smem(pk_dense_tc_kernel, plan.total);
      if (rc) return rc;
      const int nb = pkd_battles_per_tile(A), npairs = ((B + nb - 1) / nb + 1) / 2;  // a CTA works on pairs of tiles
      const int tgrid = npairs < g.sm_count ? npairs : g.sm_count;
      pk_dense_tc_kernel<<<tgrid, PKD_WARPS * 32, plan.total, st>>>(feat, B, A, prepared, lay, q, (unsigned)g.max_smem,
                                                                    g.trace);
      PK_CUDA(cudaGetLastError());
      return PK_OK;
    }
  }
  const int CPL = lay.HP / 32, AT = pk_at_for(A);
  const unsigned tail_bytes = (lay.total - lay.b[0]) * 4u;
  const unsigned per_warp = ((unsigned)AT * lay.HP + PK_DENSE_DEPTH * 32u * PK_DENSE_XS(AT)) * 4u;
  int nwarps = g.dense_warps;
  while (nwarps > 1 && tail_bytes + per_warp * nwarps > (unsigned)g.max_smem) --nwarps;
  if (tail_bytes + per_warp * nwarps > (unsigned)g.max_smem)
    return pk_fail(PK_ERR_UNSUPPORTED, "network %d x %d does not fit the small-net forward kernel", n_layers, n_channels);
  // the rest of shared memory caches the first rows of W0^T (all of them when they fit)
  int cache_rows = (int)(((unsigned)g.max_smem - tail_bytes - per_warp * nwarps) / (lay.HP * 4u));
  if (cache_rows > in_channels) cache_rows = in_channels;
  const unsigned smem = tail_bytes + per_warp * nwarps + (unsigned)cache_rows * lay.HP * 4u;
  int grid = (B + nwarps - 1) / nwarps;
  if (grid > g.sm_count) grid = g.sm_count;
#define PK_LAUNCH_DENSE(C_, A_)                                                                  \
  {                                                                                              \
    int rc_ = pk_set_smem(pk_dense_forward_kernel<C_, A_>, smem);                                \
    if (rc_) return rc_;                                                                         \
    pk_dense_forward_kernel<C_, A_><<<grid, nwarps * 32, smem, st>>>(feat, B, A, prepared, lay, cache_rows, q); \
  }
  PK_DISPATCH_CPL_AT(CPL, AT, PK_LAUNCH_DENSE);
#undef PK_LAUNCH_DENSE
  PK_CUDA(cudaGetLastError());
  return PK_OK;
}

extern "C" int pk_select(const float* q, const int32_t* mask, int B, int A, const pk_select_params* sel,
                         int32_t* action, void* stream) {
  PK_CHECK_INIT();
  PK_REQUIRE(B >= 0, "negative batch");
  PK_REQUIRE(A >= 1 && A <= PK_MAX_ACTIONS, "A=%d not in 1..%d", A, PK_MAX_ACTIONS);
  if (B == 0) return PK_OK;
  PK_REQUIRE(q != nullptr && mask != nullptr && action != nullptr, "pk_select: null argument");
  PkSelect ps;
  memset(&ps, 0, sizeof(ps));
  ps.temperature = 1.0f;
  if (sel != nullptr) {
    ps.mode = sel->mode;
    ps.eps = sel->eps;
    ps.temperature = sel->temperature;
    ps.u_explore = sel->u_explore;
    ps.u_pick = sel->u_pick;
    ps.probs = sel->probs;
  }
  PK_REQUIRE(ps.mode >= PK_SELECT_ARGMAX && ps.mode <= PK_SELECT_SOFTMAX, "unknown selection mode %d", ps.mode);
  if (ps.mode == PK_SELECT_EPSGREEDY)
    PK_REQUIRE(ps.u_explore != nullptr && ps.u_pick != nullptr, "epsilon-greedy needs u_explore and u_pick");
  if (ps.mode == PK_SELECT_SOFTMAX)
    PK_REQUIRE(ps.u_pick != nullptr && ps.temperature > 0.0f, "softmax sampling needs u_pick and temperature > 0");
  pk_select_kernel<<<(B + 255) / 256, 256, 0, (cudaStream_t)stream>>>(q, mask, B, A, ps, action);
  PK_CUDA(cudaGetLastError());
  return PK_OK;
}

// ---- DQN update ---------------------------------------------------------------------------------------
#include "pk_update_api.inc"
#include "pk_bn_api.inc"

// pk_dqn_grad without caller scratch and without any allocation: the scratch is pk_static_scratch (above).  No
// counterpart in the reference; added for callers that cannot allocate.
// Nothing in the scratch has to start at zero: every word pk_dqn_grad reads it has written in the same call (a*,
// next-state values, one gradient partial per launched CTA; the atomics fallback clears its sum with a memset node on
// the stream), so the previous call's contents are simply overwritten.
extern "C" int pk_dqn_grad_static(const float* prepared, const float* prepared_target, int n_layers, int n_channels, int A,
                                  const pk_batch* batch, float gamma, int double_dqn, float* grad, float* loss,
                                  void* stream) {
  PK_CHECK_INIT();
  PK_REQUIRE(batch != nullptr, "pk_dqn_grad_static: null batch");
  PK_REQUIRE(batch->n >= 1, "pk_dqn_grad_static: need n >= 1 (got %d)", batch->n);
  PK_REQUIRE(n_layers >= 1 && n_layers <= PK_UPD_MAX_LAYERS && n_channels >= 1, "pk_dqn_grad_static: bad network shape");
  if (g.static_scratch == nullptr || g.static_scratch_device != g.device)
    return pk_fail(PK_ERR_CUDA, "pk_dqn_grad_static: the static scratch has no device address on device %d (%s)", g.device,
                   cudaGetErrorString(g.static_scratch_rc));
  const size_t need = pk_scratch_layout(pk_make_layout(n_layers, n_channels, g.tab.C), batch->n).total;
  if (need > (size_t)PK_DQN_STATIC_SCRATCH_FLOATS)
    return pk_fail(PK_ERR_INVALID,
                   "pk_dqn_grad_static: %d transitions on a %d x %d network need %zu scratch floats, the static scratch "
                   "holds %zu; use pk_dqn_grad with caller-owned scratch",
                   batch->n, n_layers, n_channels, need, (size_t)PK_DQN_STATIC_SCRATCH_FLOATS);
  return pk_dqn_grad(prepared, prepared_target, n_layers, n_channels, A, batch, gamma, double_dqn, grad, loss,
                     g.static_scratch, stream);
}

// ---- host-buffer session ------------------------------------------------------------------------------
struct pk_session {
  int max_batch, A, materialise;
  int n_layers, n_channels;
  cudaStream_t stream;
  pk_obs* d_packed;
  float* d_feat;
  int32_t* d_mask;
  float* d_q;
  int32_t* d_action;
  float* d_u_explore;
  float* d_u_pick;
  float* d_params;
  float* d_prepared;        // owned copy (pk_session_load_weights)
  const float* prepared;    // what the kernels read: d_prepared or the caller's buffer
};

extern "C" void pk_session_destroy(pk_session* s) {
  if (!s) return;
  cudaFree(s->d_packed);
  cudaFree(s->d_feat);
  cudaFree(s->d_mask);
  cudaFree(s->d_q);
  cudaFree(s->d_action);
  cudaFree(s->d_u_explore);
  cudaFree(s->d_u_pick);
  cudaFree(s->d_params);
  cudaFree(s->d_prepared);
  if (s->stream) cudaStreamDestroy(s->stream);
  free(s);
}

extern "C" int pk_session_create(pk_session** out, int max_batch, int A, int materialise_features) {
  PK_CHECK_INIT();
  PK_REQUIRE(out != nullptr, "pk_session_create: null out");
  PK_REQUIRE(max_batch >= 1, "pk_session_create: max_batch must be positive");
  PK_REQUIRE(A >= 1 && A <= PK_MAX_ACTIONS, "A=%d not in 1..%d", A, PK_MAX_ACTIONS);
  pk_session* s = (pk_session*)calloc(1, sizeof(pk_session));
  if (!s) return pk_fail(PK_ERR_INVALID, "out of host memory");
  s->max_batch = max_batch;
  s->A = A;
  s->materialise = materialise_features;
  cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  const size_t nb = (size_t)max_batch;
  if (e == cudaSuccess) e = cudaMalloc(&s->d_packed, nb * sizeof(pk_obs));
  if (e == cudaSuccess && materialise_features) e = cudaMalloc(&s->d_feat, nb * g.tab.C * A * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_mask, nb * A * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_q, nb * A * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_action, nb * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_u_explore, nb * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&s->d_u_pick, nb * sizeof(float));
  if (e != cudaSuccess) {
    pk_session_destroy(s);
    return pk_fail(PK_ERR_CUDA, "pk_session_create: %s", cudaGetErrorString(e));
  }
  *out = s;
  return PK_OK;
}

extern "C" int pk_session_load_weights(pk_session* s, const pk_mlp* hm) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr && hm != nullptr, "pk_session_load_weights: null argument");
  PK_REQUIRE(hm->n_layers >= 1 && hm->n_layers <= PK_MAX_LAYERS && hm->n_channels >= 1, "bad network shape");
  PK_REQUIRE(hm->in_channels == g.tab.C, "in_channels %d != feature channels %d", hm->in_channels, g.tab.C);
  const PkParamLayout P = pk_make_param_layout(hm->n_layers, hm->n_channels, hm->in_channels);
  const PkLayout lay = pk_make_layout(hm->n_layers, hm->n_channels, hm->in_channels);
  cudaFree(s->d_params);
  cudaFree(s->d_prepared);
  s->d_params = s->d_prepared = nullptr;
  s->prepared = nullptr;
  PK_CUDA(cudaMalloc(&s->d_params, (size_t)P.total * sizeof(float)));
  PK_CUDA(cudaMalloc(&s->d_prepared, (size_t)lay.total_all * sizeof(float)));
  pk_mlp dm = *hm;
  const int H = hm->n_channels;
  for (int l = 0; l < hm->n_layers; ++l) {
    PK_REQUIRE(hm->w[l] && hm->b[l], "null tensor in layer %d", l);
    const size_t nw = (size_t)H * (l == 0 ? hm->in_channels : H);
    PK_CUDA(cudaMemcpyAsync(s->d_params + P.w[l], hm->w[l], nw * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    PK_CUDA(cudaMemcpyAsync(s->d_params + P.b[l], hm->b[l], H * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    dm.w[l] = s->d_params + P.w[l];
    dm.b[l] = s->d_params + P.b[l];
  }
  PK_REQUIRE(hm->w_out && hm->b_out, "null output tensor");
  PK_CUDA(cudaMemcpyAsync(s->d_params + P.wout, hm->w_out, H * sizeof(float), cudaMemcpyHostToDevice, s->stream));
  PK_CUDA(cudaMemcpyAsync(s->d_params + P.bout, hm->b_out, sizeof(float), cudaMemcpyHostToDevice, s->stream));
  dm.w_out = s->d_params + P.wout;
  dm.b_out = s->d_params + P.bout;
  int rc = pk_prepare_impl(&dm, s->d_prepared, s->stream);
  if (rc) return rc;
  PK_CUDA(cudaStreamSynchronize(s->stream));
  s->n_layers = hm->n_layers;
  s->n_channels = hm->n_channels;
  s->prepared = s->d_prepared;
  return PK_OK;
}

extern "C" int pk_session_use_prepared(pk_session* s, const float* prepared_dev, int n_layers, int n_channels) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr && prepared_dev != nullptr, "pk_session_use_prepared: null argument");
  PK_REQUIRE(n_layers >= 1 && n_layers <= PK_MAX_LAYERS && n_channels >= 1, "bad network shape");
  PK_REQUIRE(aligned16(prepared_dev), "prepared weights must be 16-byte aligned");
  s->prepared = prepared_dev;
  s->n_layers = n_layers;
  s->n_channels = n_channels;
  return PK_OK;
}

// Device-side alias of a pinned (page-locked, mapped) host buffer, or null for pageable memory.
static void* pk_pinned_alias(const void* host) {
  if (host == nullptr) return nullptr;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  if (at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) return nullptr;
  return at.devicePointer;
}

extern "C" int pk_session_wait(pk_session* s) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr, "pk_session_wait: null session");
  PK_CUDA(cudaStreamSynchronize(s->stream));
  return PK_OK;
}

extern "C" int pk_session_decide_host(pk_session* s, const pk_obs* packed_host, int B, int mode, float eps,
                                      float temperature, const float* u_explore_host, const float* u_pick_host,
                                      float* q_host, int32_t* action_host) {
  int rc = pk_session_submit_host(s, packed_host, B, mode, eps, temperature, u_explore_host, u_pick_host, q_host,
                                  action_host);
  if (rc) return rc;
  return pk_session_wait(s);
}

// Everything pk_session_decide_host does except the final synchronise: the copies and the kernel are queued on the
// session's stream and the call returns.  A driver with two sessions keeps two batches in flight (one being packed /
// read on the host while the other runs); results are valid after pk_session_wait.
extern "C" int pk_session_submit_host(pk_session* s, const pk_obs* packed_host, int B, int mode, float eps,
                                      float temperature, const float* u_explore_host, const float* u_pick_host,
                                      float* q_host, int32_t* action_host) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr && packed_host != nullptr, "pk_session_submit_host: null argument");
  PK_REQUIRE(s->prepared != nullptr, "pk_session_submit_host: no weights loaded");
  PK_REQUIRE(B >= 0 && B <= s->max_batch, "batch %d exceeds the session capacity %d", B, s->max_batch);
  if (B == 0) return PK_OK;
  // Zero-copy: when the caller's buffers are pinned the kernel reads the records over PCIe itself (each is read once,
  // one battle ahead of its use) and stores Q / actions straight into host memory: one launch and one synchronise
  // instead of three copies around it.  Pageable buffers take the staged path.
  const pk_obs* in_dev = s->d_packed;
  float* q_dev = s->d_q;
  int32_t* act_dev = s->d_action;
  bool q_direct = false, act_direct = false;
  if (g.zero_copy) {
    if (g.zero_copy == 1)  // 2 = results only: the records go through the copy engine, which overlaps another stream's kernel
      if (void* p = pk_pinned_alias(packed_host)) in_dev = static_cast<const pk_obs*>(p);
    if (void* p = pk_pinned_alias(q_host)) {
      q_dev = static_cast<float*>(p);
      q_direct = true;
    }
    if (void* p = pk_pinned_alias(action_host)) {
      act_dev = static_cast<int32_t*>(p);
      act_direct = true;
    }
  }
  if (in_dev == s->d_packed)
    PK_CUDA(cudaMemcpyAsync(s->d_packed, packed_host, (size_t)B * sizeof(pk_obs), cudaMemcpyHostToDevice, s->stream));
  pk_select_params sp;
  memset(&sp, 0, sizeof(sp));
  sp.mode = mode;
  sp.eps = eps;
  sp.temperature = temperature;
  if (mode == PK_SELECT_EPSGREEDY) {
    PK_REQUIRE(u_explore_host != nullptr, "epsilon-greedy needs u_explore");
    PK_CUDA(cudaMemcpyAsync(s->d_u_explore, u_explore_host, (size_t)B * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    sp.u_explore = s->d_u_explore;
  }
  if (mode != PK_SELECT_ARGMAX) {
    PK_REQUIRE(u_pick_host != nullptr, "this selection mode needs u_pick");
    PK_CUDA(cudaMemcpyAsync(s->d_u_pick, u_pick_host, (size_t)B * sizeof(float), cudaMemcpyHostToDevice, s->stream));
    sp.u_pick = s->d_u_pick;
  }
  int rc = pk_decide_impl(in_dev, B, s->A, s->prepared, s->n_layers, s->n_channels, s->d_feat, s->d_mask, q_dev, act_dev,
                          &sp, s->stream, g.pdl_session != 0);
  if (rc) return rc;
  if (q_host && !q_direct)
    PK_CUDA(cudaMemcpyAsync(q_host, s->d_q, (size_t)B * s->A * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
  if (action_host && !act_direct)
    PK_CUDA(cudaMemcpyAsync(action_host, s->d_action, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  return PK_OK;
}

extern "C" int pk_session_fetch(pk_session* s, int B, float* feat_host, int32_t* mask_host) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr, "pk_session_fetch: null session");
  PK_REQUIRE(B >= 0 && B <= s->max_batch, "batch %d exceeds the session capacity %d", B, s->max_batch);
  if (B == 0) return PK_OK;
  if (feat_host) {
    PK_REQUIRE(s->d_feat != nullptr, "this session does not materialise features");
    PK_CUDA(cudaMemcpyAsync(feat_host, s->d_feat, (size_t)B * g.tab.C * s->A * sizeof(float), cudaMemcpyDeviceToHost,
                            s->stream));
  }
  if (mask_host)
    PK_CUDA(cudaMemcpyAsync(mask_host, s->d_mask, (size_t)B * s->A * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  PK_CUDA(cudaStreamSynchronize(s->stream));
  return PK_OK;
}

extern "C" int pk_session_encode_host(pk_session* s, const pk_obs* packed_host, int B, float* feat_host,
                                      int32_t* mask_host) {
  PK_CHECK_INIT();
  PK_REQUIRE(s != nullptr && packed_host != nullptr, "pk_session_encode_host: null argument");
  PK_REQUIRE(s->d_feat != nullptr, "this session does not materialise features");
  PK_REQUIRE(B >= 0 && B <= s->max_batch, "batch %d exceeds the session capacity %d", B, s->max_batch);
  if (B == 0) return PK_OK;
  PK_CUDA(cudaMemcpyAsync(s->d_packed, packed_host, (size_t)B * sizeof(pk_obs), cudaMemcpyHostToDevice, s->stream));
  int rc = pk_decide_impl(s->d_packed, B, s->A, nullptr, 0, 0, s->d_feat, s->d_mask, nullptr, nullptr, nullptr, s->stream,
                          g.pdl_session != 0);
  if (rc) return rc;
  return pk_session_fetch(s, B, feat_host, mask_host);
}

extern "C" const float* pk_session_features(const pk_session* s) { return s ? s->d_feat : nullptr; }
extern "C" const int32_t* pk_session_mask(const pk_session* s) { return s ? s->d_mask : nullptr; }
#include "pk_wide_api.inc"
