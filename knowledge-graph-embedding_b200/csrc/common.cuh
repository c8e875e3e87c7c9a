// common.cuh — shared device helpers for the sm_100a KGE kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "kge_abi.h"

namespace kge {

constexpr int WARP = 32;
constexpr unsigned FULL = 0xffffffffu;
constexpr int UPD_HEAVY = 48;     // table rows with more gradient records are summed by a whole CTA

// ---- error plumbing (abi.cu owns the thread-local string) -----------------------
void set_error(const char* fmt, ...);
#define KGE_CUDA_OK(expr)                                                          \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      kge::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),       \
                     __FILE__, __LINE__);                                          \
      return KGE_ECUDA;                                                            \
    }                                                                              \
  } while (0)
#define KGE_REQUIRE(cond, code, ...)                                               \
  do {                                                                             \
    if (!(cond)) { kge::set_error(__VA_ARGS__); return (code); }                   \
  } while (0)

// Per-step scalars of a captured step graph (kge_step_graph): written by the graph's first node
// (k_set_dyn, arguments patched per launch) and read by the kernels in place of their by-value copies.
struct DynScal {
  unsigned long long step;      // Philox step counter
  unsigned long long step_next; // ... of the batch grouped ahead (kge_train_step_next)
  float lr_t;                   // Adam: lr * sqrt(1 - b2^t) / (1 - b1^t); SGD / Adagrad: lr
  uint32_t epoch;               // cross-GPU barrier epoch of the data-parallel step
  // host-side ends of a captured step (kge_step_graph_launch with pinned host buffers): read / written by the
  // graph's own kernels through mapped host memory instead of by copies queued around the graph
  const void* next_src;         // the look-ahead batch (pinned host or device memory; NULL: already staged by a copy)
  float* host_loss;             // the loss goes here as soon as it is final (NULL: not wanted)
  unsigned long long* host_seq; // ... followed by `seq`: what kge_step_graph_wait_loss spins on
  unsigned long long seq;
};
// a batch whose gradient records were grouped ahead of its step (kge_train_step_next)
struct Prefetched {
  bool valid = false;
  const void* triples = nullptr; void* ws = nullptr;
  int64_t B = 0, i_base = 0, n_ent = 0;
  uint64_t seed = 0, step = 0;
  int k = 0, side = 0, sampler = 0, set = 0, graph = 0;
};
struct kge_ctx_impl {
  int device;
  int sm_count;
  int smem_optin;
  // optional per-phase timing: ev[call][0..KGE_N_PHASES] boundaries, mask of phases hit
  cudaEvent_t* ev = nullptr;
  unsigned* ev_mask = nullptr;
  int prof_cap = 0, prof_n = 0;
  int prof_open = -1;             // slot opened by kge_train_grad: the kge_train_apply that follows records into it
  // side stream of kge_train_step (sort keys are independent of the weights, so key generation,
  // the radix sort and the segment table overlap the step kernel) + its fork / join events
  cudaStream_t aux = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_mid = nullptr, ev_heavy = nullptr;
  int opt_radix_sort = 0;         // kge_set_option(KGE_OPT_RADIX_SORT)
  int opt_no_shadow = 0;          // kge_set_option(KGE_OPT_NO_SHADOW)
  int opt_no_fused_push = 0;      // kge_set_option(KGE_OPT_NO_FUSED_PUSH)
  int opt_ring_max = 32;          // kge_set_option(KGE_OPT_RING_SLOTS): deepest TMA row ring of the step kernel (0 = none)
  unsigned* ticket = nullptr;     // device, 64 words, zero between calls (self-resetting counters)
  cudaStream_t pre_stream = nullptr;               // the next batch's grouping chain (kge_train_step_next)
  cudaEvent_t ev_pre_fork = nullptr, ev_pre_done = nullptr;
  Prefetched pre;
  cudaEvent_t* gev = nullptr;     // capturing with phase events on (kge_step_graph, KGE_OPT_GRAPH_PROFILE): [2 * KGE_N_PHASES]
  unsigned gev_mask = 0;
  int opt_graph_profile = 0;
  int opt_cluster_group = 0;      // kge_set_option(KGE_OPT_CLUSTER_GROUP): grouping by k_group_cluster (one kernel, one cluster)
  int opt_pre_early = 1;          // kge_set_option(KGE_OPT_PRE_EARLY): 1 = the look-ahead grouping starts beside the step kernel
  int opt_tc_passes = 3;          // kge_set_option(KGE_OPT_RANK_TC_PASSES): MMAs per K step of the ranking GEMM (3 or 2)
  DynScal* dyn = nullptr;         // device; used only while `capturing` (kge_step_graph_create)
  int capturing = 0;              // the call is being captured into a CUDA graph: per-step scalars come from `dyn`
};
// RAII marker: records the boundary events of one call when profiling is on
struct PhaseTimer {
  kge_ctx_impl* c; cudaStream_t st; int slot; bool g = false;
  // join: continue the slot the preceding gradient half opened (one data-parallel step = one call)
  PhaseTimer(kge_ctx_impl* c_, cudaStream_t s, bool join = false) : c(c_), st(s), slot(-1) {
    if (c->capturing) { g = c->gev != nullptr; return; }      // a captured step records into the graph's own events
    if (!c->ev) return;
    if (join && c->prof_open >= 0) { slot = c->prof_open; c->prof_open = -1; return; }
    c->prof_open = -1;
    if (c->prof_n < c->prof_cap) { slot = c->prof_n++; c->ev_mask[slot] = 0; }
  }
  void keep_open() { if (slot >= 0) c->prof_open = slot; }
  void begin(int phase) { begin(phase, st); }
  void end(int phase) { end(phase, st); }
  void begin(int phase, cudaStream_t s) {
    if (slot >= 0) cudaEventRecord(c->ev[slot * 2 * KGE_N_PHASES + 2 * phase], s);
    else if (g) cudaEventRecordWithFlags(c->gev[2 * phase], s, cudaEventRecordExternal);    // a real event-record node
  }
  void end(int phase, cudaStream_t s) {
    if (slot >= 0) { cudaEventRecord(c->ev[slot * 2 * KGE_N_PHASES + 2 * phase + 1], s); c->ev_mask[slot] |= 1u << phase; }
    else if (g) { cudaEventRecordWithFlags(c->gev[2 * phase + 1], s, cudaEventRecordExternal); c->gev_mask |= 1u << phase; }
  }
};

// ---- warp reductions (fixed butterfly order: deterministic) ---------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
// the same butterflies over aligned groups of W lanes (W = 16: two independent rows per warp, step_narrow_kernels.cuh).
// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may
// become resident while its predecessor on the stream still runs; pdl_wait() returns once that predecessor has
// completed and its writes are visible (a no-op for an ordinary launch).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// With the upper half of a warp holding zeros, warp_sum and warp_sum_w<16> return the same bits in the lower half.
template <int W> __device__ __forceinline__ float warp_sum_w(float v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
template <int W> __device__ __forceinline__ float warp_max_w(float v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(FULL, v, o));
  return v;
}
__device__ __forceinline__ void warp_sum2(float& a, float& b) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(FULL, a, o);
    b += __shfl_xor_sync(FULL, b, o);
  }
}

// ---- ids -------------------------------------------------------------------------
__device__ __forceinline__ int64_t load_id(const void* p, int64_t i, int id64) {
  return id64 ? reinterpret_cast<const int64_t*>(p)[i]
              : (int64_t) reinterpret_cast<const int32_t*>(p)[i];
}

// ---- float4 helpers ---------------------------------------------------------------
__device__ __forceinline__ float4 f4(float a, float b, float c, float d) { return make_float4(a, b, c, d); }
__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) { return f4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 operator-(float4 a, float4 b) { return f4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
__device__ __forceinline__ float4 operator*(float4 a, float4 b) { return f4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 operator*(float s, float4 a) { return f4(s * a.x, s * a.y, s * a.z, s * a.w); }
__device__ __forceinline__ float4 operator-(float4 a) { return f4(-a.x, -a.y, -a.z, -a.w); }
__device__ __forceinline__ void operator+=(float4& a, float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
__device__ __forceinline__ float hsum(float4 a) { return (a.x + a.y) + (a.z + a.w); }
__device__ __forceinline__ float dot4(float4 a, float4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ float sgn(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

// A table row spread over a warp: lane l owns 16-byte chunks l, l+32, ... (NV of them).
template <int NV>
struct Row {
  float4 v[NV];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = f4zero();
  }
};

template <int NV>
__device__ __forceinline__ void load_row(Row<NV>& r, const float* __restrict__ base, int64_t row,
                                         int ld, int cols, int lane) {
  const float4* p = reinterpret_cast<const float4*>(base + row * (int64_t) ld);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = lane + i * WARP;
    r.v[i] = (4 * c < cols) ? __ldg(p + c) : f4zero();
  }
}
template <int NV>
__device__ __forceinline__ void store_row(const Row<NV>& r, float* __restrict__ base, int64_t row,
                                          int ld, int cols, int lane) {
  float4* p = reinterpret_cast<float4*>(base + row * (int64_t) ld);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int c = lane + i * WARP;
    if (4 * c < cols) p[c] = r.v[i];
  }
}
template <int NV>
__device__ __forceinline__ float row_dot_partial(const Row<NV>& a, const Row<NV>& b) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) s += dot4(a.v[i], b.v[i]);
  return s;
}
template <int NV>
__device__ __forceinline__ void row_axpy(Row<NV>& y, float a, const Row<NV>& x) {  // y += a*x
#pragma unroll
  for (int i = 0; i < NV; ++i) y.v[i] += a * x.v[i];
}
template <int NV>
__device__ __forceinline__ void row_add(Row<NV>& y, const Row<NV>& x) {
#pragma unroll
  for (int i = 0; i < NV; ++i) y.v[i] += x.v[i];
}
// mask elements with index >= cols (used where padding could leak: Linf ties, trunc)
template <int NV>
__device__ __forceinline__ void row_mask(Row<NV>& r, int cols, int lane) {
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    int e = 4 * (lane + i * WARP);
    if (e + 0 >= cols) r.v[i].x = 0.f;
    if (e + 1 >= cols) r.v[i].y = 0.f;
    if (e + 2 >= cols) r.v[i].z = 0.f;
    if (e + 3 >= cols) r.v[i].w = 0.f;
  }
}

// ---- Philox4x32-10 (Salmon et al. SC'11; spec in oracle/philox.py) ----------------
__device__ __forceinline__ uint32_t philox_word0(uint64_t seed, uint64_t step,
                                                          uint32_t side_bit, uint64_t n) {
  uint32_t c0 = (uint32_t) n, c1 = (uint32_t)(n >> 32), c2 = (uint32_t) step,
           c3 = (uint32_t)(((step >> 32) << 1) | side_bit);
  uint32_t k0 = (uint32_t) seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    if (r) { k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
    uint64_t p0 = (uint64_t) 0xD2511F53u * c0;
    uint64_t p1 = (uint64_t) 0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
  }
  return c0;
}

// Negative entity for positive i, slot c of its k_eff negatives (reference row order).
// Returns the entity id; *is_head tells which side was corrupted.
struct SamplerDev {
  int kind;                 // kge_sampler_kind
  int side_mode;            // kge_side
  int k_eff, k_side;
  uint64_t seed, step;
  const unsigned long long* step_dyn;   // nullable: the step counter lives on the device (captured step graph)
  int64_t n_ent;
  int64_t i_base;           // global index of this call's first positive (data-parallel ranks)
  kge_sampler s;
};
__device__ __forceinline__ bool neg_is_head(const SamplerDev& sd, int c) {
  return sd.side_mode == KGE_SIDE_H || (sd.side_mode == KGE_SIDE_HT && (c & 1) == 0);
}
__device__ __forceinline__ int64_t sample_negative(const SamplerDev& sd, int64_t i, int c,
                                                   int64_t h, int64_t t) {
  bool head = neg_is_head(sd, c);
  int jj = (sd.side_mode == KGE_SIDE_HT) ? (c >> 1) : c;
  uint64_t n = (uint64_t) (i + sd.i_base) * (uint64_t) sd.k_side + (uint64_t) jj;
  const uint64_t step = sd.step_dyn ? (uint64_t) __ldg(sd.step_dyn) : sd.step;
  const uint32_t u = philox_word0(sd.seed, step, head ? 0u : 1u, n);
  if (sd.kind == KGE_NS_UNIFORM) {      // 32-bit modulo: n_ent < 2^32 is checked on the host
    if (sd.s.pool == nullptr) return (int64_t) ((uint32_t) u % (uint32_t) sd.n_ent);
    return sd.s.pool[(uint32_t) u % (uint32_t) sd.s.pool_len];
  }
  int64_t x = head ? h : t;
  int64_t ty = sd.s.ent_type[x];
  int64_t lo = sd.s.type_off[ty], len = sd.s.type_off[ty + 1] - lo;
  int64_t q = sd.s.ent_pos[x];
  int64_t eff = (q >= 0) ? len - 1 : len;
  if (eff <= 0) return x;   // host rejects singleton pools before launch
  int64_t j = (int64_t) ((uint32_t) u % (uint32_t) eff);
  if (q >= 0 && j >= q) ++j;
  return sd.s.type_members[lo + j];
}

}  // namespace kge
