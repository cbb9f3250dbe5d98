// Tower-gradient average fused with the optimizer, over NVLink peer memory.  No NCCL on this path.
//
// Replaces MirroredStrategy's packed NCCL all-reduce + per-mirror ApplyAdam
// (reference: model.py:115-119,462-467; log line "batch_all_reduce ... nccl, num_packs = 1",
// Test.ipynb:200).  Every rank maps every peer's gradient bucket, parameter bucket and signal pad
// (cudaIpc handles exchanged once at start-up), then one kernel per bucket does:
//
//   one-shot (small buckets): barrier -> each rank reads all N peers' gradients (16-byte peer
//       loads), sums in a fixed rank order in fp32, applies 1/N and the optimizer to its own
//       replica -> barrier -> clears its gradient.  One signal round-trip of latency.
//   two-shot (large buckets): barrier -> rank r reduces slice r from all peers (reduce-scatter),
//       runs the optimizer on that slice only (optimizer state is touched shard-wise), and stores
//       the updated fp32 weights straight into every peer's parameter bucket (all-gather) ->
//       barrier -> clears its gradient.  Wire bytes per GPU and direction: (N-1)/N * bucket.
//
// Signals are per-(CTA, source-rank) epoch counters written with st.release.sys and polled with
// ld.acquire.sys; epochs only ever grow, so there is no reset and no ABA, and they live in device
// memory so a captured CUDA graph replays correctly.  A spin that exceeds the communicator's time
// limit (wall-clock nanoseconds from %globaltimer, default = the host-side collective timeout
// TDL_COLLECTIVE_TIMEOUT_S, so a slow input pipeline on one tower does not kill the job after a
// couple of seconds) writes {rank, CTA, peer, epoch} into a host-mapped status word and traps, so
// the Python side can say which tower went missing (failure detection instead of a silent hang).
#include "tdl_common.cuh"

namespace tdl {
namespace p2p {

constexpr int kMaxWorld = 8;
constexpr int kMaxBlocks = 256;
constexpr int kThreads = 512;

struct Comm {
  int world, rank;
  float* grad[kMaxWorld];      // peer-mapped gradient buckets (fp32 or bf16 storage, same offsets everywhere)
  float* param[kMaxWorld];     // peer-mapped fp32 master weights
  uint32_t* flags[kMaxWorld];  // peer-mapped signal pads: [kMaxBlocks][kMaxWorld] u32
  uint32_t* epoch;             // local: [kMaxBlocks] u32, last epoch used by each CTA
  int emulate;                 // 1: all "peers" live on this GPU and ranks are launched one after another
  float* mc_grad;              // NVLS multicast address of the gradient buckets (nullptr when not available)
  float* mc_param;             // NVLS multicast address of the parameter buckets
  bf16* wire[kMaxWorld];       // peer-mapped bf16 "wire" copies of the gradient buckets (same element offsets), or null
  unsigned long long timeout_ns;  // how long a CTA waits for a peer before it reports and traps
  int* status;                 // host-mapped: [0] = 1 when a wait timed out, [1..4] = rank, CTA, peer, epoch
  uint32_t* debug_seen;        // debug builds of the protocol (TDL_P2P_DEBUG=1): [kMaxBlocks][kMaxWorld] last value seen
};

enum OptKind { kNone = 0, kSgdMomentum = 1, kAdam = 2 };

struct StepArgs {
  int64_t offset;   // first element of the bucket inside the flat buffers
  int64_t n;        // elements in the bucket
  int kind;
  int grad_is_bf16;
  int zero_grad;
  float* m1;        // momentum / Adam m (local, full-size flat buffer)
  float* m2;        // Adam v
  const float* lr_state;
  float a, b, c;    // sgd: momentum, weight_decay, -    adam: beta1, beta2, eps
  float grad_scale; // applied on top of the 1/N tower mean
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long ptx_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// CTA-level barrier across ranks: CTA b of every rank meets CTA b of every other rank.
__device__ __forceinline__ void cross_rank_barrier(const Comm& cm, uint32_t epoch) {
  __syncthreads();
  if (!cm.emulate && threadIdx.x < cm.world) {
    const int peer = threadIdx.x;
    st_release_sys(cm.flags[peer] + (size_t)blockIdx.x * kMaxWorld + cm.rank, epoch);
    const uint32_t* mine = cm.flags[cm.rank] + (size_t)blockIdx.x * kMaxWorld + peer;
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    if (cm.debug_seen) {
      // protocol assertions (race detection for the signal pads): what peer `peer` wrote for this CTA must
      // never go backwards, and a peer can be at most ONE barrier ahead of us (it cannot pass barrier e+1 before we
      // arrive at it).  A violation means a lost / reordered / foreign write into the pad: report it and stop.
      uint32_t* seen = cm.debug_seen + (size_t)blockIdx.x * kMaxWorld + peer;
      for (;;) {
        const uint32_t v = ld_acquire_sys(mine);
        const bool backwards = (int32_t)(v - *seen) < 0;
        const bool too_far = (int32_t)(v - epoch) > 1;
        if (backwards || too_far) {
          if (cm.status) {
            cm.status[1] = cm.rank; cm.status[2] = (int)blockIdx.x; cm.status[3] = peer; cm.status[4] = (int)v;
            __threadfence_system();
            cm.status[0] = backwards ? 2 : 3;
            __threadfence_system();
          }
          printf("tdl_b200 p2p DEBUG: rank %d CTA %d peer %d: signal %u after %u while waiting for epoch %u (%s)\n",
                 cm.rank, (int)blockIdx.x, peer, v, *seen, epoch, backwards ? "went backwards" : "too far ahead");
          __trap();
        }
        *seen = v;
        if ((int32_t)(v - epoch) >= 0) break;
        if ((++spins & 1023u) == 0) {
          const unsigned long long now = ptx_globaltimer();
          if (t0 == 0) t0 = now;
          if (now - t0 > cm.timeout_ns) break;  // fall through to the timeout report below
        }
      }
    }
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 1023u) != 0) continue;  // look at the wall clock every 1024 polls
      const unsigned long long now = ptx_globaltimer();
      if (t0 == 0) t0 = now;
      if (now - t0 > cm.timeout_ns) {
        if (cm.status) {
          cm.status[1] = cm.rank; cm.status[2] = (int)blockIdx.x; cm.status[3] = peer; cm.status[4] = (int)epoch;
          __threadfence_system();
          cm.status[0] = 1;
          __threadfence_system();
        }
        printf("tdl_b200 p2p: rank %d CTA %d timed out waiting for rank %d (epoch %u)\n", cm.rank, (int)blockIdx.x,
               peer, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_grad4(const Comm& cm, int r, int64_t i, int is_bf16) {
  if (is_bf16) {
    const uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(cm.grad[r]) + i);
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
    return make_float4(a.x, a.y, b.x, b.y);
  }
  return *reinterpret_cast<const float4*>(cm.grad[r] + i);
}
// Sum of `U` 4-element vectors (element offsets i[u]) over all ranks in rank order.  All U*world peer loads are
// issued before the first add so that several NVLink round trips are in flight per thread.
template <int U>
__device__ __forceinline__ void gather_sum(const Comm& cm, const int64_t* i, int nvalid, int is_bf16, float4* acc) {
#pragma unroll
  for (int u = 0; u < U; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < kMaxWorld; ++r) {
    if (r < cm.world) {
      float4 g[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        g[u] = u < nvalid ? ld_grad4(cm, r, i[u], is_bf16) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u].x += g[u].x; acc[u].y += g[u].y; acc[u].z += g[u].z; acc[u].w += g[u].w;
      }
    }
  }
}

__device__ __forceinline__ float ld_grad1(const Comm& cm, int r, int64_t i, int is_bf16) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const bf16*>(cm.grad[r])[i]) : cm.grad[r][i];
}

__device__ __forceinline__ float opt_update(const StepArgs& a, float w, float g, int64_t i, float lr, float lr_t) {
  if (a.kind == kSgdMomentum) {
    g += a.b * w;
    const float m = a.a * a.m1[i] + g;
    a.m1[i] = m;
    return w - lr * m;
  }
  if (a.kind == kAdam) {
    const float m = a.a * a.m1[i] + (1.f - a.a) * g;
    const float v = a.b * a.m2[i] + (1.f - a.b) * g * g;
    a.m1[i] = m;
    a.m2[i] = v;
    return w - lr_t * m / (sqrtf(v) + a.c);
  }
  return g;  // kNone: the "parameter" is the mean itself
}

// Clearing the gradient must mirror the READ pattern: the second cross-rank barrier is per CTA index, so CTA b may
// only clear elements that CTA b of the peers read.  `first`/`count` are in 4-element vectors; vector k of the range
// belongs to global thread (k mod nthr) -- exactly how the reduction loops assign vectors to threads.
__device__ __forceinline__ void zero_vectors(const Comm& cm, const StepArgs& a, int64_t first_vec_elem, int64_t count) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  if (a.grad_is_bf16) {
    bf16* g = reinterpret_cast<bf16*>(cm.grad[cm.rank]) + first_vec_elem;
    for (int64_t v = tid; v < count; v += nthr) *reinterpret_cast<uint2*>(g + v * 4) = make_uint2(0u, 0u);
  } else {
    float* g = cm.grad[cm.rank] + first_vec_elem;
    for (int64_t v = tid; v < count; v += nthr) *reinterpret_cast<float4*>(g + v * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
__device__ __forceinline__ void zero_tail(const Comm& cm, const StepArgs& a, int64_t first, int64_t hi) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // the tail is read by global threads 0..2
  if (first + tid < hi) {
    if (a.grad_is_bf16) reinterpret_cast<bf16*>(cm.grad[cm.rank])[first + tid] = __float2bfloat16(0.f);
    else cm.grad[cm.rank][first + tid] = 0.f;
  }
}

// ---- one-shot ----
__global__ void __launch_bounds__(kThreads)
one_shot_kernel(const Comm cm, const StepArgs a) {
  uint32_t epoch = cm.epoch[blockIdx.x];
  cross_rank_barrier(cm, ++epoch);  // every rank's gradients for this bucket are complete
  const float lr = a.lr_state ? a.lr_state[0] : 0.f;
  const float lr_t = a.lr_state ? a.lr_state[1] : 0.f;
  const float scale = a.grad_scale / (float)cm.world;
  const int64_t lo = a.offset, hi = a.offset + a.n;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  float* out = a.kind == kNone ? nullptr : cm.param[cm.rank];
  // bucket offsets are multiples of 4 elements (16 bytes fp32 / 8 bytes bf16)
  const int64_t n4 = a.n / 4;
  // kNone writes the mean into the gradient buffer itself, which peers are still reading: stage in registers
  // across the barrier is impossible for a grid-stride loop, so kNone results go to a.m1 (scratch) first.
  float* mean_out = a.kind == kNone ? a.m1 : nullptr;
  constexpr int U = 4;
  for (int64_t v0 = tid; v0 < n4; v0 += nthr * U) {
    int64_t idx[U];
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx[u] = lo + (v0 + (int64_t)u * nthr) * 4;
      if (v0 + (int64_t)u * nthr < n4) nvalid = u + 1;
    }
    float4 acc[U];
    gather_sum<U>(cm, idx, nvalid, a.grad_is_bf16, acc);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u >= nvalid) break;
      const int64_t i = idx[u];
      float4 s = acc[u];
      s.x *= scale; s.y *= scale; s.z *= scale; s.w *= scale;
      if (a.kind == kNone) {
        *reinterpret_cast<float4*>(mean_out + i) = s;
      } else {
        float4 w = *reinterpret_cast<const float4*>(out + i);
        w.x = opt_update(a, w.x, s.x, i, lr, lr_t);
        w.y = opt_update(a, w.y, s.y, i + 1, lr, lr_t);
        w.z = opt_update(a, w.z, s.z, i + 2, lr, lr_t);
        w.w = opt_update(a, w.w, s.w, i + 3, lr, lr_t);
        *reinterpret_cast<float4*>(out + i) = w;
      }
    }
  }
  for (int64_t i = lo + n4 * 4 + tid; i < hi; i += nthr) {
    float s = 0.f;
    for (int r = 0; r < cm.world; ++r) s += ld_grad1(cm, r, i, a.grad_is_bf16);
    s *= scale;
    if (a.kind == kNone) mean_out[i] = s;
    else out[i] = opt_update(a, out[i], s, i, lr, lr_t);
  }
  cross_rank_barrier(cm, ++epoch);  // all peers have finished reading my gradients
  if (a.kind == kNone) {
    // publish the mean in place (fp32 buckets only)
    float* g = cm.grad[cm.rank];
    for (int64_t v = tid; v < n4; v += nthr)  // same vector -> thread assignment as the reads above
      *reinterpret_cast<float4*>(g + lo + v * 4) = *reinterpret_cast<const float4*>(mean_out + lo + v * 4);
    if (lo + n4 * 4 + tid < hi) g[lo + n4 * 4 + tid] = mean_out[lo + n4 * 4 + tid];
  } else if (a.zero_grad) {
    zero_vectors(cm, a, lo, n4);
    zero_tail(cm, a, lo + n4 * 4, hi);
  }
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

// ---- two-shot: reduce-scatter -> shard-local optimizer -> all-gather by peer stores ----
__global__ void __launch_bounds__(kThreads)
two_shot_kernel(const Comm cm, const StepArgs a) {
  uint32_t epoch = cm.epoch[blockIdx.x];
  cross_rank_barrier(cm, ++epoch);
  const float lr = a.lr_state[0];
  const float lr_t = a.lr_state[1];
  const float scale = a.grad_scale / (float)cm.world;
  // slice boundaries in units of 4 elements so every access stays 16-byte aligned
  const int64_t n4 = a.n / 4;
  const int64_t per = (n4 + cm.world - 1) / cm.world;
  const int64_t v_lo = per * cm.rank < n4 ? per * cm.rank : n4;
  const int64_t v_hi = v_lo + per < n4 ? v_lo + per : n4;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  constexpr int U = 4;
  for (int64_t v0 = v_lo + tid; v0 < v_hi; v0 += nthr * U) {
    int64_t idx[U];
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx[u] = a.offset + (v0 + (int64_t)u * nthr) * 4;
      if (v0 + (int64_t)u * nthr < v_hi) nvalid = u + 1;
    }
    float4 acc[U];
    gather_sum<U>(cm, idx, nvalid, a.grad_is_bf16, acc);
    float4 wv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u >= nvalid) break;
      const int64_t i = idx[u];
      float4 w = *reinterpret_cast<const float4*>(cm.param[cm.rank] + i);
      w.x = opt_update(a, w.x, acc[u].x * scale, i, lr, lr_t);
      w.y = opt_update(a, w.y, acc[u].y * scale, i + 1, lr, lr_t);
      w.z = opt_update(a, w.z, acc[u].z * scale, i + 2, lr, lr_t);
      w.w = opt_update(a, w.w, acc[u].w * scale, i + 3, lr, lr_t);
      wv[u] = w;
    }
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < cm.world) {
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (u < nvalid) *reinterpret_cast<float4*>(cm.param[r] + idx[u]) = wv[u];
      }
    }
  }
  // ragged tail (< 4 elements): owned by the last rank
  if (cm.rank == cm.world - 1) {
    for (int64_t i = a.offset + n4 * 4 + tid; i < a.offset + a.n; i += nthr) {
      float s = 0.f;
      for (int r = 0; r < cm.world; ++r) s += ld_grad1(cm, r, i, a.grad_is_bf16);
      const float w = opt_update(a, cm.param[cm.rank][i], s * scale, i, lr, lr_t);
      for (int r = 0; r < cm.world; ++r) cm.param[r][i] = w;
    }
  }
  __threadfence_system();
  cross_rank_barrier(cm, ++epoch);  // every slice has landed everywhere; every gradient has been read
  if (a.zero_grad) {
    // slice s of MY bucket was read by rank s with the same (CTA, thread) -> vector assignment used above
    for (int sl = 0; sl < cm.world; ++sl) {
      const int64_t s_lo = per * sl < n4 ? per * sl : n4;
      const int64_t s_hi = s_lo + per < n4 ? s_lo + per : n4;
      zero_vectors(cm, a, a.offset + s_lo * 4, s_hi - s_lo);
    }
    zero_tail(cm, a, a.offset + n4 * 4, a.offset + a.n);
  }
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

// ---- bf16 wire format (the fused cast of the north star) ----
// Gradients are accumulated in fp32 (weight-gradient kernels red.add into the bucket) but cross NVLink as bf16:
//   phase 0 (local): every rank packs its whole fp32 bucket into its bf16 wire buffer and, in the same pass, clears the
//                    fp32 bucket for the next step -- peers only ever read the wire copy, so the clear no longer has
//                    to wait behind the second barrier;
//   barrier;
//   phase 1: rank r reduces slice r from every rank's wire copy (16-byte loads = 8 gradients, half the reduce-scatter
//            bytes of the fp32 path), accumulates in fp32 in rank order, applies 1/N and the optimizer to the fp32 master
//            slice and stores the updated fp32 weights into every peer's parameter buffer; barrier.
// Work is split in 8-element vectors; vector v of slice s belongs to global thread (v - s_lo) mod nthr in BOTH phases,
// so the per-CTA cross-rank barrier covers exactly the data a CTA's peers-with-the-same-index are about to read.
__device__ __forceinline__ void unpack_bf16x8(const uint4& raw, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

__device__ __forceinline__ void wire_slice(const Comm& cm, int64_t n8, int sl, int64_t* lo, int64_t* hi) {
  const int64_t per = (n8 + cm.world - 1) / cm.world;
  const int64_t a = per * sl < n8 ? per * sl : n8;
  *lo = a;
  *hi = a + per < n8 ? a + per : n8;
}

template <int U>
__global__ void __launch_bounds__(kThreads)
two_shot_wire_kernel(const Comm cm, const StepArgs a, int phases) {
  uint32_t epoch = cm.epoch[blockIdx.x];
  const int64_t n8 = a.n / 8;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  if (phases & 1) {
    float* g = cm.grad[cm.rank] + a.offset;
    bf16* wr = cm.wire[cm.rank] + a.offset;
    for (int sl = 0; sl < cm.world; ++sl) {
      int64_t lo, hi;
      wire_slice(cm, n8, sl, &lo, &hi);
      for (int64_t v = lo + tid; v < hi; v += nthr) {
        const float4 g0 = *reinterpret_cast<const float4*>(g + v * 8);
        const float4 g1 = *reinterpret_cast<const float4*>(g + v * 8 + 4);
        __nv_bfloat162 h[4] = {__floats2bfloat162_rn(g0.x, g0.y), __floats2bfloat162_rn(g0.z, g0.w),
                               __floats2bfloat162_rn(g1.x, g1.y), __floats2bfloat162_rn(g1.z, g1.w)};
        *reinterpret_cast<uint4*>(wr + v * 8) = *reinterpret_cast<const uint4*>(h);
        if (a.zero_grad) {
          *reinterpret_cast<float4*>(g + v * 8) = make_float4(0.f, 0.f, 0.f, 0.f);
          *reinterpret_cast<float4*>(g + v * 8 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
    // ragged tail (< 8 elements): read by the last rank's global threads 0..6
    if (n8 * 8 + tid < a.n) {
      wr[n8 * 8 + tid] = __float2bfloat16(g[n8 * 8 + tid]);
      if (a.zero_grad) g[n8 * 8 + tid] = 0.f;
    }
    __threadfence_system();
  }
  if (!(phases & 2)) return;
  cross_rank_barrier(cm, ++epoch);  // every rank's wire copy of this bucket is complete
  const float lr = a.lr_state[0];
  const float lr_t = a.lr_state[1];
  const float scale = a.grad_scale / (float)cm.world;
  int64_t v_lo, v_hi;
  wire_slice(cm, n8, cm.rank, &v_lo, &v_hi);
  for (int64_t v0 = v_lo + tid; v0 < v_hi; v0 += nthr * U) {
    int64_t idx[U];
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx[u] = a.offset + (v0 + (int64_t)u * nthr) * 8;
      if (v0 + (int64_t)u * nthr < v_hi) nvalid = u + 1;
    }
    float acc[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[u][j] = 0.f;
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < cm.world) {
        uint4 raw[U];
#pragma unroll
        for (int u = 0; u < U; ++u)  // all loads of this peer are issued before the first add
          raw[u] = u < nvalid ? *reinterpret_cast<const uint4*>(cm.wire[r] + idx[u]) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          float f[8];
          unpack_bf16x8(raw[u], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][j] += f[j];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u >= nvalid) break;
      const int64_t i = idx[u];
      float4 w0 = *reinterpret_cast<const float4*>(cm.param[cm.rank] + i);
      float4 w1 = *reinterpret_cast<const float4*>(cm.param[cm.rank] + i + 4);
      w0.x = opt_update(a, w0.x, acc[u][0] * scale, i, lr, lr_t);
      w0.y = opt_update(a, w0.y, acc[u][1] * scale, i + 1, lr, lr_t);
      w0.z = opt_update(a, w0.z, acc[u][2] * scale, i + 2, lr, lr_t);
      w0.w = opt_update(a, w0.w, acc[u][3] * scale, i + 3, lr, lr_t);
      w1.x = opt_update(a, w1.x, acc[u][4] * scale, i + 4, lr, lr_t);
      w1.y = opt_update(a, w1.y, acc[u][5] * scale, i + 5, lr, lr_t);
      w1.z = opt_update(a, w1.z, acc[u][6] * scale, i + 6, lr, lr_t);
      w1.w = opt_update(a, w1.w, acc[u][7] * scale, i + 7, lr, lr_t);
#pragma unroll
      for (int r = 0; r < kMaxWorld; ++r) {
        if (r < cm.world) {
          *reinterpret_cast<float4*>(cm.param[r] + i) = w0;
          *reinterpret_cast<float4*>(cm.param[r] + i + 4) = w1;
        }
      }
    }
  }
  if (cm.rank == cm.world - 1) {  // ragged tail: owned by the last rank
    for (int64_t i = a.offset + n8 * 8 + tid; i < a.offset + a.n; i += nthr) {
      float s = 0.f;
      for (int r = 0; r < cm.world; ++r) s += __bfloat162float(cm.wire[r][i]);
      const float w = opt_update(a, cm.param[cm.rank][i], s * scale, i, lr, lr_t);
      for (int r = 0; r < cm.world; ++r) cm.param[r][i] = w;
    }
  }
  __threadfence_system();
  cross_rank_barrier(cm, ++epoch);  // every slice has landed everywhere; every wire copy has been read
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

// Small buckets with the bf16 wire format: every rank reduces everything (one signal round trip, no second hop).
__global__ void __launch_bounds__(kThreads)
one_shot_wire_kernel(const Comm cm, const StepArgs a, int phases) {
  uint32_t epoch = cm.epoch[blockIdx.x];
  const int64_t n8 = a.n / 8;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  if (phases & 1) {
    float* g = cm.grad[cm.rank] + a.offset;
    bf16* wr = cm.wire[cm.rank] + a.offset;
    for (int64_t v = tid; v < n8; v += nthr) {
      const float4 g0 = *reinterpret_cast<const float4*>(g + v * 8);
      const float4 g1 = *reinterpret_cast<const float4*>(g + v * 8 + 4);
      __nv_bfloat162 h[4] = {__floats2bfloat162_rn(g0.x, g0.y), __floats2bfloat162_rn(g0.z, g0.w),
                             __floats2bfloat162_rn(g1.x, g1.y), __floats2bfloat162_rn(g1.z, g1.w)};
      *reinterpret_cast<uint4*>(wr + v * 8) = *reinterpret_cast<const uint4*>(h);
      if (a.zero_grad) {
        *reinterpret_cast<float4*>(g + v * 8) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(g + v * 8 + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    if (n8 * 8 + tid < a.n) {
      wr[n8 * 8 + tid] = __float2bfloat16(g[n8 * 8 + tid]);
      if (a.zero_grad) g[n8 * 8 + tid] = 0.f;
    }
    __threadfence_system();
  }
  if (!(phases & 2)) return;
  cross_rank_barrier(cm, ++epoch);
  const float lr = a.lr_state[0];
  const float lr_t = a.lr_state[1];
  const float scale = a.grad_scale / (float)cm.world;
  float* out = cm.param[cm.rank];
  for (int64_t v = tid; v < n8; v += nthr) {
    const int64_t i = a.offset + v * 8;
    uint4 raw[kMaxWorld];
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r)
      raw[r] = r < cm.world ? *reinterpret_cast<const uint4*>(cm.wire[r] + i) : make_uint4(0u, 0u, 0u, 0u);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < kMaxWorld; ++r) {
      if (r < cm.world) {
        float f[8];
        unpack_bf16x8(raw[r], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    float4 w0 = *reinterpret_cast<const float4*>(out + i), w1 = *reinterpret_cast<const float4*>(out + i + 4);
    w0.x = opt_update(a, w0.x, acc[0] * scale, i, lr, lr_t);
    w0.y = opt_update(a, w0.y, acc[1] * scale, i + 1, lr, lr_t);
    w0.z = opt_update(a, w0.z, acc[2] * scale, i + 2, lr, lr_t);
    w0.w = opt_update(a, w0.w, acc[3] * scale, i + 3, lr, lr_t);
    w1.x = opt_update(a, w1.x, acc[4] * scale, i + 4, lr, lr_t);
    w1.y = opt_update(a, w1.y, acc[5] * scale, i + 5, lr, lr_t);
    w1.z = opt_update(a, w1.z, acc[6] * scale, i + 6, lr, lr_t);
    w1.w = opt_update(a, w1.w, acc[7] * scale, i + 7, lr, lr_t);
    *reinterpret_cast<float4*>(out + i) = w0;
    *reinterpret_cast<float4*>(out + i + 4) = w1;
  }
  for (int64_t i = a.offset + n8 * 8 + tid; i < a.offset + a.n; i += nthr) {
    float s = 0.f;
    for (int r = 0; r < cm.world; ++r) s += __bfloat162float(cm.wire[r][i]);
    out[i] = opt_update(a, out[i], s * scale, i, lr, lr_t);
  }
  cross_rank_barrier(cm, ++epoch);  // all peers have finished reading my wire copy
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

// ---- NVLS: the reduction happens inside NVSwitch, the all-gather is one multicast store ----
// multimem.ld_reduce on the multicast address returns the sum over every GPU's copy of the addressed gradient
// (reduced in the switch, fp32), multimem.st writes the updated weights into every GPU's parameter bucket at once.
// Per GPU and direction the links carry about S instead of 2(W-1)/W * S for the two-shot path.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

template <int U>
__global__ void __launch_bounds__(kThreads)
nvls_kernel(const Comm cm, const StepArgs a) {
  // U switch-reduced 16-byte loads in flight per thread: a multimem round trip is about twice a peer load
  uint32_t epoch = cm.epoch[blockIdx.x];
  cross_rank_barrier(cm, ++epoch);
  const float lr = a.lr_state[0];
  const float lr_t = a.lr_state[1];
  const float scale = a.grad_scale / (float)cm.world;
  const int64_t n4 = a.n / 4;  // the host guarantees n % 4 == 0 on this path
  const int64_t per = (n4 + cm.world - 1) / cm.world;
  const int64_t v_lo = per * cm.rank < n4 ? per * cm.rank : n4;
  const int64_t v_hi = v_lo + per < n4 ? v_lo + per : n4;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
  for (int64_t v0 = v_lo + tid; v0 < v_hi; v0 += nthr * U) {
    float4 acc[U];
    int64_t idx[U];
    int nvalid = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      idx[u] = a.offset + (v0 + (int64_t)u * nthr) * 4;
      if (v0 + (int64_t)u * nthr < v_hi) nvalid = u + 1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u < nvalid) acc[u] = multimem_ld_reduce_add(cm.mc_grad + idx[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (u >= nvalid) break;
      const int64_t i = idx[u];
      float4 w = *reinterpret_cast<const float4*>(cm.param[cm.rank] + i);
      w.x = opt_update(a, w.x, acc[u].x * scale, i, lr, lr_t);
      w.y = opt_update(a, w.y, acc[u].y * scale, i + 1, lr, lr_t);
      w.z = opt_update(a, w.z, acc[u].z * scale, i + 2, lr, lr_t);
      w.w = opt_update(a, w.w, acc[u].w * scale, i + 3, lr, lr_t);
      multimem_st(cm.mc_param + i, w);
    }
  }
  __threadfence_system();
  cross_rank_barrier(cm, ++epoch);
  if (a.zero_grad) {
    for (int sl = 0; sl < cm.world; ++sl) {
      const int64_t s_lo = per * sl < n4 ? per * sl : n4;
      const int64_t s_hi = s_lo + per < n4 ? s_lo + per : n4;
      zero_vectors(cm, a, a.offset + s_lo * 4, s_hi - s_lo);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

// ---- broadcast (X2: tower `src`'s parameters become everyone's) ----
// Pull model: every rank copies [offset, offset+n) of `src`'s parameter buffer into its own over NVLink; the
// barriers keep `src` from touching the range while peers are still reading it.
__global__ void __launch_bounds__(kThreads)
broadcast_kernel(const Comm cm, int src, int64_t offset, int64_t n) {
  uint32_t epoch = cm.epoch[blockIdx.x];
  cross_rank_barrier(cm, ++epoch);
  if (cm.rank != src) {
    const float* from = cm.param[src] + offset;
    float* to = cm.param[cm.rank] + offset;
    const int64_t n4 = n / 4;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    for (int64_t v = tid; v < n4; v += nthr)
      reinterpret_cast<float4*>(to)[v] = reinterpret_cast<const float4*>(from)[v];
    for (int64_t i = n4 * 4 + tid; i < n; i += nthr) to[i] = from[i];
  }
  __threadfence_system();
  cross_rank_barrier(cm, ++epoch);
  __syncthreads();
  if (threadIdx.x == 0) cm.epoch[blockIdx.x] = epoch;
}

}  // namespace p2p
}  // namespace tdl

using namespace tdl;
using namespace tdl::p2p;

#include <map>
static std::map<Comm*, int*>& g_status_host() {  // host addresses of the communicators' status words
  static std::map<Comm*, int*> m;
  return m;
}

// ---- symmetric heap: plain cudaMalloc allocations that can be exported through cudaIpc ----
TDL_API int tdl_symm_alloc(int64_t bytes, void** out) {
  TDL_REQUIRE(bytes > 0, "bytes must be positive");
  void* p = nullptr;
  TDL_CUDA_OK(cudaMalloc(&p, (size_t)bytes));
  TDL_CUDA_OK(cudaMemset(p, 0, (size_t)bytes));
  TDL_CUDA_OK(cudaDeviceSynchronize());
  *out = p;
  return 0;
}
TDL_API int tdl_symm_free(void* p) {
  TDL_CUDA_OK(cudaFree(p));
  return 0;
}
TDL_API int tdl_ipc_export(void* p, void* handle64) {
  cudaIpcMemHandle_t h;
  TDL_CUDA_OK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
TDL_API int tdl_ipc_open(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  TDL_CUDA_OK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
TDL_API int tdl_ipc_close(void* p) {
  TDL_CUDA_OK(cudaIpcCloseMemHandle(p));
  return 0;
}
TDL_API int tdl_p2p_flag_bytes() { return kMaxBlocks * kMaxWorld * (int)sizeof(uint32_t); }

// Build a communicator from already-mapped pointers (index = rank).  `epoch_local` is kMaxBlocks u32, zeroed.
TDL_API int tdl_p2p_comm_create(int world, int rank, void** grad_ptrs, void** param_ptrs, void** flag_ptrs,
                                void* epoch_local, int emulate, void** out_comm) {
  TDL_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "bad world/rank");
  Comm* c = new Comm();
  memset(c, 0, sizeof(Comm));
  c->world = world;
  c->rank = rank;
  c->emulate = emulate;
  for (int r = 0; r < world; ++r) {
    c->grad[r] = (float*)grad_ptrs[r];
    c->param[r] = param_ptrs ? (float*)param_ptrs[r] : nullptr;
    c->flags[r] = flag_ptrs ? (uint32_t*)flag_ptrs[r] : nullptr;
  }
  c->epoch = (uint32_t*)epoch_local;
  c->timeout_ns = 600ull * 1000000000ull;
  // status word in pinned, device-mapped host memory: it survives the sticky error a trap leaves behind
  int* host_status = nullptr;
  if (cudaHostAlloc((void**)&host_status, 8 * sizeof(int), cudaHostAllocMapped) == cudaSuccess) {
    memset(host_status, 0, 8 * sizeof(int));
    int* dev_status = nullptr;
    if (cudaHostGetDevicePointer((void**)&dev_status, host_status, 0) == cudaSuccess) c->status = dev_status;
    g_status_host()[c] = host_status;
  } else {
    (void)cudaGetLastError();
  }
  *out_comm = c;
  return 0;
}
// How long a CTA waits for a peer's signal before it gives up (seconds; <= 0 keeps the current limit).
TDL_API int tdl_p2p_comm_set_timeout(void* comm, double seconds) {
  Comm* c = (Comm*)comm;
  TDL_REQUIRE(c != nullptr, "null communicator");
  if (seconds > 0) c->timeout_ns = (unsigned long long)(seconds * 1e9);
  return 0;
}
// Turns the signal-pad protocol assertions on: `seen` is a zeroed device array of kMaxBlocks * kMaxWorld u32 owned by
// the caller (NULL turns them off).  Status codes 2 / 3 (tdl_p2p_comm_status) = "signal went backwards" / "too far ahead".
TDL_API int tdl_p2p_comm_set_debug(void* comm, void* seen) {
  Comm* c = (Comm*)comm;
  TDL_REQUIRE(c != nullptr, "null communicator");
  c->debug_seen = (uint32_t*)seen;
  return 0;
}
// Reads the status word: out5 = {timed_out, rank, cta, peer, epoch}.  Plain host memory: safe to call after a CUDA
// error has poisoned the context.
TDL_API int tdl_p2p_comm_status(void* comm, int* out5) {
  auto it = g_status_host().find((Comm*)comm);
  for (int i = 0; i < 5; ++i) out5[i] = it == g_status_host().end() ? 0 : ((volatile int*)it->second)[i];
  return 0;
}
// Attach NVLS multicast addresses (from a CUDA multicast object bound to the same buffers) to a communicator.
TDL_API int tdl_p2p_comm_set_multicast(void* comm, void* mc_grad, void* mc_param) {
  Comm* c = (Comm*)comm;
  TDL_REQUIRE(c != nullptr, "null communicator");
  c->mc_grad = (float*)mc_grad;
  c->mc_param = (float*)mc_param;
  return 0;
}
// Attach the peer-mapped bf16 wire buffers (index = rank; same element offsets as the gradient buckets).
TDL_API int tdl_p2p_comm_set_wire(void* comm, void** wire_ptrs) {
  Comm* c = (Comm*)comm;
  TDL_REQUIRE(c != nullptr && wire_ptrs != nullptr, "null communicator / table");
  for (int r = 0; r < c->world; ++r) c->wire[r] = (bf16*)wire_ptrs[r];
  return 0;
}
TDL_API int tdl_p2p_comm_destroy(void* comm) {
  auto it = g_status_host().find((Comm*)comm);
  if (it != g_status_host().end()) {
    (void)cudaFreeHost(it->second);
    g_status_host().erase(it);
  }
  delete (Comm*)comm;
  return 0;
}

// Fused all-reduce(mean) + optimizer for one bucket.  mode: 1 = one-shot, 2 = two-shot, 3 = NVLS multimem
// (4 loads in flight per thread), 4 = NVLS multimem (8 in flight), 0 = pick one-/two-shot by size.
// kind: 0 = plain mean written back into the gradient bucket (m1 = fp32 staging of >= offset+n elements),
//       1 = momentum SGD (a = momentum, b = weight decay), 2 = TF Adam (a = beta1, b = beta2, c = eps).
TDL_API int tdl_p2p_fused_step(void* comm, int64_t offset, int64_t n, int mode, int kind, int grad_is_bf16,
                               int zero_grad, float* m1, float* m2, const float* lr_state, float a, float b, float c,
                               float grad_scale, int max_blocks, void* stream) {
  Comm* cm = (Comm*)comm;
  TDL_REQUIRE(cm != nullptr, "null communicator");
  TDL_REQUIRE(offset % 4 == 0, "bucket offsets must be multiples of 4 elements");
  TDL_REQUIRE(!(kind == kNone && grad_is_bf16), "plain mean supports fp32 buckets only");
  if (n <= 0) return 0;
  StepArgs s;
  s.offset = offset; s.n = n; s.kind = kind; s.grad_is_bf16 = grad_is_bf16; s.zero_grad = zero_grad;
  s.m1 = m1; s.m2 = m2; s.lr_state = lr_state; s.a = a; s.b = b; s.c = c; s.grad_scale = grad_scale;
  if (mode == 0) mode = (n * (grad_is_bf16 ? 2 : 4) <= (1 << 19) || kind == kNone) ? 1 : 2;
  if (kind == kNone) mode = 1;
  if (mode == 3 || mode == 4) {
    TDL_REQUIRE(cm->mc_grad && cm->mc_param, "NVLS mode needs multicast addresses (tdl_p2p_comm_set_multicast)");
    TDL_REQUIRE(!grad_is_bf16 && n % 4 == 0 && kind != kNone, "NVLS mode: fp32 buckets, n % 4 == 0, with optimizer");
  }
  int64_t work = mode == 1 ? n / 4 : n / 4 / cm->world;
  if (mode == 3 || mode == 4) work /= 2;  // one switch-reduced load replaces W peer loads: fewer CTAs saturate the links
  int blocks = (int)cdiv(work, (int64_t)kThreads * 2);
  if (max_blocks <= 0) max_blocks = 128;
  if (max_blocks > kMaxBlocks) max_blocks = kMaxBlocks;
  if (max_blocks > num_sms()) max_blocks = num_sms();
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 1) one_shot_kernel<<<blocks, kThreads, 0, st>>>(*cm, s);
  else if (mode == 3) nvls_kernel<4><<<blocks, kThreads, 0, st>>>(*cm, s);
  else if (mode == 4) nvls_kernel<8><<<blocks, kThreads, 0, st>>>(*cm, s);
  else two_shot_kernel<<<blocks, kThreads, 0, st>>>(*cm, s);
  TDL_LAUNCH_OK("p2p fused step");
  return 0;
}

// Same fused step with the gradients crossing NVLink as bf16 (fp32 buckets are packed into the wire buffers and cleared
// by the kernel's first phase).  mode: 0 = by size, 1 = one-shot, 2 = two-shot.  kind: 1 = SGD-momentum, 2 = Adam.
// phases: 3 = pack + reduce (normal use); 1 / 2 run one phase only (single-GPU emulation launches every rank's phase 1
// before any phase 2, because its "ranks" cannot wait for each other).
TDL_API int tdl_p2p_fused_step_wire(void* comm, int64_t offset, int64_t n, int mode, int kind, int zero_grad,
                                    float* m1, float* m2, const float* lr_state, float a, float b, float c,
                                    float grad_scale, int max_blocks, int phases, void* stream) {
  Comm* cm = (Comm*)comm;
  TDL_REQUIRE(cm != nullptr, "null communicator");
  TDL_REQUIRE(cm->wire[cm->rank] != nullptr, "no wire buffers attached (tdl_p2p_comm_set_wire)");
  TDL_REQUIRE(offset % 8 == 0, "bucket offsets must be multiples of 8 elements");
  TDL_REQUIRE(kind == kSgdMomentum || kind == kAdam, "wire format: SGD-momentum or Adam");
  TDL_REQUIRE(phases >= 1 && phases <= 3, "phases is 1, 2 or 3");
  if (n <= 0) return 0;
  StepArgs s;
  s.offset = offset; s.n = n; s.kind = kind; s.grad_is_bf16 = 0; s.zero_grad = zero_grad;
  s.m1 = m1; s.m2 = m2; s.lr_state = lr_state; s.a = a; s.b = b; s.c = c; s.grad_scale = grad_scale;
  if (mode == 0) mode = (n * 4 <= (1 << 19)) ? 1 : 2;
  // the grid must not depend on `phases`: both phases use the same (CTA, thread) -> vector assignment
  int64_t work = mode == 1 ? n / 8 : n / 8 / cm->world;
  int blocks = (int)cdiv(work, (int64_t)kThreads * 2);
  if (max_blocks <= 0) max_blocks = 128;
  if (max_blocks > kMaxBlocks) max_blocks = kMaxBlocks;
  if (max_blocks > num_sms()) max_blocks = num_sms();
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 1) one_shot_wire_kernel<<<blocks, kThreads, 0, st>>>(*cm, s, phases);
  else two_shot_wire_kernel<2><<<blocks, kThreads, 0, st>>>(*cm, s, phases);
  TDL_LAUNCH_OK("p2p fused step (bf16 wire)");
  return 0;
}

// Mirror rank `src`'s parameters [offset, offset+n) into every rank's parameter buffer (all ranks call this).
TDL_API int tdl_p2p_broadcast(void* comm, int src, int64_t offset, int64_t n, int max_blocks, void* stream) {
  Comm* cm = (Comm*)comm;
  TDL_REQUIRE(cm != nullptr, "null communicator");
  TDL_REQUIRE(src >= 0 && src < cm->world, "source rank out of range");
  TDL_REQUIRE(offset % 4 == 0, "offset must be a multiple of 4 elements");
  TDL_REQUIRE(!cm->emulate, "broadcast is not available for the single-GPU emulation communicator");
  if (n <= 0) return 0;
  int blocks = (int)cdiv(n / 4 + 1, (int64_t)kThreads * 4);
  if (max_blocks <= 0) max_blocks = 64;
  if (max_blocks > kMaxBlocks) max_blocks = kMaxBlocks;
  if (max_blocks > num_sms()) max_blocks = num_sms();
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks < 1) blocks = 1;
  broadcast_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(*cm, src, offset, n);
  TDL_LAUNCH_OK("p2p broadcast");
  return 0;
}
