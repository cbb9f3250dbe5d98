// Memory-bound spatial ops for NHWC activations: max-pool, global average pool, depthwise
// convolution, and the reference's `_upsample` (symmetric pad 1 + TF1 legacy bilinear + crop 2).
//
// reference call sites: core/resnet.py:241 (max_pool2d 3x3 s2 SAME), :246-248/:464-465 (mean pool),
// core/layers.py:34-42 + core/xception.py:102-110 (depthwise), core/layers.py:82-109 (_upsample).
// All kernels move 8 bf16 channels (16 bytes) per thread per access and accumulate in fp32.
#include "tdl_common.cuh"

namespace tdl {
namespace spatial {

constexpr int kThreads = 256;

static int grid_for(int64_t work) {
  int64_t b = cdiv(work, kThreads);
  const int64_t cap = (int64_t)num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// ------------------------------------------------------------------------------------------
// max pool (padded cells never win); forward records the winning tap so backward is a gather
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, int N, int H, int W,
                   int C, int P, int Q, int k, int stride, int pad_h, int pad_w) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c8 = C / 8;
  const int64_t total = (int64_t)N * P * Q * c8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % c8);
    int64_t t = i / c8;
    const int q = (int)(t % Q); t /= Q;
    const int p = (int)(t % P);
    const int n = (int)(t / P);
    float best[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
    for (int r = 0; r < k; ++r) {
      const int h = p * stride + r - pad_h;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < k; ++s) {
        const int w = q * stride + s - pad_w;
        if (w < 0 || w >= W) continue;
        float f[8];
        unpack8(ld8(x + (((int64_t)n * H + h) * W + w) * C + cv * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (f[j] > best[j]) { best[j] = f[j]; arg[j] = r * k + s; }
      }
    }
    st8(y + i * 8, pack8(best));
    uint2 packed;
    packed.x = (uint32_t)arg[0] | ((uint32_t)arg[1] << 8) | ((uint32_t)arg[2] << 16) | ((uint32_t)arg[3] << 24);
    packed.y = (uint32_t)arg[4] | ((uint32_t)arg[5] << 8) | ((uint32_t)arg[6] << 16) | ((uint32_t)arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + i * 8) = packed;
  }
}

__global__ void __launch_bounds__(kThreads)
maxpool_bwd_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx, bf16* __restrict__ dx, int N, int H,
                   int W, int C, int P, int Q, int k, int stride, int pad_h, int pad_w) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c8 = C / 8;
  const int64_t total = (int64_t)N * H * W * c8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % c8);
    int64_t t = i / c8;
    const int w = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int r = 0; r < k; ++r) {
      const int hp = h + pad_h - r;
      if (hp < 0 || hp % stride) continue;
      const int p = hp / stride;
      if (p >= P) continue;
      for (int s = 0; s < k; ++s) {
        const int wq = w + pad_w - s;
        if (wq < 0 || wq % stride) continue;
        const int q = wq / stride;
        if (q >= Q) continue;
        const int64_t o = ((((int64_t)n * P + p) * Q + q) * c8 + cv) * 8;
        const uint2 packed = *reinterpret_cast<const uint2*>(idx + o);
        float g[8];
        unpack8(ld8(dy + o), g);
        const int tap = r * k + s;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t a = j < 4 ? (packed.x >> (8 * j)) & 0xff : (packed.y >> (8 * (j - 4))) & 0xff;
          if ((int)a == tap) acc[j] += g[j];
        }
      }
    }
    st8(dx + i * 8, pack8(acc));
  }
}

// Compile-time window/stride variants, one thread per (pixel, 8 channels) with blockIdx.y = image: no 64-bit index
// arithmetic or runtime-stride divisions (the generic kernels above spend most of their issue slots on those), all window loads in flight
// at once.  Used for the 3x3 stride-2 pool of the ResNet root (core/resnet.py:241).
// kStats: the kernel also leaves per-CTA partial (sum, sum of squares) of the pooled values per channel in
// stats[cta][2][C] -- the batch statistics of the `postnorm` BatchNorm that follows the root pool in the reference
// network (core/resnet.py:241-242), so that BatchNorm needs no statistics pass of its own.  Requires
// kThreads % (C/8) == 0 (every thread of a CTA then keeps the same channel group for the whole CTA).
template <int K, int S, bool kStats = false>
__global__ void __launch_bounds__(kThreads)
maxpool_fwd_rows_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, int H, int W,
                        int C, int P, int Q, int pad_h, int pad_w, float* __restrict__ stats = nullptr) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c8 = C >> 3;
  const int n = blockIdx.y;
  const bf16* xn = x + (int64_t)n * H * W * C;
  const int64_t orow = (int64_t)n * P * Q * C;
  const int i = blockIdx.x * kThreads + threadIdx.x;  // (p, q, cv) inside image n, 32-bit arithmetic only
  float st_s[8], st_q[8];
  if constexpr (kStats) {
#pragma unroll
    for (int j = 0; j < 8; ++j) st_s[j] = st_q[j] = 0.f;
  }
  if (i < P * Q * c8) {
    const int pq = i / c8;
    const int cv = i - pq * c8;
    const int p = pq / Q;
    const int q = pq - p * Q;
    const int h0 = p * S - pad_h;
    const int w0 = q * S - pad_w;
    bf16x8 v[K * K];
    bool ok[K * K];
#pragma unroll
    for (int r = 0; r < K; ++r)
#pragma unroll
      for (int t = 0; t < K; ++t) {
        const int h = h0 + r, w = w0 + t;
        ok[r * K + t] = h >= 0 && h < H && w >= 0 && w < W;
        if (ok[r * K + t]) v[r * K + t] = ld8(xn + ((int64_t)h * W + w) * C + cv * 8);
      }
    float best[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
#pragma unroll
    for (int tap = 0; tap < K * K; ++tap) {
      if (!ok[tap]) continue;
      float f[8];
      unpack8(v[tap], f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f[j] > best[j]) { best[j] = f[j]; arg[j] = tap; }
    }
    st8(y + orow + (int64_t)i * 8, pack8(best));
    uint2 packed;
    packed.x = (uint32_t)arg[0] | ((uint32_t)arg[1] << 8) | ((uint32_t)arg[2] << 16) | ((uint32_t)arg[3] << 24);
    packed.y = (uint32_t)arg[4] | ((uint32_t)arg[5] << 8) | ((uint32_t)arg[6] << 16) | ((uint32_t)arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + orow + (int64_t)i * 8) = packed;
    if constexpr (kStats) {  // the maximum of bf16 values is a bf16 value: these ARE the stored numbers
#pragma unroll
      for (int j = 0; j < 8; ++j) { st_s[j] = best[j]; st_q[j] = best[j] * best[j]; }
    }
  }
  if constexpr (kStats) {
    // thread t owns channel group t % c8 (kThreads % c8 == 0): fold the kThreads / c8 pixel slots per group
    __shared__ float sm[kThreads * 16];
    float* mine = sm + threadIdx.x * 16;
#pragma unroll
    for (int j = 0; j < 8; ++j) { mine[j] = st_s[j]; mine[8 + j] = st_q[j]; }
    __syncthreads();
    const int slots = kThreads / c8;
    float* row = stats + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2 * C;
    for (int e = threadIdx.x; e < c8 * 16; e += kThreads) {
      const int cv = e >> 4, val = e & 15;
      float acc = 0.f;
      for (int t = 0; t < slots; ++t) acc += sm[(t * c8 + cv) * 16 + val];
      row[(size_t)(val >> 3) * C + cv * 8 + (val & 7)] = acc;
    }
  }
}

// Blocked form of the backward gather: a thread owns an S x S block of input pixels (aligned to the window grid)
// for 8 channels.  All pixels of the block can only be reached from the same kCells x kCells output cells, so the
// gradient + arg-max index of those cells are loaded once per block instead of once per pixel (K=3, S=2: 4 cells
// per 4 pixels instead of 16) -- the per-pixel form moved 6 bytes through L2 per byte it stored.
template <int K, int S>
__global__ void __launch_bounds__(kThreads)
maxpool_bwd_block_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx, bf16* __restrict__ dx, int H,
                         int W, int C, int P, int Q, int pad_h, int pad_w, int A, int B) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  constexpr int kCells = (K + S - 1) / S;
  const int c8 = C >> 3;
  const int n = blockIdx.y;
  const int64_t obase = (int64_t)n * P * Q * C;
  const int64_t ibase = (int64_t)n * H * W * C;
  const int i = blockIdx.x * kThreads + threadIdx.x;  // (a, b, cv): block row, block column, channel vector
  if (i >= A * B * c8) return;
  const int ab = i / c8;
  const int cv = i - ab * c8;
  const int a = ab / B;
  const int b = ab - a * B;
  bf16x8 g[kCells * kCells];
  uint2 am[kCells * kCells];
  bool okc[kCells * kCells];
#pragma unroll
  for (int j = 0; j < kCells; ++j)
#pragma unroll
    for (int k = 0; k < kCells; ++k) {
      const int p = a - j, q = b - k;
      const int c = j * kCells + k;
      okc[c] = p >= 0 && p < P && q >= 0 && q < Q;
      if (okc[c]) {
        const int64_t o = obase + (((int64_t)p * Q + q) * c8 + cv) * 8;
        g[c] = ld8(dy + o);
        am[c] = *reinterpret_cast<const uint2*>(idx + o);
      }
    }
#pragma unroll
  for (int dr = 0; dr < S; ++dr)
#pragma unroll
    for (int dc = 0; dc < S; ++dc) {
      const int h = a * S + dr - pad_h, w = b * S + dc - pad_w;
      if (h < 0 || h >= H || w < 0 || w >= W) continue;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
      for (int j = 0; j < kCells; ++j)
#pragma unroll
        for (int k = 0; k < kCells; ++k) {
          const int r = dr + S * j, t = dc + S * k;  // tap of this pixel inside the window of cell (a - j, b - k)
          const int c = j * kCells + k;
          if (r >= K || t >= K || !okc[c]) continue;
          float f[8];
          unpack8(g[c], f);
          const uint32_t tap = (uint32_t)(r * K + t);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const uint32_t win = e < 4 ? (am[c].x >> (8 * e)) & 0xff : (am[c].y >> (8 * (e - 4))) & 0xff;
            if (win == tap) acc[e] += f[e];
          }
        }
      st8(dx + ibase + (((int64_t)h * W + w) * c8 + cv) * 8, pack8(acc));
    }
}

// ------------------------------------------------------------------------------------------
// global average pool
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
gap_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int HW, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ float sm[8][32][8];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c8 = C / 8;
  const int cv = blockIdx.y * 32 + tx;
  const int n = blockIdx.x;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (cv < c8)
    for (int r = ty; r < HW; r += 8) {
      float f[8];
      unpack8(ld8(x + ((int64_t)n * HW + r) * C + cv * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += f[j];
    }
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[ty][tx][j] = s[j];
  __syncthreads();
  if (ty == 0 && cv < c8) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = 0.f;
      for (int t = 0; t < 8; ++t) a += sm[t][tx][j];
      o[j] = a / (float)HW;
    }
    st8(y + (int64_t)n * C + cv * 8, pack8(o));
  }
}

__global__ void __launch_bounds__(kThreads)
gap_bwd_kernel(const bf16* __restrict__ dy, bf16* __restrict__ dx, int64_t total_vec, int HW, int c8) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const float inv = 1.f / (float)HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % c8);
    const int64_t n = i / ((int64_t)c8 * HW);
    float g[8];
    unpack8(ld8(dy + (n * c8 + cv) * 8), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] *= inv;
    st8(dx + i * 8, pack8(g));
  }
}

// ------------------------------------------------------------------------------------------
// global average pool folded into the 1x1 convolution that follows it (the "image pooling" branch of the
// reference's ASPP: reduce_mean -> 1x1 conv, core/resnet.py:464-466).  out[n, k] = sum_c mean_hw(x[n, :, :, c]) * w[k, c].
// The pooled vector never goes to memory as an activation (only as the small fp32 [N, C] tensor the weight gradient
// needs), and the M = N "GEMM" with a handful of rows does not occupy a 128-row tensor-core tile that is ~75 % padding.
// ------------------------------------------------------------------------------------------
constexpr int kGapKT = 64;  // output channels per CTA

__global__ void __launch_bounds__(kThreads)
gap_conv1x1_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, bf16* __restrict__ y,
                       float* __restrict__ gap_out, int HW, int C, int K) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float gap[];  // [C]
  const int n = blockIdx.x;
  const int c8 = C >> 3;
  const float inv = 1.f / (float)HW;
  // phase 1: mean over the HW pixels of image n, 8 channels per thread-visit
  for (int cv = threadIdx.x; cv < c8; cv += kThreads) {
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = 0.f;
    for (int r = 0; r < HW; ++r) {
      float f[8];
      unpack8(ld8(x + ((int64_t)n * HW + r) * C + cv * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += f[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      gap[cv * 8 + j] = s[j] * inv;
      if (blockIdx.y == 0) gap_out[(int64_t)n * C + cv * 8 + j] = s[j] * inv;
    }
  }
  __syncthreads();
  // phase 2: one warp per output channel, kGapKT channels per CTA
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int kk = warp; kk < kGapKT; kk += kThreads / 32) {
    const int k = blockIdx.y * kGapKT + kk;
    if (k >= K) break;
    const float* wr = w + (int64_t)k * C;
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(gap[c], wr[c], acc);
    acc = warp_sum(acc);
    if (lane == 0) y[(int64_t)n * K + k] = __float2bfloat16(acc);
  }
}

// dx[n, r, c] = (sum_k dy[n, k] * w[k, c]) / HW
__global__ void __launch_bounds__(kThreads)
gap_conv1x1_bwd_data_kernel(const bf16* __restrict__ dy, const float* __restrict__ w, bf16* __restrict__ dx, int HW,
                            int C, int K) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ __align__(16) bf16 dg[kThreads];
  const int n = blockIdx.x;
  const int c = blockIdx.y * kThreads + threadIdx.x;
  float acc = 0.f;
  if (c < C)
    for (int k = 0; k < K; ++k) acc = fmaf(__bfloat162float(dy[(int64_t)n * K + k]), w[(int64_t)k * C + c], acc);
  dg[threadIdx.x] = __float2bfloat16(acc / (float)HW);
  __syncthreads();
  const int c0 = blockIdx.y * kThreads;
  const int nvec = min(kThreads, C - c0) >> 3;  // 16-byte vectors of this CTA's channel chunk
  for (int i = threadIdx.x; i < HW * nvec; i += kThreads) {
    const int r = i / nvec, v = i - r * nvec;
    st8(dx + ((int64_t)n * HW + r) * C + c0 + v * 8, ld8(dg + v * 8));
  }
}

// dw[k, c] += sum_n dy[n, k] * gap[n, c]
__global__ void __launch_bounds__(kThreads)
gap_conv1x1_bwd_weight_kernel(const bf16* __restrict__ dy, const float* __restrict__ gap, float* __restrict__ dw, int N,
                              int C, int K) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int k = blockIdx.x;
  const int c = blockIdx.y * kThreads + threadIdx.x;
  if (c >= C) return;
  float acc = 0.f;
  for (int n = 0; n < N; ++n) acc = fmaf(__bfloat162float(dy[(int64_t)n * K + k]), gap[(int64_t)n * C + c], acc);
  dw[(int64_t)k * C + c] += acc;
}

// ------------------------------------------------------------------------------------------
// depthwise convolution, weights [R*S, C] fp32
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
depthwise_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     bf16* __restrict__ y, int N, int H, int W, int C, int P, int Q, int R, int S, int stride,
                     int dil, int pad_h, int pad_w, int relu) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c8 = C / 8;
  const int64_t total = (int64_t)N * P * Q * c8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % c8);
    int64_t t = i / c8;
    const int q = (int)(t % Q); t /= Q;
    const int p = (int)(t % P);
    const int n = (int)(t / P);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = bias ? bias[cv * 8 + j] : 0.f;
    for (int r = 0; r < R; ++r) {
      const int h = p * stride + r * dil - pad_h;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < S; ++s) {
        const int ww = q * stride + s * dil - pad_w;
        if (ww < 0 || ww >= W) continue;
        float f[8];
        unpack8(ld8(x + (((int64_t)n * H + h) * W + ww) * C + cv * 8), f);
        const float* wp = w + (int64_t)(r * S + s) * C + cv * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(f[j], wp[j], acc[j]);
      }
    }
    if (relu) {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaxf(acc[j], 0.f);
    }
    st8(y + i * 8, pack8(acc));
  }
}

// dx = sum over taps of (dy masked by relu) * w
__global__ void __launch_bounds__(kThreads)
depthwise_bwd_data_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ y, const float* __restrict__ w,
                          bf16* __restrict__ dx, int N, int H, int W, int C, int P, int Q, int R, int S, int stride,
                          int dil, int pad_h, int pad_w, int relu) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c8 = C / 8;
  const int64_t total = (int64_t)N * H * W * c8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % c8);
    int64_t t = i / c8;
    const int ww = (int)(t % W); t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int r = 0; r < R; ++r) {
      const int hp = h + pad_h - r * dil;
      if (hp < 0 || hp % stride) continue;
      const int p = hp / stride;
      if (p >= P) continue;
      for (int s = 0; s < S; ++s) {
        const int wq = ww + pad_w - s * dil;
        if (wq < 0 || wq % stride) continue;
        const int q = wq / stride;
        if (q >= Q) continue;
        const int64_t o = ((((int64_t)n * P + p) * Q + q) * c8 + cv) * 8;
        float g[8];
        unpack8(ld8(dy + o), g);
        if (relu) {
          float yy[8];
          unpack8(ld8(y + o), yy);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = yy[j] > 0.f ? g[j] : 0.f;
        }
        const float* wp = w + (int64_t)(r * S + s) * C + cv * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(g[j], wp[j], acc[j]);
      }
    }
    st8(dx + i * 8, pack8(acc));
  }
}

// dw[tap, c] += sum_pixels dyv * x ; dbias[c] += sum_pixels dyv.  3x3 only (taps <= 9).
__global__ void __launch_bounds__(kThreads)
depthwise_bwd_weight_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, const bf16* __restrict__ y,
                            float* __restrict__ dw, float* __restrict__ dbias, int N, int H, int W, int C, int P,
                            int Q, int R, int S, int stride, int dil, int pad_h, int pad_w, int relu, int tx, int ty) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float sm[];  // [ty][tx][8]
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int c8 = C / 8;
  const int cv = blockIdx.y * tx + cx;
  const int taps = R * S;
  float acc[9][8];
  float accb[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) accb[j] = 0.f;
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
  const int64_t pixels = (int64_t)N * P * Q;
  if (cv < c8) {
    for (int64_t px = (int64_t)blockIdx.x * ty + ry; px < pixels; px += (int64_t)gridDim.x * ty) {
      const int q = (int)(px % Q);
      const int p = (int)((px / Q) % P);
      const int n = (int)(px / ((int64_t)Q * P));
      float g[8];
      unpack8(ld8(dy + px * C + cv * 8), g);
      if (relu) {
        float yy[8];
        unpack8(ld8(y + px * C + cv * 8), yy);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = yy[j] > 0.f ? g[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) accb[j] += g[j];
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        if (t < taps) {
          const int r = t / S, s = t - (t / S) * S;
          const int h = p * stride + r * dil - pad_h;
          const int ww = q * stride + s * dil - pad_w;
          if (h >= 0 && h < H && ww >= 0 && ww < W) {
            float f[8];
            unpack8(ld8(x + (((int64_t)n * H + h) * W + ww) * C + cv * 8), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(g[j], f[j], acc[t][j]);
          }
        }
      }
    }
  }
  // reduce over ry through shared memory, one statistic (tap or bias) at a time
  for (int t = 0; t <= taps; ++t) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v = accb[j];
#pragma unroll
      for (int u = 0; u < 9; ++u)
        if (u == t) v = acc[u][j];
      sm[((size_t)ry * tx + cx) * 8 + j] = v;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < tx * 8; idx += kThreads) {
      const int col = idx >> 3, j = idx & 7;
      const int cvv = blockIdx.y * tx + col;
      if (cvv >= c8) continue;
      float a = 0.f;
      for (int u = 0; u < ty; ++u) a += sm[((size_t)u * tx + col) * 8 + j];
      if (t < taps) atomicAdd(dw + (int64_t)t * C + cvv * 8 + j, a);
      else if (dbias) atomicAdd(dbias + cvv * 8 + j, a);
    }
  }
}

// ------------------------------------------------------------------------------------------
// `_upsample`: symmetric pad 1 -> legacy bilinear to (OH+4, OW+4) -> crop 2
// ------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16(v); }

struct UpCoord {
  int lo, hi;
  float w;
};
// o: output index in the cropped frame; in: unpadded input size; out: requested output size
__device__ __forceinline__ UpCoord up_coord(int o, int in, int out) {
  const int inp = in + 2;
  const float scale = (float)inp / (float)(out + 4);
  const float src = (float)(o + 2) * scale;
  int lo = (int)floorf(src);
  if (lo > inp - 1) lo = inp - 1;
  int hi = lo + 1 < inp - 1 ? lo + 1 : inp - 1;
  UpCoord c;
  c.w = src - (float)lo;
  // padded index -> source index (symmetric pad by one == edge replicate)
  c.lo = min(max(lo - 1, 0), in - 1);
  c.hi = min(max(hi - 1, 0), in - 1);
  return c;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
upsample_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H, int W, int C, int OH, int OW,
                    int64_t ldy) {  // ldy: elements between consecutive output pixels (C, or a wider concat buffer's)
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)N * OH * OW * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t t = i / C;
    const int ox = (int)(t % OW); t /= OW;
    const int oy = (int)(t % OH);
    const int n = (int)(t / OH);
    const UpCoord cy = up_coord(oy, H, OH), cx = up_coord(ox, W, OW);
    const T* base = x + (int64_t)n * H * W * C + c;
    const float tl = to_f<T>(base[((int64_t)cy.lo * W + cx.lo) * C]);
    const float tr = to_f<T>(base[((int64_t)cy.lo * W + cx.hi) * C]);
    const float bl = to_f<T>(base[((int64_t)cy.hi * W + cx.lo) * C]);
    const float br = to_f<T>(base[((int64_t)cy.hi * W + cx.hi) * C]);
    const float top = tl + (tr - tl) * cx.w;
    const float bot = bl + (br - bl) * cx.w;
    y[(i / C) * ldy + c] = from_f<T>(top + (bot - top) * cy.w);
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
upsample_bwd_kernel(const T* __restrict__ dy, float* __restrict__ dx, int N, int H, int W, int C, int OH, int OW,
                    int64_t ld_dy) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)N * OH * OW * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t t = i / C;
    const int ox = (int)(t % OW); t /= OW;
    const int oy = (int)(t % OH);
    const int n = (int)(t / OH);
    const UpCoord cy = up_coord(oy, H, OH), cx = up_coord(ox, W, OW);
    const float g = to_f<T>(dy[(i / C) * ld_dy + c]);
    float* base = dx + (int64_t)n * H * W * C + c;
    atomicAdd(base + ((int64_t)cy.lo * W + cx.lo) * C, g * (1.f - cy.w) * (1.f - cx.w));
    atomicAdd(base + ((int64_t)cy.lo * W + cx.hi) * C, g * (1.f - cy.w) * cx.w);
    atomicAdd(base + ((int64_t)cy.hi * W + cx.lo) * C, g * cy.w * (1.f - cx.w));
    atomicAdd(base + ((int64_t)cy.hi * W + cx.hi) * C, g * cy.w * cx.w);
  }
}

}  // namespace spatial
}  // namespace tdl

using namespace tdl;
using namespace tdl::spatial;

TDL_API int tdl_maxpool_fwd(const void* x, void* y, void* idx_u8, int N, int H, int W, int C, int P, int Q, int k,
                            int stride, int pad_h, int pad_w, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  TDL_REQUIRE(k * k <= 255, "window too large");
  const int64_t work = (int64_t)N * P * Q * (C / 8);
  if (work == 0) return 0;
  if (k == 3 && stride == 2 && N <= 65535 && (int64_t)P * Q * C < (1ll << 31))
    {
      cudaError_t _le = tdl::launch_pdl(maxpool_fwd_rows_kernel<3, 2, false>,
          dim3(dim3((unsigned)cdiv((int64_t)P * Q * (C / 8), kThreads), N)), dim3(kThreads), 0, (cudaStream_t)stream,
          (const bf16*)x, (bf16*)y, (uint8_t*)idx_u8, H, W, C, P, Q, pad_h, pad_w, (float*)nullptr);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "maxpool_fwd_rows_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(maxpool_fwd_kernel, dim3(grid_for(work)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const bf16*)x, (bf16*)y, (uint8_t*)idx_u8, N, H, W, C, P, Q, k, stride, pad_h,
          pad_w);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "maxpool_fwd_kernel");
    }
  TDL_LAUNCH_OK("maxpool_fwd_kernel");
  return 0;
}

// 3x3 stride-2 max-pool that also leaves the partial BatchNorm statistics of its output: stats_partial is
// [*nparts][2][C] fp32 with *nparts = N * ceil(P*Q*(C/8) / 256) rows (tdl_maxpool_stats_rows); feed both to
// tdl_bn_finalize_partials.  Needs 256 % (C/8) == 0.
TDL_API int tdl_maxpool_stats_rows(int N, int C, int P, int Q) {
  return N * (int)cdiv((int64_t)P * Q * (C / 8), kThreads);
}
TDL_API int tdl_maxpool_fwd_stats(const void* x, void* y, void* idx_u8, int N, int H, int W, int C, int P, int Q,
                                  int pad_h, int pad_w, float* stats_partial, int* nparts, void* stream) {
  TDL_REQUIRE(C % 8 == 0 && kThreads % (C / 8) == 0, "C/8 must divide 256");
  TDL_REQUIRE(N >= 1 && N <= 65535 && (int64_t)P * Q * C < (1ll << 31), "shape out of range for the fused pool");
  TDL_REQUIRE(stats_partial != nullptr && nparts != nullptr, "statistics buffers required");
  const unsigned gx = (unsigned)cdiv((int64_t)P * Q * (C / 8), kThreads);
  cudaError_t e = tdl::launch_pdl(maxpool_fwd_rows_kernel<3, 2, true>, dim3(gx, N), dim3(kThreads), 0,
                                  (cudaStream_t)stream, (const bf16*)x, (bf16*)y, (uint8_t*)idx_u8, H, W, C, P, Q,
                                  pad_h, pad_w, stats_partial);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "maxpool_fwd_rows_kernel<stats>");
  TDL_LAUNCH_OK("maxpool_fwd_rows_kernel<stats>");
  *nparts = (int)gx * N;
  return 0;
}

TDL_API int tdl_maxpool_bwd(const void* dy, const void* idx_u8, void* dx, int N, int H, int W, int C, int P, int Q,
                            int k, int stride, int pad_h, int pad_w, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  const int64_t work = (int64_t)N * H * W * (C / 8);
  if (work == 0) return 0;
  if (k == 3 && stride == 2 && N <= 65535 && (int64_t)H * W * C < (1ll << 31)) {
    const int A = (H + pad_h + 1) / 2, B = (W + pad_w + 1) / 2;  // 2x2 pixel blocks on the window grid
    {
      cudaError_t _le = tdl::launch_pdl(maxpool_bwd_block_kernel<3, 2>,
          dim3(dim3((unsigned)cdiv((int64_t)A * B * (C / 8), kThreads), N)), dim3(kThreads), 0, (cudaStream_t)stream,
          (const bf16*)dy, (const uint8_t*)idx_u8, (bf16*)dx, H, W, C, P, Q, pad_h, pad_w, A, B);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "maxpool_bwd_block_kernel");
    }
  } else
    {
      cudaError_t _le = tdl::launch_pdl(maxpool_bwd_kernel, dim3(grid_for(work)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const bf16*)dy, (const uint8_t*)idx_u8, (bf16*)dx, N, H, W, C, P, Q, k, stride,
          pad_h, pad_w);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "maxpool_bwd_kernel");
    }
  TDL_LAUNCH_OK("maxpool_bwd_kernel");
  return 0;
}

TDL_API int tdl_gap_fwd(const void* x, void* y, int N, int HW, int C, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  if (N == 0) return 0;
  dim3 grid(N, (unsigned)cdiv(C / 8, 32));
  {
    cudaError_t _le = tdl::launch_pdl(gap_fwd_kernel, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream,
        (const bf16*)x, (bf16*)y, HW, C);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "gap_fwd_kernel");
  }
  TDL_LAUNCH_OK("gap_fwd_kernel");
  return 0;
}

TDL_API int tdl_gap_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  const int64_t total = (int64_t)N * HW * (C / 8);
  if (total == 0) return 0;
  {
    cudaError_t _le = tdl::launch_pdl(gap_bwd_kernel, dim3(grid_for(total)), dim3(kThreads), 0, (cudaStream_t)stream,
        (const bf16*)dy, (bf16*)dx, total, HW, C / 8);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "gap_bwd_kernel");
  }
  TDL_LAUNCH_OK("gap_bwd_kernel");
  return 0;
}

// y[N, K] (bf16) = mean_hw(x[N, HW, C]) . w[K, C]^T (fp32 master weights); gap_out [N, C] fp32 keeps the pooled vectors
// for the weight gradient.  C % 8 == 0, C <= 12288 (the pooled vector lives in shared memory).
TDL_API int tdl_gap_conv1x1_fwd(const void* x, const float* w, void* y, float* gap_out, int N, int HW, int C, int K,
                                void* stream) {
  TDL_REQUIRE(C % 8 == 0 && C > 0 && C <= 12288, "C must be a multiple of 8, at most 12288");
  TDL_REQUIRE(N >= 1 && K >= 1 && HW >= 1 && gap_out != nullptr, "bad shape");
  dim3 grid(N, (unsigned)cdiv(K, kGapKT));
  cudaError_t e = tdl::launch_pdl(gap_conv1x1_fwd_kernel, grid, dim3(kThreads), (size_t)C * sizeof(float),
                                  (cudaStream_t)stream, (const bf16*)x, w, (bf16*)y, gap_out, HW, C, K);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "gap_conv1x1_fwd_kernel");
  TDL_LAUNCH_OK("gap_conv1x1_fwd_kernel");
  return 0;
}
// dx [N, HW, C] bf16 (written) and dw [K, C] fp32 (ACCUMULATED into) of the fused pool + 1x1; either may be NULL.
TDL_API int tdl_gap_conv1x1_bwd(const void* dy, const float* w, const float* gap, void* dx, float* dw, int N, int HW,
                                int C, int K, void* stream) {
  TDL_REQUIRE(C % 8 == 0 && C > 0, "C must be a multiple of 8");
  cudaStream_t st = (cudaStream_t)stream;
  if (dx) {
    dim3 grid(N, (unsigned)cdiv(C, kThreads));
    cudaError_t e = tdl::launch_pdl(gap_conv1x1_bwd_data_kernel, grid, dim3(kThreads), 0, st, (const bf16*)dy, w,
                                    (bf16*)dx, HW, C, K);
    if (e != cudaSuccess) return tdl::cuda_fail(e, "gap_conv1x1_bwd_data_kernel");
    TDL_LAUNCH_OK("gap_conv1x1_bwd_data_kernel");
  }
  if (dw) {
    TDL_REQUIRE(gap != nullptr, "the pooled vectors are needed for the weight gradient");
    dim3 grid(K, (unsigned)cdiv(C, kThreads));
    cudaError_t e = tdl::launch_pdl(gap_conv1x1_bwd_weight_kernel, grid, dim3(kThreads), 0, st, (const bf16*)dy, gap,
                                    dw, N, C, K);
    if (e != cudaSuccess) return tdl::cuda_fail(e, "gap_conv1x1_bwd_weight_kernel");
    TDL_LAUNCH_OK("gap_conv1x1_bwd_weight_kernel");
  }
  return 0;
}

TDL_API int tdl_depthwise_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C,
                              int P, int Q, int R, int S, int stride, int dil, int pad_h, int pad_w, int relu,
                              void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  const int64_t work = (int64_t)N * P * Q * (C / 8);
  if (work == 0) return 0;
  {
    cudaError_t _le = tdl::launch_pdl(depthwise_fwd_kernel, dim3(grid_for(work)), dim3(kThreads), 0,
        (cudaStream_t)stream, (const bf16*)x, w, bias, (bf16*)y, N, H, W, C, P, Q, R, S, stride, dil, pad_h, pad_w,
        relu);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise_fwd_kernel");
  }
  TDL_LAUNCH_OK("depthwise_fwd_kernel");
  return 0;
}

// dx written; dw[R*S, C] and dbias[C] are accumulated into (caller zeroes them).
TDL_API int tdl_depthwise_bwd(const void* x, const float* w, const void* dy, const void* y, void* dx, float* dw,
                              float* dbias, int N, int H, int W, int C, int P, int Q, int R, int S, int stride,
                              int dil, int pad_h, int pad_w, int relu, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  TDL_REQUIRE(R * S <= 9, "depthwise weight gradient supports up to 9 taps");
  cudaStream_t st = (cudaStream_t)stream;
  if (dx) {
    const int64_t work = (int64_t)N * H * W * (C / 8);
    if (work)
      {
        cudaError_t _le = tdl::launch_pdl(depthwise_bwd_data_kernel, dim3(grid_for(work)), dim3(kThreads), 0, st,
            (const bf16*)dy, (const bf16*)y, w, (bf16*)dx, N, H, W, C, P, Q, R, S, stride, dil, pad_h, pad_w, relu);
        if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise_bwd_data_kernel");
      }
    TDL_LAUNCH_OK("depthwise_bwd_data_kernel");
  }
  if (dw) {
    const int c8 = C / 8;
    int tx = 1;
    while (tx < 32 && tx < c8) tx <<= 1;
    const int ty = kThreads / tx;
    const int gy = (int)cdiv(c8, tx);
    const int64_t pixels = (int64_t)N * P * Q;
    int64_t gx = cdiv(pixels, (int64_t)ty * 8);
    const int64_t cap = (int64_t)num_sms() * 4 / gy > 0 ? (int64_t)num_sms() * 4 / gy : 1;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid((unsigned)gx, gy);
    {
      cudaError_t _le = tdl::launch_pdl(depthwise_bwd_weight_kernel, dim3(grid), dim3(kThreads),
          kThreads * 8 * sizeof(float), st, (const bf16*)x, (const bf16*)dy, (const bf16*)y, dw, dbias, N, H, W, C, P,
          Q, R, S, stride, dil, pad_h, pad_w, relu, tx, ty);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise_bwd_weight_kernel");
    }
    TDL_LAUNCH_OK("depthwise_bwd_weight_kernel");
  }
  return 0;
}

// ldy: elements between consecutive output pixels of y (>= C): y may be a channel slice of a wider buffer, so the
// result lands where a concat would have copied it.
TDL_API int tdl_upsample_fwd_ld(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int64_t ldy,
                                int is_bf16, void* stream) {
  TDL_REQUIRE(ldy >= C, "ldy must be at least C");
  const int64_t total = (int64_t)N * OH * OW * C;
  if (total == 0) return 0;
  if (is_bf16)
    {
      cudaError_t _le = tdl::launch_pdl(upsample_fwd_kernel<bf16>, dim3(grid_for(total)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const bf16*)x, (bf16*)y, N, H, W, C, OH, OW, ldy);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "upsample_fwd_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(upsample_fwd_kernel<float>, dim3(grid_for(total)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const float*)x, (float*)y, N, H, W, C, OH, OW, ldy);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "upsample_fwd_kernel");
    }
  TDL_LAUNCH_OK("upsample_fwd_kernel");
  return 0;
}

TDL_API int tdl_upsample_fwd(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int is_bf16,
                             void* stream) {
  return tdl_upsample_fwd_ld(x, y, N, H, W, C, OH, OW, C, is_bf16, stream);
}

// dx (fp32, caller-zeroed) += scatter of dy through the same interpolation weights; ld_dy as ldy above
TDL_API int tdl_upsample_bwd_ld(const void* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int64_t ld_dy,
                                int is_bf16, void* stream) {
  TDL_REQUIRE(ld_dy >= C, "ld_dy must be at least C");
  const int64_t total = (int64_t)N * OH * OW * C;
  if (total == 0) return 0;
  if (is_bf16)
    {
      cudaError_t _le = tdl::launch_pdl(upsample_bwd_kernel<bf16>, dim3(grid_for(total)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const bf16*)dy, dx, N, H, W, C, OH, OW, ld_dy);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "upsample_bwd_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(upsample_bwd_kernel<float>, dim3(grid_for(total)), dim3(kThreads), 0,
          (cudaStream_t)stream, (const float*)dy, dx, N, H, W, C, OH, OW, ld_dy);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "upsample_bwd_kernel");
    }
  TDL_LAUNCH_OK("upsample_bwd_kernel");
  return 0;
}
TDL_API int tdl_upsample_bwd(const void* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int is_bf16,
                             void* stream) {
  return tdl_upsample_bwd_ld(dy, dx, N, H, W, C, OH, OW, C, is_bf16, stream);
}

// ------------------------------------------------------------------------------------------
// input pipeline tail on the device: uint8 NHWC -> normalised bf16 NHWC (optionally channel-padded)
// ------------------------------------------------------------------------------------------
namespace tdl {
namespace spatial {
__global__ void __launch_bounds__(kThreads)
u8_to_bf16_kernel(const uint8_t* __restrict__ x, bf16* __restrict__ y, int64_t pixels, int C, int Cpad,
                  const float* __restrict__ mean, const float* __restrict__ inv_std) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = pixels * Cpad;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cpad);
    const int64_t px = i / Cpad;
    float v = 0.f;
    if (c < C) v = ((float)x[px * C + c] * (1.f / 255.f) - mean[c]) * inv_std[c];
    y[i] = __float2bfloat16(v);
  }
}
// Unpadded images (Cpad == C <= 4, element count a multiple of 16): 16 bytes in, two 16-byte vectors out per thread;
// the channel of element j is (first + j) % C with the per-channel constants in registers.
template <int C>
__global__ void __launch_bounds__(kThreads)
u8_to_bf16_vec_kernel(const uint8_t* __restrict__ x, bf16* __restrict__ y, int64_t groups,
                      const float* __restrict__ mean, const float* __restrict__ inv_std) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  float m[C], s[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    m[c] = mean[c];
    s[c] = inv_std[c];
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
    const uint4 in = __ldg(reinterpret_cast<const uint4*>(x) + i);
    const uint32_t wds[4] = {in.x, in.y, in.z, in.w};
    int c = (int)((i * 16) % C);
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float u = (float)((wds[j >> 2] >> (8 * (j & 3))) & 0xffu);
      float mc = m[0], sc = s[0];
#pragma unroll
      for (int k = 1; k < C; ++k)
        if (c == k) {
          mc = m[k];
          sc = s[k];
        }
      v[j] = (u * (1.f / 255.f) - mc) * sc;  // the same expression as the scalar kernel: identical rounding
      if (++c == C) c = 0;
    }
    st8(y + i * 16, pack8(v));
    st8(y + i * 16 + 8, pack8(v + 8));
  }
}
}  // namespace spatial
}  // namespace tdl

TDL_API int tdl_u8_to_bf16_normalize(const void* x_u8, void* y_bf16, int64_t pixels, int C, int Cpad,
                                     const float* mean, const float* inv_std, void* stream) {
  TDL_REQUIRE(Cpad >= C, "Cpad must be >= C");
  const int64_t total = pixels * Cpad;
  if (total == 0) return 0;
  if (Cpad == C && C >= 1 && C <= 4 && total % 16 == 0 && (reinterpret_cast<uintptr_t>(x_u8) & 15) == 0 &&
      (reinterpret_cast<uintptr_t>(y_bf16) & 15) == 0) {
    const int64_t groups = total / 16;
    const dim3 grid(tdl::spatial::grid_for(groups)), block(tdl::spatial::kThreads);
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t _le;
    switch (C) {
      case 1: _le = tdl::launch_pdl(tdl::spatial::u8_to_bf16_vec_kernel<1>, grid, block, 0, st, (const uint8_t*)x_u8, (tdl::bf16*)y_bf16, groups, mean, inv_std); break;
      case 2: _le = tdl::launch_pdl(tdl::spatial::u8_to_bf16_vec_kernel<2>, grid, block, 0, st, (const uint8_t*)x_u8, (tdl::bf16*)y_bf16, groups, mean, inv_std); break;
      case 3: _le = tdl::launch_pdl(tdl::spatial::u8_to_bf16_vec_kernel<3>, grid, block, 0, st, (const uint8_t*)x_u8, (tdl::bf16*)y_bf16, groups, mean, inv_std); break;
      default: _le = tdl::launch_pdl(tdl::spatial::u8_to_bf16_vec_kernel<4>, grid, block, 0, st, (const uint8_t*)x_u8, (tdl::bf16*)y_bf16, groups, mean, inv_std); break;
    }
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "tdl::spatial::u8_to_bf16_vec_kernel");
    TDL_LAUNCH_OK("u8_to_bf16_vec_kernel");
    return 0;
  }
  {
    cudaError_t _le = tdl::launch_pdl(tdl::spatial::u8_to_bf16_kernel, dim3(tdl::spatial::grid_for(total)),
        dim3(tdl::spatial::kThreads), 0, (cudaStream_t)stream, (const uint8_t*)x_u8, (tdl::bf16*)y_bf16, pixels, C,
        Cpad, mean, inv_std);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "tdl::spatial::u8_to_bf16_kernel");
  }
  TDL_LAUNCH_OK("u8_to_bf16_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------
// bias gradient: out[c] += sum over rows of dy[rows, C] (bf16 in, fp32 accumulate, red.add per CTA)
// ------------------------------------------------------------------------------------------
namespace tdl {
namespace spatial {
__global__ void __launch_bounds__(kThreads)
colsum_kernel(const bf16* __restrict__ dy, float* __restrict__ out, int64_t rows, int C, int tx, int ty) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float sm[];  // [ty][tx][8]
  const int c8 = C >> 3;
  const int cx = threadIdx.x % tx, ry = threadIdx.x / tx;
  const int cv = blockIdx.y * tx + cx;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  if (cv < c8) {
    const int64_t step = (int64_t)gridDim.x * ty;
    int64_t r = (int64_t)blockIdx.x * ty + ry;
    for (; r + step < rows; r += 2 * step) {  // two independent 16-byte loads in flight
      float a[8], b[8];
      unpack8(ld8(dy + r * C + cv * 8), a);
      unpack8(ld8(dy + (r + step) * C + cv * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += a[j] + b[j];
    }
    for (; r < rows; r += step) {
      float a[8];
      unpack8(ld8(dy + r * C + cv * 8), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += a[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) sm[(ry * tx + cx) * 8 + j] = s[j];
  __syncthreads();
  for (int idx = threadIdx.x; idx < tx * 8; idx += kThreads) {
    const int col = idx >> 3;
    const int cvv = blockIdx.y * tx + col;
    if (cvv >= c8) continue;
    float acc = 0.f;
    for (int t = 0; t < ty; ++t) acc += sm[(t * tx + col) * 8 + (idx & 7)];
    atomicAdd(out + cvv * 8 + (idx & 7), acc);
  }
}
}  // namespace spatial
}  // namespace tdl

// out[C] (fp32) += column sums of dy[rows, C] (bf16): the bias gradient of a convolution.  Caller zeroes `out` if needed.
TDL_API int tdl_colsum_bf16(const void* dy, float* out, int64_t rows, int C, void* stream) {
  using namespace tdl::spatial;
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  if (rows <= 0) return 0;
  const int c8 = C / 8;
  int tx = 1;
  while (tx < 32 && tx < c8) tx <<= 1;
  const int ty = kThreads / tx;
  const int gy = (int)tdl::cdiv(c8, tx);
  int64_t gx = tdl::cdiv(rows, (int64_t)ty * 8);
  const int64_t cap = (int64_t)tdl::num_sms() * 4 / gy;
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  {
    cudaError_t _le = tdl::launch_pdl(colsum_kernel, dim3(dim3((unsigned)gx, (unsigned)gy)), dim3(kThreads),
        kThreads * 8 * sizeof(float), (cudaStream_t)stream, (const tdl::bf16*)dy, out, rows, C, tx, ty);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "colsum_kernel");
  }
  TDL_LAUNCH_OK("colsum_kernel");
  return 0;
}
