// 3x3 depthwise convolution for NHWC bf16 (forward, data gradient, weight gradient), memory-roofline oriented.
//
// reference call sites: core/xception.py:102-110 (every Xception separable conv, stride 1/2, rate from the atrous
// bookkeeping) and core/layers.py:34-42 (ASPP, rates 2/4/8, bias + ReLU).  Depthwise conv has no reuse across
// channels, so tensor cores are pointless; the job is to touch each activation byte once and keep the instruction
// count per element under the ~20 lane-instructions that the HBM roofline allows:
//   * a thread owns 4 channels for all of its pixels: the 9x4 filter taps (and bias) sit in registers
//   * it produces strips of 4 horizontally adjacent outputs, so for stride 1 / rate 1 each input row is loaded
//     as 6 vectors that feed 12 tap uses (instead of 12 loads)
//   * the Xception pre-activation ReLU is applied on load (`relu_in`), the ASPP ReLU on store (`relu_out`)
//   * weight gradient: same strip walk, 36+4 accumulators per thread, block reduction in shared memory, one fp32
//     atomic per (tap, channel) per CTA.
#include "tdl_common.cuh"

namespace tdl {
namespace dw {

constexpr int kThreads = 256;
constexpr int kStrip = 4;

struct Shape {
  int N, H, W, C, P, Q, stride, dil, pad_h, pad_w;
};

__device__ __forceinline__ void ld4(const bf16* p, float* f, bool relu) {
  const uint2 raw = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
  if (relu) {
#pragma unroll
    for (int j = 0; j < 4; ++j) f[j] = fmaxf(f[j], 0.f);
  }
}
__device__ __forceinline__ void zero4(float* f) { f[0] = f[1] = f[2] = f[3] = 0.f; }
__device__ __forceinline__ void st4(bf16* p, const float* f) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]), b = __floats2bfloat162_rn(f[2], f[3]);
  uint2 raw;
  raw.x = *reinterpret_cast<const uint32_t*>(&a);
  raw.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = raw;
}

// kDense: stride 1 and dilation 1 (input columns of neighbouring outputs overlap -> 6 loads per row per strip)
// out = relu_out?( bias + sum_taps w[tap] * relu_in?(x) ), optionally zeroed where mask <= 0 (data-gradient use)
template <bool kDense>
__global__ void __launch_bounds__(kThreads)
depthwise3x3_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    const bf16* __restrict__ mask, bf16* __restrict__ y, Shape s, int tx, int ty, int relu_in,
                    int relu_out) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int cx = threadIdx.x % tx, sy = threadIdx.x / tx;
  const int c4 = s.C / 4;
  const int cv = blockIdx.y * tx + cx;
  if (cv >= c4) return;
  float wt[9][4], b4[4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4 v = *reinterpret_cast<const float4*>(w + (size_t)t * s.C + cv * 4);
    wt[t][0] = v.x; wt[t][1] = v.y; wt[t][2] = v.z; wt[t][3] = v.w;
  }
  if (bias) {
    const float4 v = *reinterpret_cast<const float4*>(bias + cv * 4);
    b4[0] = v.x; b4[1] = v.y; b4[2] = v.z; b4[3] = v.w;
  } else {
    zero4(b4);
  }
  const int qs = (s.Q + kStrip - 1) / kStrip;
  const int64_t strips = (int64_t)s.N * s.P * qs;
  for (int64_t st = (int64_t)blockIdx.x * ty + sy; st < strips; st += (int64_t)gridDim.x * ty) {
    const int q0 = (int)(st % qs) * kStrip;
    const int p = (int)((st / qs) % s.P);
    const int n = (int)(st / ((int64_t)qs * s.P));
    float acc[kStrip][4];
#pragma unroll
    for (int j = 0; j < kStrip; ++j) {
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[j][c] = b4[c];
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = p * s.stride + r * s.dil - s.pad_h;
      if (h < 0 || h >= s.H) continue;
      const bf16* row = x + (((int64_t)n * s.H + h) * s.W) * s.C + cv * 4;
      if (kDense) {
        float v[kStrip + 2][4];
#pragma unroll
        for (int c = 0; c < kStrip + 2; ++c) {
          const int col = q0 + c - s.pad_w;
          if (col >= 0 && col < s.W) ld4(row + (int64_t)col * s.C, v[c], relu_in);
          else zero4(v[c]);
        }
#pragma unroll
        for (int j = 0; j < kStrip; ++j)
#pragma unroll
          for (int t = 0; t < 3; ++t)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[j][c] = fmaf(v[j + t][c], wt[r * 3 + t][c], acc[j][c]);
      } else {
#pragma unroll
        for (int j = 0; j < kStrip; ++j)
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            const int col = (q0 + j) * s.stride + t * s.dil - s.pad_w;
            if (col >= 0 && col < s.W) {
              float v[4];
              ld4(row + (int64_t)col * s.C, v, relu_in);
#pragma unroll
              for (int c = 0; c < 4; ++c) acc[j][c] = fmaf(v[c], wt[r * 3 + t][c], acc[j][c]);
            }
          }
      }
    }
#pragma unroll
    for (int j = 0; j < kStrip; ++j) {
      const int q = q0 + j;
      if (q >= s.Q) break;
      const int64_t o = ((((int64_t)n * s.P + p) * s.Q + q) * s.C) + cv * 4;
      if (relu_out) {
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[j][c] = fmaxf(acc[j][c], 0.f);
      }
      if (mask) {
        float m[4];
        ld4(mask + o, m, false);
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[j][c] = m[c] > 0.f ? acc[j][c] : 0.f;
      }
      st4(y + o, acc[j]);
    }
  }
}

// dw[tap][c] += sum_pixels dyv * relu_in?(x shifted);  dbias[c] += sum dyv;  dyv = dy * (ymask > 0 if given)
template <bool kDense>
__global__ void __launch_bounds__(kThreads)
depthwise3x3_wgrad_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, const bf16* __restrict__ ymask,
                          float* __restrict__ dw, float* __restrict__ dbias, Shape s, int tx, int ty, int relu_in) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float sm[];  // [ty][tx][4]
  const int cx = threadIdx.x % tx, sy = threadIdx.x / tx;
  const int c4 = s.C / 4;
  const int cv = blockIdx.y * tx + cx;
  float acc[9][4], accb[4];
#pragma unroll
  for (int t = 0; t < 9; ++t) zero4(acc[t]);
  zero4(accb);
  if (cv < c4) {
    const int qs = (s.Q + kStrip - 1) / kStrip;
    const int64_t strips = (int64_t)s.N * s.P * qs;
    for (int64_t st = (int64_t)blockIdx.x * ty + sy; st < strips; st += (int64_t)gridDim.x * ty) {
      const int q0 = (int)(st % qs) * kStrip;
      const int p = (int)((st / qs) % s.P);
      const int n = (int)(st / ((int64_t)qs * s.P));
      float g[kStrip][4];
#pragma unroll
      for (int j = 0; j < kStrip; ++j) {
        const int q = q0 + j;
        if (q < s.Q) {
          const int64_t o = ((((int64_t)n * s.P + p) * s.Q + q) * s.C) + cv * 4;
          ld4(dy + o, g[j], false);
          if (ymask) {
            float m[4];
            ld4(ymask + o, m, false);
#pragma unroll
            for (int c = 0; c < 4; ++c) g[j][c] = m[c] > 0.f ? g[j][c] : 0.f;
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) accb[c] += g[j][c];
        } else {
          zero4(g[j]);
        }
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int h = p * s.stride + r * s.dil - s.pad_h;
        if (h < 0 || h >= s.H) continue;
        const bf16* row = x + (((int64_t)n * s.H + h) * s.W) * s.C + cv * 4;
        if (kDense) {
          float v[kStrip + 2][4];
#pragma unroll
          for (int c = 0; c < kStrip + 2; ++c) {
            const int col = q0 + c - s.pad_w;
            if (col >= 0 && col < s.W) ld4(row + (int64_t)col * s.C, v[c], relu_in);
            else zero4(v[c]);
          }
#pragma unroll
          for (int j = 0; j < kStrip; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t)
#pragma unroll
              for (int c = 0; c < 4; ++c) acc[r * 3 + t][c] = fmaf(g[j][c], v[j + t][c], acc[r * 3 + t][c]);
        } else {
#pragma unroll
          for (int j = 0; j < kStrip; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              const int col = (q0 + j) * s.stride + t * s.dil - s.pad_w;
              if (col >= 0 && col < s.W) {
                float v[4];
                ld4(row + (int64_t)col * s.C, v, relu_in);
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r * 3 + t][c] = fmaf(g[j][c], v[c], acc[r * 3 + t][c]);
              }
            }
        }
      }
    }
  }
  // reduce the ty strip-threads of every channel vector through shared memory, one statistic at a time
  for (int t = 0; t <= 9; ++t) {
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v = accb[c];
#pragma unroll
      for (int u = 0; u < 9; ++u)
        if (u == t) v = acc[u][c];
      sm[((size_t)sy * tx + cx) * 4 + c] = v;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < tx * 4; idx += kThreads) {
      const int col = idx >> 2, c = idx & 3;
      const int cvv = blockIdx.y * tx + col;
      if (cvv >= c4) continue;
      float a = 0.f;
      for (int u = 0; u < ty; ++u) a += sm[((size_t)u * tx + col) * 4 + c];
      if (t < 9) atomicAdd(dw + (size_t)t * s.C + cvv * 4 + c, a);
      else if (dbias) atomicAdd(dbias + cvv * 4 + c, a);
    }
  }
}

struct Launch {
  int tx, ty, gx, gy;
};
static Launch plan(const Shape& s, int blocks_per_sm) {
  Launch l;
  const int c4 = s.C / 4;
  l.tx = 1;
  while (l.tx < 64 && l.tx < c4) l.tx <<= 1;
  l.ty = kThreads / l.tx;
  l.gy = (int)cdiv(c4, l.tx);
  const int64_t strips = (int64_t)s.N * s.P * cdiv(s.Q, kStrip);
  int64_t want = cdiv(strips, (int64_t)l.ty * 2);
  int64_t cap = (int64_t)num_sms() * blocks_per_sm / l.gy;
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  l.gx = (int)want;
  return l;
}

// Data gradient of a STRIDED 3x3 depthwise conv (dilation 1): a gather over the <= ceil(3/S)^2 output pixels whose
// window covers this input pixel, one thread per (input pixel, 8 channels), blockIdx.y = image.  The pre-activation
// ReLU mask (xmask <= 0 -> 0) is applied on the store, so no separate mask pass runs (core/xception.py:184-210).
template <int S>
__global__ void __launch_bounds__(kThreads)
depthwise3x3_dgrad_strided_kernel(const bf16* __restrict__ dy, const float* __restrict__ w, const bf16* __restrict__ xmask,
                                  bf16* __restrict__ dx, int H, int W, int C, int P, int Q, int pad_h, int pad_w) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  constexpr int kCells = (3 + S - 1) / S;
  const int c8 = C >> 3;
  const int n = blockIdx.y;
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= H * W * c8) return;
  const int hw = i / c8;
  const int cv = i - hw * c8;
  const int h = hw / W;
  const int wq = hw - h * W;
  const int hp = h + pad_h, wp = wq + pad_w;
  const int p_hi = hp / S, q_hi = wp / S;
  const bf16* dyn = dy + (int64_t)n * P * Q * C + cv * 8;
  const int64_t xoff = ((int64_t)n * H * W + hw) * C + cv * 8;
  bf16x8 g[kCells * kCells];
  int tap[kCells * kCells];
#pragma unroll
  for (int a = 0; a < kCells; ++a)
#pragma unroll
    for (int b = 0; b < kCells; ++b) {
      const int p = p_hi - a, q = q_hi - b;
      const int r = hp - p * S, t = wp - q * S;
      const bool ok = p >= 0 && p < P && q >= 0 && q < Q && r < 3 && t < 3;
      tap[a * kCells + b] = ok ? r * 3 + t : -1;
      if (ok) g[a * kCells + b] = ld8(dyn + ((int64_t)p * Q + q) * C);
    }
  bf16x8 xm;
  if (xmask) xm = ld8(xmask + xoff);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
  for (int c = 0; c < kCells * kCells; ++c) {
    if (tap[c] < 0) continue;
    float gv[8];
    unpack8(g[c], gv);
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + (int64_t)tap[c] * C + cv * 8));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + (int64_t)tap[c] * C + cv * 8 + 4));
    acc[0] = fmaf(gv[0], w0.x, acc[0]); acc[1] = fmaf(gv[1], w0.y, acc[1]);
    acc[2] = fmaf(gv[2], w0.z, acc[2]); acc[3] = fmaf(gv[3], w0.w, acc[3]);
    acc[4] = fmaf(gv[4], w1.x, acc[4]); acc[5] = fmaf(gv[5], w1.y, acc[5]);
    acc[6] = fmaf(gv[6], w1.z, acc[6]); acc[7] = fmaf(gv[7], w1.w, acc[7]);
  }
  if (xmask) {
    float mv[8];
    unpack8(xm, mv);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = mv[j] > 0.f ? acc[j] : 0.f;
  }
  st8(dx + xoff, pack8(acc));
}

}  // namespace dw
}  // namespace tdl

using namespace tdl;
using namespace tdl::dw;

// y = relu_out?(bias + depthwise3x3(relu_in?(x)));  w is [9, C] fp32 (tap-major), C % 4 == 0.
// `mask` (same shape as y, may be NULL) zeroes outputs where mask <= 0: used when this kernel computes a data
// gradient (x := dy, w := flipped taps, mask := the forward input of a pre-activation ReLU).
TDL_API int tdl_depthwise3x3_fwd(const void* x, const float* w, const float* bias, const void* mask, void* y, int N,
                                 int H, int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w,
                                 int relu_in, int relu_out, void* stream) {
  TDL_REQUIRE(C % 4 == 0, "C must be a multiple of 4");
  if ((int64_t)N * P * Q == 0) return 0;
  Shape s = {N, H, W, C, P, Q, stride, dil, pad_h, pad_w};
  Launch l = plan(s, 8);
  dim3 grid(l.gx, l.gy);
  cudaStream_t st = (cudaStream_t)stream;
  if (stride == 1 && dil == 1)
    {
      cudaError_t _le = tdl::launch_pdl(depthwise3x3_kernel<true>, dim3(grid), dim3(kThreads), 0, st, (const bf16*)x,
          w, bias, (const bf16*)mask, (bf16*)y, s, l.tx, l.ty, relu_in, relu_out);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(depthwise3x3_kernel<false>, dim3(grid), dim3(kThreads), 0, st, (const bf16*)x,
          w, bias, (const bf16*)mask, (bf16*)y, s, l.tx, l.ty, relu_in, relu_out);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_kernel");
    }
  TDL_LAUNCH_OK("depthwise3x3_kernel");
  return 0;
}

// dx[N,H,W,C] = data gradient of a stride-2 3x3 depthwise conv (dilation 1) with weights w [9, C] (NOT flipped);
// xmask (may be NULL) zeroes dx where the forward input of a pre-activation ReLU was <= 0.
TDL_API int tdl_depthwise3x3_dgrad_s2(const void* dy, const float* w, const void* xmask, void* dx, int N, int H, int W,
                                      int C, int P, int Q, int pad_h, int pad_w, void* stream) {
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  TDL_REQUIRE(N <= 65535 && (int64_t)H * W * C < (1ll << 31), "shape too large for the strided depthwise dgrad kernel");
  if ((int64_t)N * H * W == 0) return 0;
  dim3 grid((unsigned)cdiv((int64_t)H * W * (C / 8), kThreads), (unsigned)N);
  {
    cudaError_t _le = tdl::launch_pdl(depthwise3x3_dgrad_strided_kernel<2>, dim3(grid), dim3(kThreads), 0,
        (cudaStream_t)stream, (const bf16*)dy, w, (const bf16*)xmask, (bf16*)dx, H, W, C, P, Q, pad_h, pad_w);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_dgrad_strided_kernel");
  }
  TDL_LAUNCH_OK("depthwise3x3_dgrad_strided_kernel");
  return 0;
}

// dw [9, C] and dbias [C] (may be NULL) are accumulated into (caller zeroes them); ymask (may be NULL) masks dy where
// the forward output was <= 0 (fused output ReLU).
TDL_API int tdl_depthwise3x3_wgrad(const void* x, const void* dy, const void* ymask, float* dw, float* dbias, int N,
                                   int H, int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w,
                                   int relu_in, void* stream) {
  TDL_REQUIRE(C % 4 == 0, "C must be a multiple of 4");
  if ((int64_t)N * P * Q == 0) return 0;
  Shape s = {N, H, W, C, P, Q, stride, dil, pad_h, pad_w};
  Launch l = plan(s, 4);
  dim3 grid(l.gx, l.gy);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)kThreads * 4 * sizeof(float);
  if (stride == 1 && dil == 1)
    {
      cudaError_t _le = tdl::launch_pdl(depthwise3x3_wgrad_kernel<true>, dim3(grid), dim3(kThreads), smem, st,
          (const bf16*)x, (const bf16*)dy, (const bf16*)ymask, dw, dbias, s, l.tx, l.ty, relu_in);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_wgrad_kernel");
    }
  else
    {
      cudaError_t _le = tdl::launch_pdl(depthwise3x3_wgrad_kernel<false>, dim3(grid), dim3(kThreads), smem, st,
          (const bf16*)x, (const bf16*)dy, (const bf16*)ymask, dw, dbias, s, l.tx, l.ty, relu_in);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_wgrad_kernel");
    }
  TDL_LAUNCH_OK("depthwise3x3_wgrad_kernel");
  return 0;
}

// ================================================================================================================
// TMA-tiled variant (stride 1/2, dilation 1/2): the haloed input patch of a (TH x TW) output tile and 64 channels is
// fetched by ONE 4-D TMA box into 128B-swizzled shared memory (out-of-bounds zero fill = the padding), double
// buffered; 256 threads then compute from shared memory -- 8 channels x a strip of 4 outputs per thread, the 9x8
// filter taps of the thread's channels in registers.  All global traffic is TMA loads + 16-byte stores.
// ================================================================================================================
#include <cudaTypedefs.h>

namespace tdl {
namespace dwt {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}
static int encode_nhwc(CUtensorMap* m, const void* base, int N, int H, int W, int C, int bw, int bh) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return tdl::fail(-3, "%s%s", "cuTensorMapEncodeTiled entry point unavailable", "");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * W * 2, (cuuint64_t)C * W * H * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, str, box, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {  // first driver call of this thread: bind the primary context, encode again
    cudaFree(nullptr);
    r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, str, box, es,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) return tdl::fail(-4, "%s%s", "cuTensorMapEncodeTiled failed (depthwise)", "");
  return 0;
}

// Output tile (TH x TW, TW a multiple of 4, TH*TW <= 128 so that 32 thread groups x 4-wide strips cover it): pick the
// shape that needs the fewest tiles for a P x Q map -- 19x19 is 4 tiles of 5x20 instead of 6 tiles of 8x16.
static void choose_tile(int P, int Q, int stride, int* th, int* tw) {
  const int max_out = stride == 1 ? 128 : 64;  // the strided input box is ~4x larger per output
  long best = -1;
  int bh = 8, bw = 16;
  for (int w = 8; w <= 32; w += 4) {
    int h = max_out / w;
    if (h > P) h = P;
    if (h < 1) continue;
    // a shorter tile that still needs the same number of tile rows wastes fewer threads
    const int rows = (int)tdl::cdiv(P, h);
    h = (int)tdl::cdiv(P, rows);
    const long tiles = (long)rows * tdl::cdiv(Q, w);
    const long cost = tiles * 1000 + (h * w);  // fewest tiles first, then the smaller box
    if (best < 0 || cost < best) {
      best = cost;
      bh = h;
      bw = w;
    }
  }
  *th = bh;
  *tw = bw;
}

struct TShape {
  int N, H, W, C, P, Q, stride, dil, pad_h, pad_w;
  int TH, TW, BH, BW;          // output tile, input box
  int tiles_h, tiles_w;
  int flip;                    // forward kernel: read the taps in reverse order (data gradient of a stride-1 layer)
};

constexpr int kT = 512;  // 16 warps: 16 channel-quads x 32 strip groups

// 4 channels (8 bytes) of pixel-row `lrow` of a 128B-swizzled [rows][64 ch] tile; cq = channel quad 0..15
__device__ __forceinline__ void lds4(const uint8_t* tile, int lrow, int cq, float* f, bool relu) {
  const uint2 raw = *reinterpret_cast<const uint2*>(tile + lrow * 128 + (((cq >> 1) ^ (lrow & 7)) << 4) + ((cq & 1) << 3));
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float2 v = __bfloat1622float2(h[i]);
    f[2 * i] = relu ? fmaxf(v.x, 0.f) : v.x;
    f[2 * i + 1] = relu ? fmaxf(v.y, 0.f) : v.y;
  }
}

// Same access through a precomputed 32-bit shared-space address (tile base + loop-invariant offset): no 64-bit
// generic address arithmetic and no swizzle math in the tile loop; the pre-activation ReLU is one packed bf16x2 max.
__device__ __forceinline__ uint32_t swz_off(int lrow, int cq) {
  return (uint32_t)(lrow * 128 + (((cq >> 1) ^ (lrow & 7)) << 4) + ((cq & 1) << 3));
}
__device__ __forceinline__ void lds4_at(uint32_t addr, float* f, bool relu) {
  uint32_t a, b;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
  if (relu) {
    const __nv_bfloat162 z = __floats2bfloat162_rn(0.f, 0.f);
    __nv_bfloat162 ha = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), z);
    __nv_bfloat162 hb = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&b), z);
    a = *reinterpret_cast<uint32_t*>(&ha);
    b = *reinterpret_cast<uint32_t*>(&hb);
  }
  f[0] = __uint_as_float(a << 16);
  f[1] = __uint_as_float(a & 0xffff0000u);
  f[2] = __uint_as_float(b << 16);
  f[3] = __uint_as_float(b & 0xffff0000u);
}

constexpr int kMaxStages = 4;

template <int STRIDE, int DIL>
__global__ void __launch_bounds__(kT)
depthwise3x3_tma_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_mask,
                        const float* __restrict__ w, const float* __restrict__ bias, int has_mask,
                        bf16* __restrict__ y, const TShape s, int stages, int relu_in, int relu_out,
                        float* __restrict__ stats, const bf16* __restrict__ addend) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int x_bytes = ((s.BH * s.BW * 128 + 1023) / 1024) * 1024;
  const int m_bytes = has_mask ? ((s.TH * s.TW * 128 + 1023) / 1024) * 1024 : 0;
  const int stage_bytes = x_bytes + m_bytes;
  __shared__ uint64_t full[kMaxStages];
  const int tid = threadIdx.x;
  const int cvec = tid & 15, grp = tid >> 4;
  const int c0 = blockIdx.y * 64;
  const int ch = c0 + cvec * 4;
  const bool ch_ok = ch < s.C;
  if (tid == 0) {
    ptx::prefetch_tmap(&tm_x);
    for (int i = 0; i < kMaxStages; ++i) ptx::mbar_init(&full[i], 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  float wt[9][4], b8[4];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int c = 0; c < 4; ++c) wt[t][c] = ch_ok ? w[(size_t)(s.flip ? 8 - t : t) * s.C + ch + c] : 0.f;
#pragma unroll
  for (int c = 0; c < 4; ++c) b8[c] = (bias && ch_ok) ? bias[ch + c] : 0.f;
  // optional fused BatchNorm statistics of the (bf16-rounded) output: a thread's four channels never change, so the
  // sums live in registers for the whole kernel and are folded once at the end into this CTA's partial row
  float ssum[4] = {0.f, 0.f, 0.f, 0.f}, ssq[4] = {0.f, 0.f, 0.f, 0.f};

  const int tiles = s.N * s.tiles_h * s.tiles_w;
  const uint32_t tx_bytes = (uint32_t)(s.BH * s.BW + (has_mask ? s.TH * s.TW : 0)) * 128u;
  auto issue = [&](int tile, int stage) {
    const int tw_i = tile % s.tiles_w;
    const int th_i = (tile / s.tiles_w) % s.tiles_h;
    const int n = tile / (s.tiles_w * s.tiles_h);
    ptx::mbar_expect_tx(&full[stage], tx_bytes);
    ptx::tma_load_4d(smem + stage * stage_bytes, &tm_x, &full[stage], c0, tw_i * s.TW * STRIDE - s.pad_w,
                     th_i * s.TH * STRIDE - s.pad_h, n);
    if (has_mask)
      ptx::tma_load_4d(smem + stage * stage_bytes + x_bytes, &tm_mask, &full[stage], c0, tw_i * s.TW, th_i * s.TH, n);
  };
  if (tid == 0)
    for (int i = 0; i < stages - 1; ++i) {
      const int t0 = blockIdx.x + i * gridDim.x;
      if (t0 < tiles) issue(t0, i);
    }
  const int strips_w = s.TW / 4;
  const int strips = s.TH * strips_w;
  // everything that depends only on the thread's strip is loop invariant: tile-local row indices of the taps
  const int ph = grp / strips_w, pw0 = (grp % strips_w) * 4;
  const bool active = grp < strips && ch_ok;
  // dense case: the 18 input and 4 mask addresses of this thread relative to the stage base never change
  uint32_t xoff[3][6], moff[4];
  if (STRIDE == 1 && DIL == 1) {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c) xoff[r][c] = swz_off((ph + r) * s.BW + pw0 + c, cvec);
#pragma unroll
    for (int j = 0; j < 4; ++j) moff[j] = (uint32_t)x_bytes + swz_off(ph * s.TW + pw0 + j, cvec);
  }
  const uint32_t smem_base = ptx::smem_u32(smem);
  int it = 0, stage = 0;
  uint32_t parity = 0;
  // tile coordinates advance by the grid stride without a division per tile
  const int per_img = s.tiles_w * s.tiles_h;
  const int step_n = (int)gridDim.x / per_img, step_r = (int)gridDim.x % per_img;
  const int step_h = step_r / s.tiles_w, step_w = step_r % s.tiles_w;
  int cur_n = (int)blockIdx.x / per_img, cur_h = ((int)blockIdx.x % per_img) / s.tiles_w,
      cur_w = ((int)blockIdx.x % per_img) % s.tiles_w;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
    const int tw_i = cur_w, th_i = cur_h, n = cur_n;
    cur_w += step_w;
    if (cur_w >= s.tiles_w) { cur_w -= s.tiles_w; ++cur_h; }
    cur_h += step_h;
    if (cur_h >= s.tiles_h) { cur_h -= s.tiles_h; ++cur_n; }
    cur_n += step_n;
    const int ahead = tile + (stages - 1) * gridDim.x;
    if (tid == 0 && ahead < tiles) issue(ahead, stage == 0 ? stages - 1 : stage - 1);  // consumed in iteration it-1
    ptx::mbar_wait(&full[stage], parity);
    const uint8_t* in = smem + stage * stage_bytes;
    const uint8_t* min_ = in + x_bytes;
    const uint32_t in_a = smem_base + (uint32_t)(stage * stage_bytes);
    if (++stage == stages) {
      stage = 0;
      parity ^= 1;
    }
    if (active) {
      const int p = th_i * s.TH + ph;
      const int q0 = tw_i * s.TW + pw0;
      const int64_t oidx = ((((int64_t)n * s.P + p) * s.Q + q0) * s.C) + ch;
      // data-gradient use: the other consumer's gradient of this strip, requested NOW so that the loads are in flight
      // while the taps are computed (issued next to their use they cost a memory round trip per output pixel)
      uint2 ad[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) ad[j] = make_uint2(0u, 0u);
      if (addend != nullptr && p < s.P) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (q0 + j < s.Q) ad[j] = __ldg(reinterpret_cast<const uint2*>(addend + oidx + (int64_t)j * s.C));
      }
      float acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[j][c] = b8[c];
      if (STRIDE == 1 && DIL == 1) {
        // dense case on packed fp32 FMAs (fma.rn.f32x2, two channels per instruction): the inner product is 144 scalar
        // FMAs per thread and tile, the largest single share of this issue-bound kernel's instructions
        float2 a2[4][2];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          a2[j][0] = make_float2(b8[0], b8[1]);
          a2[j][1] = make_float2(b8[2], b8[3]);
        }
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          float2 v2[6][2];
#pragma unroll
          for (int c = 0; c < 6; ++c) {
            float v[4];
            lds4_at(in_a + xoff[r][c], v, relu_in);
            v2[c][0] = make_float2(v[0], v[1]);
            v2[c][1] = make_float2(v[2], v[3]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              a2[j][0] = __ffma2_rn(v2[j + t][0], make_float2(wt[r * 3 + t][0], wt[r * 3 + t][1]), a2[j][0]);
              a2[j][1] = __ffma2_rn(v2[j + t][1], make_float2(wt[r * 3 + t][2], wt[r * 3 + t][3]), a2[j][1]);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc[j][0] = a2[j][0].x; acc[j][1] = a2[j][0].y; acc[j][2] = a2[j][1].x; acc[j][3] = a2[j][1].y;
        }
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int lrow0 = (ph * STRIDE + r * DIL) * s.BW + pw0 * STRIDE;
        if (STRIDE == 1 && DIL == 1) {
          // (done above)
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              float v[4];
              lds4(in, lrow0 + j * STRIDE + t * DIL, cvec, v, relu_in);
#pragma unroll
              for (int c = 0; c < 4; ++c) acc[j][c] = fmaf(v[c], wt[r * 3 + t][c], acc[j][c]);
            }
        }
      }
      if (p < s.P) {
        bf16* orow = y + oidx;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (q0 + j >= s.Q) break;
          if (relu_out) {
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[j][c] = fmaxf(acc[j][c], 0.f);
          }
          if (has_mask) {
            float m[4];
            if (STRIDE == 1 && DIL == 1) lds4_at(in_a + moff[j], m, false);
            else lds4(min_, ph * s.TW + pw0 + j, cvec, m, false);
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[j][c] = m[c] > 0.f ? acc[j][c] : 0.f;
          }
          if (addend) {  // added unmasked, after the input-ReLU mask
            acc[j][0] += __uint_as_float(ad[j].x << 16);
            acc[j][1] += __uint_as_float(ad[j].x & 0xffff0000u);
            acc[j][2] += __uint_as_float(ad[j].y << 16);
            acc[j][3] += __uint_as_float(ad[j].y & 0xffff0000u);
          }
          tdl::dw::st4(orow + (int64_t)j * s.C, acc[j]);
          if (stats) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const float v = __bfloat162float(__float2bfloat16_rn(acc[j][c]));  // the value as stored
              ssum[c] += v;
              ssq[c] = fmaf(v, v, ssq[c]);
            }
          }
        }
      }
    }
    __syncthreads();  // everyone is done reading this stage before it is refilled
  }
  if (stats) {
    // 32 pixel groups x 16 channel quads -> 64 channel sums, twice (sum, sum of squares); the stages are idle now
    float* red = reinterpret_cast<float*>(smem);
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      __syncthreads();
#pragma unroll
      for (int c = 0; c < 4; ++c) red[(grp * 16 + cvec) * 4 + c] = which ? ssq[c] : ssum[c];
      __syncthreads();
      if (tid < 64 && c0 + tid < s.C) {
        float a = 0.f;
        for (int gidx = 0; gidx < kT / 16; ++gidx) a += red[gidx * 64 + tid];
        stats[((size_t)blockIdx.x * 2 + which) * s.C + c0 + tid] = a;
      }
    }
  }
}

// Weight gradient on the same tiling: the haloed x patch and the dy tile arrive by TMA (zero fill outside the
// maps, so no bounds logic), every thread keeps 9x8 + 8 accumulators for its 8 channels across all of its CTA's
// tiles, and one shared-memory reduction + fp32 atomics per CTA finish the job.
template <int STRIDE, int DIL>
__global__ void __launch_bounds__(kT)
depthwise3x3_tma_wgrad_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_dy,
                              float* __restrict__ dw, float* __restrict__ dbias, const TShape s, int stages,
                              int relu_in) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int x_bytes = ((s.BH * s.BW * 128 + 1023) / 1024) * 1024;
  const int g_bytes = ((s.TH * s.TW * 128 + 1023) / 1024) * 1024;
  const int stage_bytes = x_bytes + g_bytes;
  __shared__ uint64_t full[kMaxStages];
  const int tid = threadIdx.x;
  const int cvec = tid & 15, grp = tid >> 4;
  const int c0 = blockIdx.y * 64;
  const int ch = c0 + cvec * 4;
  const bool ch_ok = ch < s.C;
  if (tid == 0) {
    ptx::prefetch_tmap(&tm_x);
    ptx::prefetch_tmap(&tm_dy);
    for (int i = 0; i < kMaxStages; ++i) ptx::mbar_init(&full[i], 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  float acc[9][4], accb[4];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[t][c] = 0.f;
#pragma unroll
  for (int c = 0; c < 4; ++c) accb[c] = 0.f;

  const int tiles = s.N * s.tiles_h * s.tiles_w;
  const uint32_t tx_bytes = (uint32_t)(s.BH * s.BW + s.TH * s.TW) * 128u;
  auto issue = [&](int tile, int stage) {
    const int tw_i = tile % s.tiles_w;
    const int th_i = (tile / s.tiles_w) % s.tiles_h;
    const int n = tile / (s.tiles_w * s.tiles_h);
    ptx::mbar_expect_tx(&full[stage], tx_bytes);
    ptx::tma_load_4d(smem + stage * stage_bytes, &tm_x, &full[stage], c0, tw_i * s.TW * STRIDE - s.pad_w,
                     th_i * s.TH * STRIDE - s.pad_h, n);
    ptx::tma_load_4d(smem + stage * stage_bytes + x_bytes, &tm_dy, &full[stage], c0, tw_i * s.TW, th_i * s.TH, n);
  };
  int it = 0;
  if (tid == 0)
    for (int i = 0; i < stages - 1; ++i) {
      const int t0 = blockIdx.x + i * gridDim.x;
      if (t0 < tiles) issue(t0, i);
    }
  const int strips_w = s.TW / 4;
  const int strips = s.TH * strips_w;
  const int ph = grp / strips_w, pw0 = (grp % strips_w) * 4;
  // loop-invariant shared-memory offsets of this thread's gradient strip and (dense case) input patch
  uint32_t xoff[3][6], goff[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) goff[j] = (uint32_t)x_bytes + swz_off(ph * s.TW + pw0 + j, cvec);
  if (STRIDE == 1 && DIL == 1) {
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 6; ++c) xoff[r][c] = swz_off((ph + r) * s.BW + pw0 + c, cvec);
  }
  const uint32_t smem_base = ptx::smem_u32(smem);
  int stage = 0;
  uint32_t parity = 0;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
    const int ahead = tile + (stages - 1) * gridDim.x;
    if (tid == 0 && ahead < tiles) issue(ahead, stage == 0 ? stages - 1 : stage - 1);
    ptx::mbar_wait(&full[stage], parity);
    const uint8_t* in = smem + stage * stage_bytes;
    const uint32_t in_a = smem_base + (uint32_t)(stage * stage_bytes);
    if (++stage == stages) {
      stage = 0;
      parity ^= 1;
    }
    if (grp < strips && ch_ok) {
      float g[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        lds4_at(in_a + goff[j], g[j], false);
#pragma unroll
        for (int c = 0; c < 4; ++c) accb[c] += g[j][c];
      }
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int lrow0 = (ph * STRIDE + r * DIL) * s.BW + pw0 * STRIDE;
        if (STRIDE == 1 && DIL == 1) {
          // packed fp32 FMAs (two channels per instruction), as in the forward kernel
          float2 v2[6][2], a2[3][2];
#pragma unroll
          for (int c = 0; c < 6; ++c) {
            float v[4];
            lds4_at(in_a + xoff[r][c], v, relu_in);
            v2[c][0] = make_float2(v[0], v[1]);
            v2[c][1] = make_float2(v[2], v[3]);
          }
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            a2[t][0] = make_float2(acc[r * 3 + t][0], acc[r * 3 + t][1]);
            a2[t][1] = make_float2(acc[r * 3 + t][2], acc[r * 3 + t][3]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 g0 = make_float2(g[j][0], g[j][1]), g1 = make_float2(g[j][2], g[j][3]);
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              a2[t][0] = __ffma2_rn(g0, v2[j + t][0], a2[t][0]);
              a2[t][1] = __ffma2_rn(g1, v2[j + t][1], a2[t][1]);
            }
          }
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            acc[r * 3 + t][0] = a2[t][0].x; acc[r * 3 + t][1] = a2[t][0].y;
            acc[r * 3 + t][2] = a2[t][1].x; acc[r * 3 + t][3] = a2[t][1].y;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              float v[4];
              lds4(in, lrow0 + j * STRIDE + t * DIL, cvec, v, relu_in);
#pragma unroll
              for (int c = 0; c < 4; ++c) acc[r * 3 + t][c] = fmaf(g[j][c], v[c], acc[r * 3 + t][c]);
            }
        }
      }
    }
    __syncthreads();
  }
  // CTA reduction: for each of the 10 statistics, 32 groups x 64 channels -> 64 sums -> one atomic each
  float* red = reinterpret_cast<float*>(smem);  // 512 x 4 floats = 8 KB (the stages are idle now)
  for (int t = 0; t <= 9; ++t) {
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v = accb[c];
#pragma unroll
      for (int u = 0; u < 9; ++u)
        if (u == t) v = acc[u][c];
      red[(grp * 16 + cvec) * 4 + c] = v;
    }
    __syncthreads();
    if (tid < 64) {
      const int cc = c0 + tid;  // channel
      if (cc < s.C) {
        float a = 0.f;
        for (int gidx = 0; gidx < 32; ++gidx) a += red[gidx * 64 + tid];
        if (t < 9) atomicAdd(dw + (size_t)t * s.C + cc, a);
        else if (dbias) atomicAdd(dbias + cc, a);
      }
    }
  }
}

}  // namespace dwt
}  // namespace tdl

// dw [9, C] / dbias [C] accumulated into; TMA-tiled (no output-ReLU mask on this path).
TDL_API int tdl_depthwise3x3_tma_wgrad(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W,
                                       int C, int P, int Q, int stride, int dil, int pad_h, int pad_w, int relu_in,
                                       void* stream) {
  using namespace tdl::dwt;
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  TDL_REQUIRE((stride == 1 || stride == 2) && (dil == 1 || dil == 2) && !(stride == 2 && dil == 2),
              "TMA depthwise supports stride 1/2 with dilation 1, or stride 1 with dilation 2");
  if ((int64_t)N * P * Q == 0) return 0;
  TShape s;
  s.N = N; s.H = H; s.W = W; s.C = C; s.P = P; s.Q = Q; s.stride = stride; s.dil = dil; s.pad_h = pad_h; s.pad_w = pad_w;
  s.flip = 0;
  choose_tile(P, Q, stride, &s.TH, &s.TW);
  s.BW = (s.TW - 1) * stride + 2 * dil + 1;
  s.BH = (s.TH - 1) * stride + 2 * dil + 1;
  s.tiles_h = (int)tdl::cdiv(P, s.TH);
  s.tiles_w = (int)tdl::cdiv(Q, s.TW);
  CUtensorMap tmx, tmg;
  int rc = encode_nhwc(&tmx, x, N, H, W, C, s.BW, s.BH);
  if (rc) return rc;
  rc = encode_nhwc(&tmg, dy, N, P, Q, C, s.TW, s.TH);
  if (rc) return rc;
  const int x_bytes = ((s.BH * s.BW * 128 + 1023) / 1024) * 1024;
  const int g_bytes = ((s.TH * s.TW * 128 + 1023) / 1024) * 1024;
  int stages = (150 * 1024) / (x_bytes + g_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) stages = 2;
  int smem = stages * (x_bytes + g_bytes) + 1024;
  if (smem < 8192 + 1024) smem = 8192 + 1024;
  const int tiles = N * s.tiles_h * s.tiles_w;
  const int gy = (int)tdl::cdiv(C, 64);
  int gx = (tdl::num_sms() * 2) / gy;
  if (gx < 1) gx = 1;
  if (gx > tiles) gx = tiles;
  dim3 grid(gx, gy);
  cudaStream_t st = (cudaStream_t)stream;
#define TDL_DWW_LAUNCH(S, D)                                                                                         \
  do {                                                                                                               \
    static bool attr = false;                                                                                        \
    if (!attr) {                                                                                                     \
      TDL_CUDA_OK(cudaFuncSetAttribute(depthwise3x3_tma_wgrad_kernel<S, D>,                                          \
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));                    \
      attr = true;                                                                                                   \
    }                                                                                                                \
    cudaError_t _le = tdl::launch_pdl(depthwise3x3_tma_wgrad_kernel<S, D>, dim3(grid), dim3(kT), smem, st, tmx, tmg, dw,  \
                                      dbias, s, stages, relu_in);                                                    \
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_tma_wgrad_kernel");                             \
  } while (0)
  if (stride == 1 && dil == 1) TDL_DWW_LAUNCH(1, 1);
  else if (stride == 1) TDL_DWW_LAUNCH(1, 2);
  else TDL_DWW_LAUNCH(2, 1);
#undef TDL_DWW_LAUNCH
  TDL_LAUNCH_OK("depthwise3x3_tma_wgrad_kernel");
  return 0;
}

static int depthwise3x3_tma_fwd_impl(const void* x, const float* w, const float* bias, const void* mask, void* y,
                                     int N, int H, int W, int C, int P, int Q, int stride, int dil, int pad_h,
                                     int pad_w, int relu_in, int relu_out, float* stats, int* nparts, void* stream,
                                     int flip_taps = 0, const void* addend = nullptr) {
  using namespace tdl::dwt;
  TDL_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  TDL_REQUIRE((stride == 1 || stride == 2) && (dil == 1 || dil == 2) && !(stride == 2 && dil == 2),
              "TMA depthwise supports stride 1/2 with dilation 1, or stride 1 with dilation 2");
  if ((int64_t)N * P * Q == 0) return 0;
  TShape s;
  s.N = N; s.H = H; s.W = W; s.C = C; s.P = P; s.Q = Q; s.stride = stride; s.dil = dil; s.pad_h = pad_h; s.pad_w = pad_w;
  s.flip = flip_taps ? 1 : 0;
  choose_tile(P, Q, stride, &s.TH, &s.TW);
  s.BW = (s.TW - 1) * stride + 2 * dil + 1;
  s.BH = (s.TH - 1) * stride + 2 * dil + 1;
  s.tiles_h = (int)tdl::cdiv(P, s.TH);
  s.tiles_w = (int)tdl::cdiv(Q, s.TW);
  CUtensorMap tm, tmm;
  int rc = encode_nhwc(&tm, x, N, H, W, C, s.BW, s.BH);
  if (rc) return rc;
  const int has_mask = mask != nullptr;
  rc = encode_nhwc(&tmm, has_mask ? mask : x, N, has_mask ? P : H, has_mask ? Q : W, C, s.TW, s.TH);
  if (rc) return rc;
  const int x_bytes = ((s.BH * s.BW * 128 + 1023) / 1024) * 1024;
  const int m_bytes = has_mask ? ((s.TH * s.TW * 128 + 1023) / 1024) * 1024 : 0;
  int stages = (150 * 1024) / (x_bytes + m_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) stages = 2;
  const int smem = stages * (x_bytes + m_bytes) + 1024;
  const int tiles = N * s.tiles_h * s.tiles_w;
  const int gy = (int)tdl::cdiv(C, 64);
  int gx = (tdl::num_sms() * 2) / gy;
  if (gx < 1) gx = 1;
  if (gx > tiles) gx = tiles;
  dim3 grid(gx, gy);
  if (nparts) *nparts = gx;
  cudaStream_t st = (cudaStream_t)stream;
#define TDL_DW_LAUNCH(S, D)                                                                                          \
  do {                                                                                                               \
    static bool attr = false;                                                                                        \
    if (!attr) {                                                                                                     \
      TDL_CUDA_OK(cudaFuncSetAttribute(depthwise3x3_tma_kernel<S, D>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                       200 * 1024));                                                                 \
      attr = true;                                                                                                   \
    }                                                                                                                \
    cudaError_t _le = tdl::launch_pdl(depthwise3x3_tma_kernel<S, D>, dim3(grid), dim3(kT), smem, st, tm, tmm, w, bias,     \
                                      has_mask, (tdl::bf16*)y, s, stages, relu_in, relu_out, stats,                  \
                                      (const tdl::bf16*)addend);                                                     \
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "depthwise3x3_tma_kernel");                                   \
  } while (0)
  if (stride == 1 && dil == 1) TDL_DW_LAUNCH(1, 1);
  else if (stride == 1) TDL_DW_LAUNCH(1, 2);
  else TDL_DW_LAUNCH(2, 1);
#undef TDL_DW_LAUNCH
  TDL_LAUNCH_OK("depthwise3x3_tma_kernel");
  return 0;
}

// Same contract as tdl_depthwise3x3_fwd; requires C % 8 == 0, stride in {1,2}, dilation in {1,2}.
TDL_API int tdl_depthwise3x3_tma_fwd(const void* x, const float* w, const float* bias, const void* mask, void* y,
                                     int N, int H, int W, int C, int P, int Q, int stride, int dil, int pad_h,
                                     int pad_w, int relu_in, int relu_out, void* stream) {
  return depthwise3x3_tma_fwd_impl(x, w, bias, mask, y, N, H, W, C, P, Q, stride, dil, pad_h, pad_w, relu_in, relu_out,
                                   nullptr, nullptr, stream);
}

// Data gradient of a stride-1 depthwise layer = the forward kernel on dy with the taps read in reverse order
// (`w` is the layer's own [9][C] weight tensor: no flipped copy), optionally masked by the input's ReLU.
// `addend` (optional, same shape as dx): added AFTER the mask -- the gradient that reaches the layer's input through
// another consumer (an Xception module's skip connection), so no separate elementwise add runs.
TDL_API int tdl_depthwise3x3_tma_dgrad(const void* dy, const float* w, const void* mask, const void* addend, void* dx,
                                       int N, int H, int W, int C, int dil, int pad_h, int pad_w, void* stream) {
  TDL_REQUIRE(addend == nullptr || (reinterpret_cast<uintptr_t>(addend) & 7) == 0, "addend must be 8-byte aligned");
  return depthwise3x3_tma_fwd_impl(dy, w, nullptr, mask, dx, N, H, W, C, H, W, 1, dil, pad_h, pad_w, 0, 0, nullptr,
                                   nullptr, stream, 1, addend);
}

// Forward + per-CTA partial BatchNorm statistics of the bf16 output (for the BatchNorm that follows the depthwise conv in
// the Xception separable convolution): stats_partial is [2 * tdl_num_sms()][2][C] fp32, *nparts_host receives the number
// of valid leading rows; feed both to tdl_bn_finalize_partials.
TDL_API int tdl_depthwise3x3_tma_fwd_stats(const void* x, const float* w, const float* bias, void* y, int N, int H,
                                           int W, int C, int P, int Q, int stride, int dil, int pad_h, int pad_w,
                                           int relu_in, int relu_out, float* stats_partial, int* nparts_host,
                                           void* stream) {
  TDL_REQUIRE(stats_partial != nullptr && nparts_host != nullptr, "stats buffers required");
  return depthwise3x3_tma_fwd_impl(x, w, bias, nullptr, y, N, H, W, C, P, Q, stride, dil, pad_h, pad_w, relu_in,
                                   relu_out, stats_partial, nparts_host, stream);
}
