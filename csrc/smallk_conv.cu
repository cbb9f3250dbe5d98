// Convolution with a handful of output channels (K <= 4): the decoder's final 3x3 conv to ONE logit channel
// (reference core/resnet.py:487-492).  With N = 1 a 128xN tensor-core tile would be >98 % padding, so
// this is a CUDA-core reduction: the GEMM degenerates to one dot product of length R*S*C per output pixel.
//
//   forward      one warp per output pixel, lanes stride over 8-channel vectors, weights (bf16) in shared memory
//   data grad    one warp per input pixel: dx[c] = sum_taps dy[tap] * w[tap][c]  (stride 1 only)
//   weight grad  a thread owns one 8-channel vector and all taps in registers, walks a slice of the pixels,
//                block-reduces through shared memory, then red.add into the fp32 gradient (bias gradient included)
//
// x, y, dy, dx are NHWC bf16; the weight is the fp32 master [K, R, S, C] (rounded to bf16 on load so the numerics
// equal the tcgen05 path, which multiplies bf16 weights).
#include "tdl_common.cuh"

namespace tdl {
namespace smallk {

constexpr int kMaxK = 4;
constexpr int kMaxTaps = 9;
constexpr int kThreads = 256;

struct Geo {
  int N, H, W, C, K, R, S, dil, pt, pl, P, Q;
};

__device__ __forceinline__ void load_weights_smem(const float* __restrict__ w, __nv_bfloat16* sw, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) sw[i] = __float2bfloat16_rn(w[i]);
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads)
fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
           __nv_bfloat16* __restrict__ y, const Geo g) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __nv_bfloat16* sw = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  const int taps = g.R * g.S;
  load_weights_smem(w, sw, g.K * taps * g.C);
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int64_t npix = (int64_t)g.N * g.P * g.Q;
  const int cv_n = g.C >> 3;
  for (int64_t pix = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); pix < npix;
       pix += (int64_t)gridDim.x * warps_per_block) {
    const int q = (int)(pix % g.Q);
    const int p = (int)((pix / g.Q) % g.P);
    const int n = (int)(pix / ((int64_t)g.Q * g.P));
    float acc[kMaxK];
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) acc[k] = 0.f;
    for (int t = 0; t < taps; ++t) {
      const int hh = p - g.pt + (t / g.S) * g.dil;
      const int ww = q - g.pl + (t % g.S) * g.dil;
      if (hh < 0 || hh >= g.H || ww < 0 || ww >= g.W) continue;
      const __nv_bfloat16* xp = x + (((int64_t)n * g.H + hh) * g.W + ww) * g.C;
      for (int cv = lane; cv < cv_n; cv += 32) {
        float xv[8];
        unpack8(ld8(xp + cv * 8), xv);
#pragma unroll
        for (int k = 0; k < kMaxK; ++k) {
          if (k >= g.K) break;
          float wv[8];
          unpack8(ld8(sw + ((int64_t)k * taps + t) * g.C + cv * 8), wv);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[k] = fmaf(xv[j], wv[j], acc[k]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      if (k >= g.K) break;
      float v = acc[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) y[pix * g.K + k] = __float2bfloat16_rn(v + (bias ? bias[k] : 0.f));
    }
  }
}

__global__ void __launch_bounds__(kThreads)
bwd_data_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ w, __nv_bfloat16* __restrict__ dx,
                const Geo g) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __nv_bfloat16* sw = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  const int taps = g.R * g.S;
  load_weights_smem(w, sw, g.K * taps * g.C);
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int64_t npix = (int64_t)g.N * g.H * g.W;
  const int cv_n = g.C >> 3;
  for (int64_t pix = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); pix < npix;
       pix += (int64_t)gridDim.x * warps_per_block) {
    const int ww = (int)(pix % g.W);
    const int hh = (int)((pix / g.W) % g.H);
    const int n = (int)(pix / ((int64_t)g.W * g.H));
    // the (k, tap) gradients that touch this input pixel; warp-uniform, so the loads broadcast
    float gv[kMaxK * kMaxTaps];
#pragma unroll
    for (int i = 0; i < kMaxK * kMaxTaps; ++i) gv[i] = 0.f;
    bool any = false;
#pragma unroll
    for (int t = 0; t < kMaxTaps; ++t) {
      if (t >= taps) break;
      const int p = hh + g.pt - (t / g.S) * g.dil;
      const int q = ww + g.pl - (t % g.S) * g.dil;
      if (p < 0 || p >= g.P || q < 0 || q >= g.Q) continue;
      const __nv_bfloat16* gp = dy + (((int64_t)n * g.P + p) * g.Q + q) * g.K;
#pragma unroll
      for (int k = 0; k < kMaxK; ++k) {
        if (k >= g.K) break;
        gv[k * kMaxTaps + t] = __bfloat162float(gp[k]);
        any = true;
      }
    }
    for (int cv = lane; cv < cv_n; cv += 32) {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      if (any) {
#pragma unroll
        for (int k = 0; k < kMaxK; ++k) {
          if (k >= g.K) break;
#pragma unroll
          for (int t = 0; t < kMaxTaps; ++t) {
            if (t >= taps) break;
            const float gg = gv[k * kMaxTaps + t];
            if (gg == 0.f) continue;
            float wv[8];
            unpack8(ld8(sw + ((int64_t)k * taps + t) * g.C + cv * 8), wv);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = fmaf(gg, wv[j], acc[j]);
          }
        }
      }
      st8(dx + pix * g.C + cv * 8, pack8(acc));
    }
  }
}

// dw[k, tap, c] += sum_pix dy[pix, k] * x[pix shifted by tap, c];  dbias[k] += sum_pix dy[pix, k]
__global__ void __launch_bounds__(kThreads)
bwd_weight_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw,
                  float* __restrict__ dbias, const Geo g) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sacc = reinterpret_cast<float*>(smem_raw);  // [taps][C]
  const int taps = g.R * g.S;
  const int cv_n = g.C >> 3;                           // <= kThreads (checked on the host)
  const int groups = kThreads / cv_n;                  // pixel sub-groups inside the CTA
  const int cv = threadIdx.x % cv_n;
  const int grp = threadIdx.x / cv_n;
  const bool active = grp < groups;
  const int64_t npix = (int64_t)g.N * g.P * g.Q;
  for (int k = 0; k < g.K; ++k) {
    for (int i = threadIdx.x; i < taps * g.C; i += kThreads) sacc[i] = 0.f;
    __syncthreads();
    float acc[kMaxTaps][8];
#pragma unroll
    for (int t = 0; t < kMaxTaps; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
    float gsum = 0.f;
    if (active) {
      for (int64_t pix = (int64_t)blockIdx.x * groups + grp; pix < npix; pix += (int64_t)gridDim.x * groups) {
        const float gg = __bfloat162float(dy[pix * g.K + k]);
        if (gg == 0.f) continue;
        gsum += gg;
        const int q = (int)(pix % g.Q);
        const int p = (int)((pix / g.Q) % g.P);
        const int n = (int)(pix / ((int64_t)g.Q * g.P));
#pragma unroll
        for (int t = 0; t < kMaxTaps; ++t) {
          if (t >= taps) break;
          const int hh = p - g.pt + (t / g.S) * g.dil;
          const int ww = q - g.pl + (t % g.S) * g.dil;
          if (hh < 0 || hh >= g.H || ww < 0 || ww >= g.W) continue;
          float xv[8];
          unpack8(ld8(x + (((int64_t)n * g.H + hh) * g.W + ww) * g.C + cv * 8), xv);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(gg, xv[j], acc[t][j]);
        }
      }
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t) {
        if (t >= taps) break;
#pragma unroll
        for (int j = 0; j < 8; ++j) atomicAdd(&sacc[t * g.C + cv * 8 + j], acc[t][j]);
      }
      if (dbias != nullptr && cv == 0 && gsum != 0.f) atomicAdd(dbias + k, gsum);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < taps * g.C; i += kThreads) {
      const float v = sacc[i];
      if (v != 0.f) atomicAdd(dw + (int64_t)k * taps * g.C + i, v);
    }
    __syncthreads();
  }
}

static int check_geo(const Geo& g) {
  TDL_REQUIRE(g.K >= 1 && g.K <= kMaxK, "small-K conv handles 1..4 output channels");
  TDL_REQUIRE(g.R * g.S <= kMaxTaps, "small-K conv handles at most 9 taps");
  TDL_REQUIRE(g.C % 8 == 0 && g.C >= 8, "C must be a multiple of 8");
  return 0;
}

}  // namespace smallk
}  // namespace tdl

using namespace tdl;
using namespace tdl::smallk;

static Geo make_geo(int N, int H, int W, int C, int K, int R, int S, int dil, int pad_h, int pad_w, int P, int Q) {
  Geo g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.K = K; g.R = R; g.S = S; g.dil = dil; g.pt = pad_h; g.pl = pad_w; g.P = P; g.Q = Q;
  return g;
}

static int grid_for(int64_t items, int per_block) {
  int64_t b = (items + per_block - 1) / per_block;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

TDL_API int tdl_conv2d_smallk_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C,
                                  int K, int R, int S, int dil, int pad_h, int pad_w, int P, int Q, void* stream) {
  const Geo g = make_geo(N, H, W, C, K, R, S, dil, pad_h, pad_w, P, Q);
  if (check_geo(g)) return 1;
  const size_t smem = (size_t)K * R * S * C * sizeof(__nv_bfloat16);
  TDL_REQUIRE(smem <= 200 * 1024, "weights do not fit in shared memory");
  TDL_CUDA_OK(cudaFuncSetAttribute(fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    cudaError_t _le = tdl::launch_pdl(fwd_kernel, dim3(grid_for((int64_t)N * P * Q, kThreads / 32)), dim3(kThreads),
        smem, (cudaStream_t)stream, (const __nv_bfloat16*)x, w, bias, (__nv_bfloat16*)y, g);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "fwd_kernel");
  }
  TDL_LAUNCH_OK("smallk conv fwd");
  return 0;
}

TDL_API int tdl_conv2d_smallk_bwd_data(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int K,
                                       int R, int S, int dil, int pad_h, int pad_w, int P, int Q, void* stream) {
  const Geo g = make_geo(N, H, W, C, K, R, S, dil, pad_h, pad_w, P, Q);
  if (check_geo(g)) return 1;
  const size_t smem = (size_t)K * R * S * C * sizeof(__nv_bfloat16);
  TDL_REQUIRE(smem <= 200 * 1024, "weights do not fit in shared memory");
  TDL_CUDA_OK(cudaFuncSetAttribute(bwd_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  {
    cudaError_t _le = tdl::launch_pdl(bwd_data_kernel, dim3(grid_for((int64_t)N * H * W, kThreads / 32)),
        dim3(kThreads), smem, (cudaStream_t)stream, (const __nv_bfloat16*)dy, w, (__nv_bfloat16*)dx, g);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bwd_data_kernel");
  }
  TDL_LAUNCH_OK("smallk conv bwd data");
  return 0;
}

// dw (fp32 [K, R, S, C]) and dbias (fp32 [K], may be null) are accumulated into: zero them first if needed.
TDL_API int tdl_conv2d_smallk_bwd_weight(const void* x, const void* dy, float* dw, float* dbias, int N, int H, int W,
                                         int C, int K, int R, int S, int dil, int pad_h, int pad_w, int P, int Q,
                                         void* stream) {
  const Geo g = make_geo(N, H, W, C, K, R, S, dil, pad_h, pad_w, P, Q);
  if (check_geo(g)) return 1;
  TDL_REQUIRE(C / 8 <= kThreads, "small-K weight gradient handles C <= 2048");
  const size_t smem = (size_t)R * S * C * sizeof(float);
  TDL_CUDA_OK(cudaFuncSetAttribute(bwd_weight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int groups = kThreads / (C / 8);
  int blocks = grid_for((int64_t)N * P * Q, groups * 16);
  if (blocks > num_sms() * 2) blocks = num_sms() * 2;
  {
    cudaError_t _le = tdl::launch_pdl(bwd_weight_kernel, dim3(blocks), dim3(kThreads), smem, (cudaStream_t)stream,
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, dw, dbias, g);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bwd_weight_kernel");
  }
  TDL_LAUNCH_OK("smallk conv bwd weight");
  return 0;
}
