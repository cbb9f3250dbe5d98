// Fused BatchNorm (+residual) (+ReLU) for NHWC bf16 activations, forward and backward.
//
// Replaces slim.batch_norm + ReLU + the residual add/ReLU of the reference's units
// (reference: core/resnet.py:71,87,126,148,242,373-387; decay/eps from model.py:16-17).
// Semantics follow TF's fused batch norm: biased variance for normalisation, unbiased
// variance into the moving average, moving = decay*moving + (1-decay)*batch.
//
// Everything is memory bound, so the structure is: one column-reduction pass that reads the
// tensor once (16-byte vector loads, per-CTA partials in a scratch buffer, deterministic
// second-level reduce in fp64), and one elementwise pass.
#include "tdl_common.cuh"

namespace tdl {
namespace bn {

constexpr int kThreads = 256;

struct ReduceShape {
  int tx;       // threads across channel-vectors
  int ty;       // threads across rows
  int gx;       // CTAs across rows
  int gy;       // CTAs across channel-vectors
};

static ReduceShape reduce_shape(int64_t rows, int C, int ctas_per_sm = 4) {
  ReduceShape s;
  const int c8 = C / 8;
  s.tx = 1;
  while (s.tx < 32 && s.tx < c8) s.tx <<= 1;
  s.ty = kThreads / s.tx;
  s.gy = (int)cdiv(c8, s.tx);
  int64_t want = cdiv(rows, (int64_t)s.ty * 8);
  int64_t cap = (int64_t)num_sms() * ctas_per_sm / s.gy;  // one full wave of resident CTAs
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  s.gx = (int)want;
  return s;
}

// The one definition of the normalised+activated value, shared by forward and backward so the
// recomputed ReLU mask in backward matches the forward bit for bit.
__device__ __forceinline__ float bn_affine(float x, float scale, float shift) { return fmaf(x, scale, shift); }

// ---- forward statistics: per-CTA partial (sum, sumsq) -> scratch[gx][2][C] ----
__global__ void __launch_bounds__(kThreads)
bn_partial_stats_kernel(const bf16* __restrict__ x, int64_t rows, int C, int tx, int ty, float* __restrict__ partial) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float sm[];  // [ty][tx][16]
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int c8 = C / 8;
  const int cv = blockIdx.y * tx + cx;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  if (cv < c8) {
    const int64_t step = (int64_t)gridDim.x * ty;
    const bf16* col = x + cv * 8;
    int64_t r = (int64_t)blockIdx.x * ty + ry;
    for (; r + 3 * step < rows; r += 4 * step) {  // four independent 16-byte loads in flight
      const bf16x8 v0 = ld8(col + r * C), v1 = ld8(col + (r + step) * C), v2 = ld8(col + (r + 2 * step) * C),
                   v3 = ld8(col + (r + 3 * step) * C);
      float f[8];
#define TDL_ACC(v)                    \
  unpack8(v, f);                      \
  _Pragma("unroll") for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
      TDL_ACC(v0) TDL_ACC(v1) TDL_ACC(v2) TDL_ACC(v3)
    }
    for (; r < rows; r += step) {
      float f[8];
      const bf16x8 v0 = ld8(col + r * C);
      TDL_ACC(v0)
    }
#undef TDL_ACC
  }
  float* mine = sm + (size_t)(ry * tx + cx) * 16;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
  __syncthreads();
  // tx*16 (column, value) pairs are each reduced over the ty row-threads by one thread
  for (int idx = threadIdx.x; idx < tx * 16; idx += kThreads) {
    const int col = idx >> 4;
    const int val = idx & 15;
    const int cvv = blockIdx.y * tx + col;
    if (cvv >= c8) continue;
    float acc = 0.f;
    for (int t = 0; t < ty; ++t) acc += sm[(size_t)(t * tx + col) * 16 + val];
    const int which = val >> 3;  // 0 = first statistic, 1 = second
    const int ch = cvv * 8 + (val & 7);
    partial[((size_t)blockIdx.x * 2 + which) * C + ch] = acc;
  }
}

// Second-level reduce: kFinCh channels x kFinSlices slices of the partial list per CTA, fp64 accumulation.  The
// partial list is short (<= a few hundred rows) so the kernel is pure latency: many slices and four independent
// loads in flight per thread keep it to a couple of memory round trips.
constexpr int kFinThreads = 512;
constexpr int kFinCh = 8;
constexpr int kFinSlices = kFinThreads / kFinCh;

__device__ __forceinline__ void reduce_partials(const float* __restrict__ partial, int nparts, int C, int c,
                                                double* s_out, double* q_out) {
  __shared__ double sm_s[kFinSlices][kFinCh], sm_q[kFinSlices][kFinCh];
  const int lane = threadIdx.x % kFinCh, slice = threadIdx.x / kFinCh;
  double s = 0.0, q = 0.0;
  if (c < C) {
    int p = slice;
    for (; p + 3 * kFinSlices < nparts; p += 4 * kFinSlices) {
      float a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a[u] = partial[((size_t)(p + u * kFinSlices) * 2) * C + c];
        b[u] = partial[((size_t)(p + u * kFinSlices) * 2 + 1) * C + c];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { s += (double)a[u]; q += (double)b[u]; }
    }
    for (; p < nparts; p += kFinSlices) {
      s += (double)partial[((size_t)p * 2) * C + c];
      q += (double)partial[((size_t)p * 2 + 1) * C + c];
    }
  }
  sm_s[slice][lane] = s;
  sm_q[slice][lane] = q;
  __syncthreads();
  s = 0.0; q = 0.0;
  if (slice == 0)
#pragma unroll
    for (int t = 0; t < kFinSlices; ++t) { s += sm_s[t][lane]; q += sm_q[t][lane]; }
  *s_out = s;
  *q_out = q;
}

__global__ void __launch_bounds__(kFinThreads)
bn_finalize_stats_kernel(const float* __restrict__ partial, int nparts, int C, double inv_n, double unbias,
                         const float* __restrict__ gamma, const float* __restrict__ beta,
                         float* __restrict__ running_mean, float* __restrict__ running_var, float decay, float eps,
                         float* __restrict__ mean_out, float* __restrict__ invstd_out, float* __restrict__ scale_out,
                         float* __restrict__ shift_out) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c = blockIdx.x * kFinCh + (threadIdx.x % kFinCh);
  double s, q;
  reduce_partials(partial, nparts, C, c, &s, &q);
  if (threadIdx.x >= kFinCh || c >= C) return;
  const double m = s * inv_n;
  double var = q * inv_n - m * m;
  if (var < 0.0) var = 0.0;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.f;
  const float b = beta ? beta[c] : 0.f;
  const float scale = g * invstd;
  mean_out[c] = (float)m;
  invstd_out[c] = invstd;
  scale_out[c] = scale;
  shift_out[c] = b - (float)m * scale;
  if (running_mean) running_mean[c] = decay * running_mean[c] + (1.f - decay) * (float)m;
  if (running_var) running_var[c] = decay * running_var[c] + (1.f - decay) * (float)(var * unbias);
}

__global__ void bn_infer_scale_shift_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                            const float* __restrict__ rm, const float* __restrict__ rv, float eps,
                                            float* __restrict__ scale, float* __restrict__ shift, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(rv[c] + eps);
  const float g = gamma ? gamma[c] : 1.f;
  const float s = g * invstd;
  scale[c] = s;
  shift[c] = (beta ? beta[c] : 0.f) - rm[c] * s;
}

// ---- forward apply: y = relu(x*scale + shift (+ residual)) ----
// Every thread owns one 8-channel column for all of its rows: the per-channel constants are loaded once into
// registers, the row loop keeps two independent 16-byte loads in flight.
__global__ void __launch_bounds__(kThreads)
bn_apply_kernel(const bf16* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                const bf16* __restrict__ residual, bf16* __restrict__ y, uint8_t* __restrict__ mask_out, int64_t rows,
                int C, int tx, int ty, int relu) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int cv = blockIdx.y * tx + cx;
  if (cv >= C / 8) return;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = scale[cv * 8 + j]; sh[j] = shift[cv * 8 + j]; }
  const int64_t step = (int64_t)gridDim.x * ty;
  const int64_t col = (int64_t)cv * 8;
  auto one = [&](int64_t r, const bf16x8& xv, const bf16x8* rv) {
    float f[8];
    unpack8(xv, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = bn_affine(f[j], sc[j], sh[j]);
    if (rv) {
      float rr[8];
      unpack8(*rv, rr);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += rr[j];
    }
    if (relu) {
      if (mask_out) {  // one bit per element: the backward pass reads this instead of the whole activation
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) m |= (f[j] > 0.f ? 1u : 0u) << j;
        mask_out[r * (C >> 3) + cv] = (uint8_t)m;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
    }
    st8(y + r * C + col, pack8(f));
  };
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + step < rows; r += 2 * step) {
    const bf16x8 x0 = ld8(x + r * C + col), x1 = ld8(x + (r + step) * C + col);
    if (residual) {
      const bf16x8 r0 = ld8(residual + r * C + col), r1 = ld8(residual + (r + step) * C + col);
      one(r, x0, &r0);
      one(r + step, x1, &r1);
    } else {
      one(r, x0, nullptr);
      one(r + step, x1, nullptr);
    }
  }
  for (; r < rows; r += step) {
    const bf16x8 x0 = ld8(x + r * C + col);
    if (residual) {
      const bf16x8 r0 = ld8(residual + r * C + col);
      one(r, x0, &r0);
    } else {
      one(r, x0, nullptr);
    }
  }
}

// ---- backward reduce: dbeta = sum(dy), dgamma = sum(dy * xhat), dy = dz * relu_mask ----
__global__ void __launch_bounds__(kThreads)
bn_bwd_partial_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ x, const bf16* __restrict__ z,
                      const float* __restrict__ scale, const float* __restrict__ shift,
                      const float* __restrict__ mean, const float* __restrict__ invstd, int64_t rows, int C, int tx,
                      int ty, int relu, int z_bits, float* __restrict__ partial) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ float sm[];
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int c8 = C / 8;
  const int cv = blockIdx.y * tx + cx;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  if (cv < c8) {
    float sc[8], sh[8], mu[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = scale[cv * 8 + j];
      sh[j] = shift[cv * 8 + j];
      mu[j] = mean[cv * 8 + j];
    }
    const int64_t step = (int64_t)gridDim.x * ty;
    const int64_t col = (int64_t)cv * 8;
    // q accumulates dy * (x - mean); the invstd factor is applied once at the end
    auto one = [&](const bf16x8& gv, const bf16x8& xv, const bf16x8* zv, int mb) {
      float g[8], f[8];
      unpack8(gv, g);
      unpack8(xv, f);
      if (relu) {
        if (mb >= 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = ((mb >> j) & 1) ? g[j] : 0.f;
        } else if (zv) {
          float zz[8];
          unpack8(*zv, zz);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = zz[j] > 0.f ? g[j] : 0.f;
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = bn_affine(f[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += g[j];
        q[j] = fmaf(g[j], f[j] - mu[j], q[j]);
      }
    };
    int64_t r = (int64_t)blockIdx.x * ty + ry;
    for (; r + step < rows; r += 2 * step) {
      const bf16x8 g0 = ld8(dz + r * C + col), g1 = ld8(dz + (r + step) * C + col);
      const bf16x8 x0 = ld8(x + r * C + col), x1 = ld8(x + (r + step) * C + col);
      if (relu && z && z_bits) {
        const uint8_t* zb = reinterpret_cast<const uint8_t*>(z);
        const int m0 = zb[r * c8 + cv], m1 = zb[(r + step) * c8 + cv];
        one(g0, x0, nullptr, m0);
        one(g1, x1, nullptr, m1);
      } else if (relu && z) {
        const bf16x8 z0 = ld8(z + r * C + col), z1 = ld8(z + (r + step) * C + col);
        one(g0, x0, &z0, -1);
        one(g1, x1, &z1, -1);
      } else {
        one(g0, x0, nullptr, -1);
        one(g1, x1, nullptr, -1);
      }
    }
    for (; r < rows; r += step) {
      const bf16x8 g0 = ld8(dz + r * C + col), x0 = ld8(x + r * C + col);
      if (relu && z && z_bits) {
        one(g0, x0, nullptr, (int)reinterpret_cast<const uint8_t*>(z)[r * c8 + cv]);
      } else if (relu && z) {
        const bf16x8 z0 = ld8(z + r * C + col);
        one(g0, x0, &z0, -1);
      } else {
        one(g0, x0, nullptr, -1);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] *= invstd[cv * 8 + j];
  }
  float* mine = sm + (size_t)(ry * tx + cx) * 16;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mine[j] = s[j];
    mine[8 + j] = q[j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < tx * 16; idx += kThreads) {
    const int col = idx >> 4;
    const int val = idx & 15;
    const int cvv = blockIdx.y * tx + col;
    if (cvv >= c8) continue;
    float acc = 0.f;
    for (int t = 0; t < ty; ++t) acc += sm[(size_t)(t * tx + col) * 16 + val];
    const int which = val >> 3;
    const int ch = cvv * 8 + (val & 7);
    partial[((size_t)blockIdx.x * 2 + which) * C + ch] = acc;
  }
}

__global__ void __launch_bounds__(kFinThreads)
bn_bwd_finalize_kernel(const float* __restrict__ partial, int nparts, int C, float* __restrict__ dbeta,
                       float* __restrict__ dgamma, float* __restrict__ acc_dbeta, float* __restrict__ acc_dgamma) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int c = blockIdx.x * kFinCh + (threadIdx.x % kFinCh);
  double s, q;
  reduce_partials(partial, nparts, C, c, &s, &q);
  if (threadIdx.x >= kFinCh || c >= C) return;
  dbeta[c] = (float)s;
  dgamma[c] = (float)q;
  if (acc_dbeta) acc_dbeta[c] += (float)s;
  if (acc_dgamma) acc_dgamma[c] += (float)q;
}

// ---- backward apply: dx = gamma*invstd*(dy - dbeta/n - xhat*dgamma/n); dres = dy ----
// Folded per channel into dx = A*dy + B*x + D so the row loop is three FMAs per element.
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ x, const bf16* __restrict__ z,
                    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean,
                    const float* __restrict__ invstd, const float* __restrict__ dbeta,
                    const float* __restrict__ dgamma, float inv_n, bf16* __restrict__ dx, bf16* __restrict__ dres,
                    int64_t rows, int C, int tx, int ty, int relu, int z_bits) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int cv = blockIdx.y * tx + cx;
  if (cv >= C / 8) return;
  float sc[8], sh[8], B[8], D[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cv * 8 + j;
    sc[j] = scale[c];  // == gamma * invstd == A
    sh[j] = shift[c];
    const float k = sc[j] * invstd[c] * dgamma[c] * inv_n;
    B[j] = -k;
    D[j] = k * mean[c] - sc[j] * dbeta[c] * inv_n;
  }
  const int64_t step = (int64_t)gridDim.x * ty;
  const int64_t col = (int64_t)cv * 8;
  auto one = [&](int64_t r, const bf16x8& gv, const bf16x8& xv, const bf16x8* zv, int mb) {
    float g[8], f[8], o[8];
    unpack8(gv, g);
    unpack8(xv, f);
    if (relu) {
      if (mb >= 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = ((mb >> j) & 1) ? g[j] : 0.f;
      } else if (zv) {
        float zz[8];
        unpack8(*zv, zz);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = zz[j] > 0.f ? g[j] : 0.f;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = bn_affine(f[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(sc[j], g[j], fmaf(B[j], f[j], D[j]));
    st8(dx + r * C + col, pack8(o));
    if (dres) st8(dres + r * C + col, pack8(g));
  };
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + step < rows; r += 2 * step) {
    const bf16x8 g0 = ld8(dz + r * C + col), g1 = ld8(dz + (r + step) * C + col);
    const bf16x8 x0 = ld8(x + r * C + col), x1 = ld8(x + (r + step) * C + col);
    if (relu && z && z_bits) {
      const uint8_t* zb = reinterpret_cast<const uint8_t*>(z);
      const int m0 = zb[r * (C >> 3) + cv], m1 = zb[(r + step) * (C >> 3) + cv];
      one(r, g0, x0, nullptr, m0);
      one(r + step, g1, x1, nullptr, m1);
    } else if (relu && z) {
      const bf16x8 z0 = ld8(z + r * C + col), z1 = ld8(z + (r + step) * C + col);
      one(r, g0, x0, &z0, -1);
      one(r + step, g1, x1, &z1, -1);
    } else {
      one(r, g0, x0, nullptr, -1);
      one(r + step, g1, x1, nullptr, -1);
    }
  }
  for (; r < rows; r += step) {
    const bf16x8 g0 = ld8(dz + r * C + col), x0 = ld8(x + r * C + col);
    if (relu && z && z_bits) {
      one(r, g0, x0, nullptr, (int)reinterpret_cast<const uint8_t*>(z)[r * (C >> 3) + cv]);
    } else if (relu && z) {
      const bf16x8 z0 = ld8(z + r * C + col);
      one(r, g0, x0, &z0, -1);
    } else {
      one(r, g0, x0, nullptr, -1);
    }
  }
}

// grid for the "thread owns a channel column" elementwise kernels
static ReduceShape apply_shape(int64_t rows, int C) {
  ReduceShape s = reduce_shape(rows, C);
  int64_t want = cdiv(rows, (int64_t)s.ty * 4);
  int64_t cap = (int64_t)num_sms() * 8 / s.gy;
  if (cap < 1) cap = 1;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  s.gx = (int)want;
  return s;
}

static int ew_blocks(int64_t nvec) {
  int64_t b = cdiv(nvec, kThreads);
  const int64_t cap = (int64_t)num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// Resident CTAs per SM of a reduction kernel (registers + its 16 KB of shared memory decide): the grid is sized to
// exactly one wave of them, a 4-per-SM grid of a 3-per-SM kernel would run a second, one-third-full wave.
constexpr int kMaxReduceCtasPerSm = 8;
template <typename Kern>
static int resident_ctas(Kern kernel, int* cache) {
  if (*cache == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, kThreads * 16 * sizeof(float)) != cudaSuccess)
      n = 0;
    if (n < 1) n = 1;
    if (n > kMaxReduceCtasPerSm) n = kMaxReduceCtasPerSm;
    *cache = n;
  }
  return *cache;
}
static int g_fwd_stats_ctas = 0, g_bwd_partial_ctas = 0;

}  // namespace bn
}  // namespace tdl

using namespace tdl;
using namespace tdl::bn;

// Number of fp32 scratch elements tdl_bn_fwd_stats / tdl_bn_bwd_reduce need for this shape.
TDL_API int64_t tdl_bn_scratch_floats(int64_t rows, int C) {
  if (C <= 0 || C % 8) return -1;
  ReduceShape s = reduce_shape(rows, C, kMaxReduceCtasPerSm);  // upper bound over the reduction kernels
  return (int64_t)s.gx * 2 * C;
}

// Batch statistics of x[rows, C] -> mean, invstd, and the folded scale/shift; updates the moving averages.
TDL_API int tdl_bn_fwd_stats(const void* x, int64_t rows, int C, float* scratch, int64_t scratch_floats,
                             const float* gamma, const float* beta, float* running_mean, float* running_var,
                             float decay, float eps, float* mean, float* invstd, float* scale, float* shift,
                             void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(rows > 0, "empty input");
  ReduceShape s = reduce_shape(rows, C, resident_ctas(bn_partial_stats_kernel, &g_fwd_stats_ctas));
  TDL_REQUIRE(scratch_floats >= (int64_t)s.gx * 2 * C, "scratch too small (see tdl_bn_scratch_floats)");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_partial_stats_kernel, dim3(grid), dim3(kThreads),
        kThreads * 16 * sizeof(float), st, (const bf16*)x, rows, C, s.tx, s.ty, scratch);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_partial_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_partial_stats_kernel");
  const double unbias = rows > 1 ? (double)rows / (double)(rows - 1) : 1.0;
  {
    cudaError_t _le = tdl::launch_pdl(bn_finalize_stats_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        st, scratch, s.gx, C, 1.0 / (double)rows, unbias, gamma, beta, running_mean, running_var, decay, eps, mean,
        invstd, scale, shift);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}

// Finish BatchNorm statistics from partial (sum, sumsq) rows produced elsewhere (tdl_conv2d_fprop_stats_bf16):
// partial is [nparts][2][C] fp32 and `rows` is the number of pixels the sums cover.
TDL_API int tdl_bn_finalize_partials(const float* partial, int nparts, int64_t rows, int C, const float* gamma,
                                     const float* beta, float* running_mean, float* running_var, float decay,
                                     float eps, float* mean, float* invstd, float* scale, float* shift, void* stream) {
  TDL_REQUIRE(C > 0 && nparts > 0 && rows > 0, "bad arguments");
  const double unbias = rows > 1 ? (double)rows / (double)(rows - 1) : 1.0;
  {
    cudaError_t _le = tdl::launch_pdl(bn_finalize_stats_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        (cudaStream_t)stream, partial, nparts, C, 1.0 / (double)rows, unbias, gamma, beta, running_mean, running_var,
        decay, eps, mean, invstd, scale, shift);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}

TDL_API int tdl_bn_infer_scale_shift(const float* gamma, const float* beta, const float* running_mean,
                                     const float* running_var, float eps, float* scale, float* shift, int C,
                                     void* stream) {
  {
    cudaError_t _le = tdl::launch_pdl(bn_infer_scale_shift_kernel, dim3((C + 127) / 128), dim3(128), 0,
        (cudaStream_t)stream, gamma, beta, running_mean, running_var, eps, scale, shift, C);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_infer_scale_shift_kernel");
  }
  TDL_LAUNCH_OK("bn_infer_scale_shift_kernel");
  return 0;
}

// y = relu?(x*scale + shift (+ residual))
// mask_out (may be NULL; only meaningful with relu): [rows][C/8] bytes, bit j of byte (r, c/8) = (pre-ReLU value of
// channel 8*(c/8)+j was > 0).  Pass it as `z` with z_is_bitmask = 1 to the backward entry points.
TDL_API int tdl_bn_apply(const void* x, const float* scale, const float* shift, const void* residual, void* y,
                         void* mask_out, int64_t rows, int C, int relu, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  if (rows == 0) return 0;
  ReduceShape s = apply_shape(rows, C);
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_apply_kernel, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream,
        (const bf16*)x, scale, shift, (const bf16*)residual, (bf16*)y, (uint8_t*)mask_out, rows, C, s.tx, s.ty, relu);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_apply_kernel");
  }
  TDL_LAUNCH_OK("bn_apply_kernel");
  return 0;
}

// dbeta/dgamma for this layer (written), optionally also accumulated into acc_* (the flat gradient slots).
TDL_API int tdl_bn_bwd_reduce(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                              const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                              float* scratch,
                              int64_t scratch_floats, float* dgamma, float* dbeta, float* acc_dgamma,
                              float* acc_dbeta, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  ReduceShape s = reduce_shape(rows, C, resident_ctas(bn_bwd_partial_kernel, &g_bwd_partial_ctas));
  TDL_REQUIRE(scratch_floats >= (int64_t)s.gx * 2 * C, "scratch too small (see tdl_bn_scratch_floats)");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_bwd_partial_kernel, dim3(grid), dim3(kThreads),
        kThreads * 16 * sizeof(float), st, (const bf16*)dz, (const bf16*)x, (const bf16*)z, scale, shift, mean,
        invstd, rows, C, s.tx, s.ty, relu, z_is_bitmask, scratch);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_bwd_partial_kernel");
  }
  TDL_LAUNCH_OK("bn_bwd_partial_kernel");
  {
    cudaError_t _le = tdl::launch_pdl(bn_bwd_finalize_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        st, scratch, s.gx, C, dbeta, dgamma, acc_dbeta, acc_dgamma);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_bwd_finalize_kernel");
  }
  TDL_LAUNCH_OK("bn_bwd_finalize_kernel");
  return 0;
}

// BatchNorm-backward sums that were already reduced per CTA elsewhere (the data-gradient kernel's epilogue,
// tdl_conv2d_dgrad_bnstats_bf16): partial is [nparts][2][C] = (sum dz, sum dz * xhat).
TDL_API int tdl_bn_bwd_finalize_partials(const float* partial, int nparts, int C, float* dgamma, float* dbeta,
                                         float* acc_dgamma, float* acc_dbeta, void* stream) {
  TDL_REQUIRE(C > 0 && nparts > 0, "bad shape");
  {
    cudaError_t _le = tdl::launch_pdl(bn_bwd_finalize_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        (cudaStream_t)stream, partial, nparts, C, dbeta, dgamma, acc_dbeta, acc_dgamma);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_bwd_finalize_kernel");
  }
  TDL_LAUNCH_OK("bn_bwd_finalize_kernel");
  return 0;
}

TDL_API int tdl_bn_bwd_apply(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                             const float* shift, const float* mean, const float* invstd, const float* dgamma, const float* dbeta,
                             void* dx, void* dres, int64_t rows, int C, int relu, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  if (rows == 0) return 0;
  ReduceShape s = apply_shape(rows, C);
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_bwd_apply_kernel, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream,
        (const bf16*)dz, (const bf16*)x, (const bf16*)z, scale, shift, mean, invstd, dbeta, dgamma, 1.f / (float)rows,
        (bf16*)dx, (bf16*)dres, rows, C, s.tx, s.ty, relu, z_is_bitmask);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_bwd_apply_kernel");
  }
  TDL_LAUNCH_OK("bn_bwd_apply_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of tdl_bn_fwd_stats.
//
// Some users of this library cannot hand us scratch memory because they cannot obtain any: processes whose
// whole device-memory budget was reserved up front by their own arena or by a serving runtime (a late
// cudaMalloc either fails or is forbidden by policy), and latency-critical loops in which cudaMalloc/cudaFree
// are unacceptable because they synchronise the whole device.  For them the per-CTA partial sums live in a
// fixed array that is part of the library image itself.
//
//   * Limit: kStaticScratchFloats floats (4 MiB).  Every GPU that loads the library carries this array from
//     module load until the process exits, whether or not the entry point is ever called.
//   * The array's name is only a host-side handle; the device address is looked up with cudaGetSymbolAddress
//     once per device, on that device's first call (with that device current).  That first call therefore does
//     a little set-up work and should be made before any stream capture begins; later calls only launch.
//   * All calls share the one array and it keeps whatever the previous call left behind.  Nothing in it has to
//     start at zero: the first kernel overwrites every partial slot that the second kernel reads, so no clearing
//     pass is needed.  Two calls whose kernels can overlap on the GPU (different streams without an ordering
//     between them, or two host threads) race on the array and produce garbage -- serialise them.
//   * No cudaMalloc, cudaFree or allocating library routine is called on this path.
// ------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace tdl {
namespace bn {

constexpr int64_t kStaticScratchFloats = 1 << 20;
static __device__ __align__(256) float g_bn_static_scratch[kStaticScratchFloats];

constexpr int kMaxDevices = 64;

// Device address of g_bn_static_scratch for the current device (looked up once per device).
static int static_scratch_address(float** out) {
  static std::mutex mu;
  static float* cache[kMaxDevices] = {nullptr};
  int dev = 0;
  TDL_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return tdl::fail(-2, "%s%s", "tdl_bn_fwd_stats_noalloc: device ordinal out of range", "");
  std::lock_guard<std::mutex> lock(mu);
  if (cache[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess) return tdl::cuda_fail(e, "cudaGetSymbolAddress(g_bn_static_scratch)");
    if (p == nullptr) return tdl::fail(-2, "%s%s", "cudaGetSymbolAddress returned a null address", "");
    cache[dev] = static_cast<float*>(p);
  }
  *out = cache[dev];
  return 0;
}

}  // namespace bn
}  // namespace tdl

// Same result as tdl_bn_fwd_stats, but the partial sums go to the library's own fixed array instead of caller
// scratch.  Shapes whose partials need more than 1,048,576 floats are refused.  See the block comment above for
// the sharing / first-call rules.
TDL_API int tdl_bn_fwd_stats_noalloc(const void* x, int64_t rows, int C, const float* gamma, const float* beta,
                                     float* running_mean, float* running_var, float decay, float eps, float* mean,
                                     float* invstd, float* scale, float* shift, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(rows > 0, "empty input");
  ReduceShape s = reduce_shape(rows, C);
  const int64_t need = (int64_t)s.gx * 2 * C;
  if (need > kStaticScratchFloats) {
    snprintf(tdl::err_buf(), 512,
             "tdl_bn_fwd_stats_noalloc: this shape (rows=%lld, C=%d) needs %lld floats of scratch but the built-in "
             "array holds %lld; use tdl_bn_fwd_stats with caller-provided scratch",
             (long long)rows, C, (long long)need, (long long)kStaticScratchFloats);
    return -5;
  }
  float* scratch = nullptr;
  int rc = static_scratch_address(&scratch);
  if (rc != 0) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_partial_stats_kernel, dim3(grid), dim3(kThreads),
        kThreads * 16 * sizeof(float), st, (const bf16*)x, rows, C, s.tx, s.ty, scratch);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_partial_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_partial_stats_kernel");
  const double unbias = rows > 1 ? (double)rows / (double)(rows - 1) : 1.0;
  {
    cudaError_t _le = tdl::launch_pdl(bn_finalize_stats_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        st, scratch, s.gx, C, 1.0 / (double)rows, unbias, gamma, beta, running_mean, running_var, decay, eps, mean,
        invstd, scale, shift);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}
