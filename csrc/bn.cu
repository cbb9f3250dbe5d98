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
                         float* __restrict__ shift_out, const float* __restrict__ mean_offset) {
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
  // mean_offset: a per-channel constant the producer left OUT of the tensor because this normalisation removes it
  // anyway (a folded convolution bias); the moving mean still has to see it
  const float m_true = (float)m + (mean_offset ? mean_offset[c] : 0.f);
  if (running_mean) running_mean[c] = decay * running_mean[c] + (1.f - decay) * m_true;
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
                int C, int tx, int ty, int relu, int64_t ldy) {  // ldy: row pitch of y in elements (C, or a concat buffer's)
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
    st8(y + r * ldy + col, pack8(f));
  };
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  if (residual == nullptr) {
    // Four 16-byte loads in flight per thread, as the residual form has (x and residual of two rows): with two, the
    // 4 resident CTAs keep 32 KB per SM in flight and the kernel sits at 73 % of copy bandwidth waiting on memory
    // (ncu: stall_long_sb at the first use of the load); the residual form measured 93 %.
    for (; r + 3 * step < rows; r += 4 * step) {
      const bf16x8 x0 = ld8(x + r * C + col), x1 = ld8(x + (r + step) * C + col);
      const bf16x8 x2 = ld8(x + (r + 2 * step) * C + col), x3 = ld8(x + (r + 3 * step) * C + col);
      one(r, x0, nullptr);
      one(r + step, x1, nullptr);
      one(r + 2 * step, x2, nullptr);
      one(r + 3 * step, x3, nullptr);
    }
  }
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
                      int ty, int relu, int z_bits, float* __restrict__ partial, int64_t ld_dz) {
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
      const bf16x8 g0 = ld8(dz + r * ld_dz + col), g1 = ld8(dz + (r + step) * ld_dz + col);
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
      const bf16x8 g0 = ld8(dz + r * ld_dz + col), x0 = ld8(x + r * C + col);
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
                    int64_t rows, int C, int tx, int ty, int relu, int z_bits, int64_t ld_dz,
                    const bf16* __restrict__ addend, int mask_by_x) {
  // addend (may be null): a second gradient of the SAME tensor x (the identity shortcut of a residual unit takes an alias
  // of the BatchNorm's input), added here instead of by a separate elementwise kernel.  mask_by_x: x is itself the
  // output of a ReLU (x >= 0); its backward -- zero where x == 0 -- is applied to the total, so the producer of x does
  // not need a pass of its own for it.
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
    if (addend) {
      float a[8];
      unpack8(ld8(addend + r * C + col), a);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] += a[j];
    }
    if (mask_by_x) {
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = f[j] > 0.f ? o[j] : 0.f;
    }
    st8(dx + r * C + col, pack8(o));
    if (dres) st8(dres + r * C + col, pack8(g));
  };
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + step < rows; r += 2 * step) {
    const bf16x8 g0 = ld8(dz + r * ld_dz + col), g1 = ld8(dz + (r + step) * ld_dz + col);
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
    const bf16x8 g0 = ld8(dz + r * ld_dz + col), x0 = ld8(x + r * C + col);
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
// per_sm = CTAs of the kernel that fit on one SM: the grid is capped at ONE wave of them (threads loop over rows), a
// grid of 8 per SM for a 4-per-SM kernel ran a second, partly filled wave of tiny CTAs on the small layers
static ReduceShape apply_shape(int64_t rows, int C, int per_sm = 4) {
  ReduceShape s = reduce_shape(rows, C);
  int64_t want = cdiv(rows, (int64_t)s.ty * 4);
  int64_t cap = (int64_t)num_sms() * per_sm / s.gy;
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
        invstd, scale, shift, (const float*)nullptr);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}

// Finish BatchNorm statistics from partial (sum, sumsq) rows produced elsewhere (tdl_conv2d_fprop_stats_bf16):
// partial is [nparts][2][C] fp32 and `rows` is the number of pixels the sums cover.
TDL_API int tdl_bn_finalize_partials_off(const float* partial, int nparts, int64_t rows, int C, const float* gamma,
                                         const float* beta, float* running_mean, float* running_var, float decay,
                                         float eps, float* mean, float* invstd, float* scale, float* shift,
                                         const float* mean_offset, void* stream) {
  TDL_REQUIRE(C > 0 && nparts > 0 && rows > 0, "bad arguments");
  const double unbias = rows > 1 ? (double)rows / (double)(rows - 1) : 1.0;
  {
    cudaError_t _le = tdl::launch_pdl(bn_finalize_stats_kernel, dim3((C + kFinCh - 1) / kFinCh), dim3(kFinThreads), 0,
        (cudaStream_t)stream, partial, nparts, C, 1.0 / (double)rows, unbias, gamma, beta, running_mean, running_var,
        decay, eps, mean, invstd, scale, shift, mean_offset);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}
TDL_API int tdl_bn_finalize_partials(const float* partial, int nparts, int64_t rows, int C, const float* gamma,
                                     const float* beta, float* running_mean, float* running_var, float decay,
                                     float eps, float* mean, float* invstd, float* scale, float* shift, void* stream) {
  return tdl_bn_finalize_partials_off(partial, nparts, rows, C, gamma, beta, running_mean, running_var, decay, eps,
                                      mean, invstd, scale, shift, nullptr, stream);
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
// ldy = elements between consecutive rows of y (>= C, multiple of 8): y may be a channel slice of a wider buffer, which
// is how concatenations disappear (every ASPP branch normalises straight into its slice).
TDL_API int tdl_bn_apply_ld(const void* x, const float* scale, const float* shift, const void* residual, void* y,
                            void* mask_out, int64_t rows, int C, int relu, int64_t ldy, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(ldy >= C && ldy % 8 == 0, "ldy must be a multiple of 8, at least C");
  if (rows == 0) return 0;
  ReduceShape s = apply_shape(rows, C);
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_apply_kernel, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream,
        (const bf16*)x, scale, shift, (const bf16*)residual, (bf16*)y, (uint8_t*)mask_out, rows, C, s.tx, s.ty, relu,
        ldy);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_apply_kernel");
  }
  TDL_LAUNCH_OK("bn_apply_kernel");
  return 0;
}
TDL_API int tdl_bn_apply(const void* x, const float* scale, const float* shift, const void* residual, void* y,
                         void* mask_out, int64_t rows, int C, int relu, void* stream) {
  return tdl_bn_apply_ld(x, scale, shift, residual, y, mask_out, rows, C, relu, C, stream);
}

// dbeta/dgamma for this layer (written), optionally also accumulated into acc_* (the flat gradient slots).
// ld_dz = elements between consecutive rows of dz (>= C, multiple of 8): dz may be a channel slice of the gradient of a
// concat buffer.
TDL_API int tdl_bn_bwd_reduce_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                                 const float* shift, const float* mean, const float* invstd, int64_t rows, int C,
                                 int relu, int64_t ld_dz, float* scratch, int64_t scratch_floats, float* dgamma,
                                 float* dbeta, float* acc_dgamma, float* acc_dbeta, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(ld_dz >= C && ld_dz % 8 == 0, "ld_dz must be a multiple of 8, at least C");
  ReduceShape s = reduce_shape(rows, C, resident_ctas(bn_bwd_partial_kernel, &g_bwd_partial_ctas));
  TDL_REQUIRE(scratch_floats >= (int64_t)s.gx * 2 * C, "scratch too small (see tdl_bn_scratch_floats)");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(s.gx, s.gy);
  {
    cudaError_t _le = tdl::launch_pdl(bn_bwd_partial_kernel, dim3(grid), dim3(kThreads),
        kThreads * 16 * sizeof(float), st, (const bf16*)dz, (const bf16*)x, (const bf16*)z, scale, shift, mean,
        invstd, rows, C, s.tx, s.ty, relu, z_is_bitmask, scratch, ld_dz);
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

TDL_API int tdl_bn_bwd_reduce(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                              const float* shift, const float* mean, const float* invstd, int64_t rows, int C, int relu,
                              float* scratch,
                              int64_t scratch_floats, float* dgamma, float* dbeta, float* acc_dgamma,
                              float* acc_dbeta, void* stream) {
  return tdl_bn_bwd_reduce_ld(dz, x, z, z_is_bitmask, scale, shift, mean, invstd, rows, C, relu, C, scratch,
                              scratch_floats, dgamma, dbeta, acc_dgamma, acc_dbeta, stream);
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

// ------------------------------------------------------------------------------------------------------------
// Finalize folded into the apply kernels (thread-block clusters + distributed shared memory).
//
// A BatchNorm layer used to cost a statistics pass (or the producing convolution's epilogue), a tiny "finalize" launch
// that folds the per-CTA partial rows into per-channel constants, and the apply pass -- and the same again in
// backward.  The ~106 finalize launches of a ResNet-50 step are pure latency (about half a millisecond per step).
// Here the apply kernels do the fold themselves: the grid is launched as clusters of up to 8 CTAs along the row
// dimension, the first CTA of every cluster reduces the partial rows of its channel group in fp64 (fixed order, so
// the result is the same in every cluster and from run to run), leaves the per-channel constants in its shared
// memory, and the other CTAs of the cluster read them through DSMEM after one cluster barrier.  The redundancy is
// grid/8 instead of grid, which keeps the extra L2 reads far below the tensor traffic.  The first cluster also writes
// the per-channel outputs (mean / invstd / scale / shift, moving averages; dgamma / dbeta) exactly once.
// ------------------------------------------------------------------------------------------------------------
namespace tdl {
namespace bn {

constexpr int kMaxCluster = 8;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 4 consecutive floats from the shared memory of CTA `rank` of this cluster (the address is this CTA's own address of
// the same variable)
__device__ __forceinline__ float4 ld_dsmem_f32x4(const float* local_ptr, uint32_t rank) {
  uint32_t ra;
  float4 v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(ptx::smem_u32(local_ptr)), "r"(rank));
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(ra)
               : "memory");
  return v;
}

// fp64 column sums of partial[nparts][2][C] over all rows for channels [c_lo, c_lo + nch); nch divides kThreads.
// Result valid in threads < nch (thread c owns channel c_lo + c).
__device__ __forceinline__ void fold_partial_rows(const float* __restrict__ partial, int nparts, int C, int c_lo,
                                                  int nch, double* red, double* s_out, double* q_out) {
  const int tid = threadIdx.x;
  const int slices = kThreads / nch;
  const int c = tid % nch, sl = tid / nch;
  const int ch = c_lo + c;
  double s = 0.0, q = 0.0;
  if (ch < C) {
    int p = sl;
    for (; p + 3 * slices < nparts; p += 4 * slices) {  // four independent row pairs in flight
      float a[4], b[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a[u] = partial[((size_t)(p + u * slices) * 2) * C + ch];
        b[u] = partial[((size_t)(p + u * slices) * 2 + 1) * C + ch];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { s += (double)a[u]; q += (double)b[u]; }
    }
    for (; p < nparts; p += slices) {
      s += (double)partial[((size_t)p * 2) * C + ch];
      q += (double)partial[((size_t)p * 2 + 1) * C + ch];
    }
  }
  red[tid] = s;
  red[kThreads + tid] = q;
  __syncthreads();
  if (sl == 0) {
    for (int t = 1; t < slices; ++t) {
      s += red[t * nch + c];
      q += red[kThreads + t * nch + c];
    }
  }
  *s_out = s;
  *q_out = q;
}

// ---- forward: finalize + y = relu?(x*scale + shift (+ residual)) ----
template <bool kRes, bool kRelu, bool kMask>
__global__ void __launch_bounds__(kThreads)
bn_apply_fin_kernel(const bf16* __restrict__ x, const float* __restrict__ partial, int nparts, double inv_n,
                    double unbias, const float* __restrict__ gamma, const float* __restrict__ beta,
                    float* __restrict__ running_mean, float* __restrict__ running_var, float decay, float eps,
                    float* __restrict__ mean_out, float* __restrict__ invstd_out, float* __restrict__ scale_out,
                    float* __restrict__ shift_out, const bf16* __restrict__ residual, bf16* __restrict__ y,
                    uint8_t* __restrict__ mask_out, int64_t rows, int C, int tx, int ty) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ double red[2 * kThreads];
  __shared__ __align__(16) float s_sc[kThreads], s_sh[kThreads];
  const int nch = tx * 8;
  const int c_lo = blockIdx.y * nch;
  if (cluster_ctarank() == 0) {
    double s, q;
    fold_partial_rows(partial, nparts, C, c_lo, nch, red, &s, &q);
    if (threadIdx.x < nch) {
      const int ch = c_lo + threadIdx.x;
      float scale = 0.f, shift = 0.f;
      if (ch < C) {
        const double m = s * inv_n;
        double var = q * inv_n - m * m;
        if (var < 0.0) var = 0.0;
        const float invstd = (float)(1.0 / sqrt(var + (double)eps));
        const float g = gamma ? gamma[ch] : 1.f;
        const float b = beta ? beta[ch] : 0.f;
        scale = g * invstd;
        shift = b - (float)m * scale;
        if (blockIdx.x == 0) {  // the per-channel outputs are written once, by the first cluster
          mean_out[ch] = (float)m;
          invstd_out[ch] = invstd;
          scale_out[ch] = scale;
          shift_out[ch] = shift;
          if (running_mean) running_mean[ch] = decay * running_mean[ch] + (1.f - decay) * (float)m;
          if (running_var) running_var[ch] = decay * running_var[ch] + (1.f - decay) * (float)(var * unbias);
        }
      }
      s_sc[threadIdx.x] = scale;
      s_sh[threadIdx.x] = shift;
    }
  }
  cluster_sync_all();
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int cv = blockIdx.y * tx + cx;
  float sc[8], sh[8];
  {
    const float4 a0 = ld_dsmem_f32x4(s_sc + cx * 8, 0), a1 = ld_dsmem_f32x4(s_sc + cx * 8 + 4, 0);
    const float4 b0 = ld_dsmem_f32x4(s_sh + cx * 8, 0), b1 = ld_dsmem_f32x4(s_sh + cx * 8 + 4, 0);
    sc[0] = a0.x; sc[1] = a0.y; sc[2] = a0.z; sc[3] = a0.w; sc[4] = a1.x; sc[5] = a1.y; sc[6] = a1.z; sc[7] = a1.w;
    sh[0] = b0.x; sh[1] = b0.y; sh[2] = b0.z; sh[3] = b0.w; sh[4] = b1.x; sh[5] = b1.y; sh[6] = b1.z; sh[7] = b1.w;
  }
  cluster_sync_all();  // the first CTA's shared memory must outlive every read of it
  if (cv >= C / 8) return;
  const int64_t step = (int64_t)gridDim.x * ty;
  const int64_t col = (int64_t)cv * 8;
  const int c8 = C >> 3;
  auto one = [&](int64_t r, const bf16x8& xv, const bf16x8& rv) {
    float f[8];
    unpack8(xv, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = bn_affine(f[j], sc[j], sh[j]);
    if constexpr (kRes) {
      float rr[8];
      unpack8(rv, rr);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] += rr[j];
    }
    if constexpr (kRelu) {
      if constexpr (kMask) {
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) m |= (f[j] > 0.f ? 1u : 0u) << j;
        mask_out[r * c8 + cv] = (uint8_t)m;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
    }
    st8(y + r * C + col, pack8(f));
  };
  constexpr int U = kRes ? 2 : 4;  // rows in flight per thread (16-byte loads: U, or 2U with a residual)
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + (U - 1) * step < rows; r += U * step) {
    bf16x8 xv[U], rv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      xv[u] = ld8(x + (r + u * step) * C + col);
      if constexpr (kRes) rv[u] = ld8(residual + (r + u * step) * C + col);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) one(r + u * step, xv[u], rv[u]);
  }
  for (; r < rows; r += step) {
    bf16x8 xv = ld8(x + r * C + col), rv;
    if constexpr (kRes) rv = ld8(residual + r * C + col);
    one(r, xv, rv);
  }
}

// ReLU mask source of the backward kernels
enum MaskMode { kNoRelu = 0, kRecompute = 1, kBits = 2, kFromZ = 3 };

template <int MODE>
__device__ __forceinline__ void apply_mask(float* g, const float* f, const float* sc, const float* sh, int mb,
                                           const bf16x8& zv) {
  if constexpr (MODE == kBits) {
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = ((mb >> j) & 1) ? g[j] : 0.f;
  } else if constexpr (MODE == kFromZ) {
    float zz[8];
    unpack8(zv, zz);
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = zz[j] > 0.f ? g[j] : 0.f;
  } else if constexpr (MODE == kRecompute) {
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = bn_affine(f[j], sc[j], sh[j]) > 0.f ? g[j] : 0.f;
  }
}

// ---- backward reduce, specialised per mask source, four rows in flight ----
// (four resident CTAs per SM for the bit-mask variant: it fits 64 registers without spilling and gains ~5 % from the
//  extra loads in flight; the recompute variant spills at 64 and measured slower, profiles/bn_fold_experiment.txt)
template <int MODE>
__global__ void __launch_bounds__(kThreads, MODE == 2 ? 4 : 1)
bn_bwd_partial_t_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ x, const void* __restrict__ zraw,
                        const float* __restrict__ scale, const float* __restrict__ shift,
                        const float* __restrict__ mean, const float* __restrict__ invstd, int64_t rows, int C, int tx,
                        int ty, float* __restrict__ partial, int64_t ld_dz) {
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
      mu[j] = mean[cv * 8 + j];
      if constexpr (MODE == kRecompute) {
        sc[j] = scale[cv * 8 + j];
        sh[j] = shift[cv * 8 + j];
      }
    }
    const int64_t step = (int64_t)gridDim.x * ty;
    const int64_t col = (int64_t)cv * 8;
    const uint8_t* zb = reinterpret_cast<const uint8_t*>(zraw);
    const bf16* zt = reinterpret_cast<const bf16*>(zraw);
    auto one = [&](const bf16x8& gv, const bf16x8& xv, int mb, const bf16x8& zv) {
      float g[8], f[8];
      unpack8(gv, g);
      unpack8(xv, f);
      apply_mask<MODE>(g, f, sc, sh, mb, zv);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += g[j];
        q[j] = fmaf(g[j], f[j] - mu[j], q[j]);
      }
    };
    constexpr int U = MODE == kFromZ ? 2 : 4;
    int64_t r = (int64_t)blockIdx.x * ty + ry;
    for (; r + (U - 1) * step < rows; r += U * step) {
      bf16x8 gv[U], xv[U], zv[U];
      int mb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        gv[u] = ld8(dz + (r + u * step) * ld_dz + col);
        xv[u] = ld8(x + (r + u * step) * C + col);
        if constexpr (MODE == kBits) mb[u] = zb[(r + u * step) * c8 + cv];
        if constexpr (MODE == kFromZ) zv[u] = ld8(zt + (r + u * step) * C + col);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) one(gv[u], xv[u], mb[u], zv[u]);
    }
    for (; r < rows; r += step) {
      bf16x8 gv = ld8(dz + r * ld_dz + col), xv = ld8(x + r * C + col), zv;
      int mb = 0;
      if constexpr (MODE == kBits) mb = zb[r * c8 + cv];
      if constexpr (MODE == kFromZ) zv = ld8(zt + r * C + col);
      one(gv, xv, mb, zv);
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

// ---- backward: finalize (dbeta, dgamma) + dx = A*dy + B*x + D; dres = dy ----
template <int MODE, bool kDres>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_fin_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ x, const void* __restrict__ zraw,
                        const float* __restrict__ scale, const float* __restrict__ shift,
                        const float* __restrict__ mean, const float* __restrict__ invstd,
                        const float* __restrict__ partial, int nparts, float inv_n, float* __restrict__ dgamma_out,
                        float* __restrict__ dbeta_out, float* __restrict__ acc_dgamma, float* __restrict__ acc_dbeta,
                        bf16* __restrict__ dx, bf16* __restrict__ dres, int64_t rows, int C, int tx, int ty) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ double red[2 * kThreads];
  __shared__ __align__(16) float s_A[kThreads], s_B[kThreads], s_D[kThreads], s_S[kThreads];
  const int nch = tx * 8;
  const int c_lo = blockIdx.y * nch;
  if (cluster_ctarank() == 0) {
    double s, q;
    fold_partial_rows(partial, nparts, C, c_lo, nch, red, &s, &q);
    if (threadIdx.x < nch) {
      const int ch = c_lo + threadIdx.x;
      float A = 0.f, B = 0.f, D = 0.f, S = 0.f;
      if (ch < C) {
        const float dbeta = (float)s, dgamma = (float)q;
        A = scale[ch];  // == gamma * invstd
        S = shift[ch];
        const float k = A * invstd[ch] * dgamma * inv_n;
        B = -k;
        D = k * mean[ch] - A * dbeta * inv_n;
        if (blockIdx.x == 0) {
          dbeta_out[ch] = dbeta;
          dgamma_out[ch] = dgamma;
          if (acc_dbeta) acc_dbeta[ch] += dbeta;
          if (acc_dgamma) acc_dgamma[ch] += dgamma;
        }
      }
      s_A[threadIdx.x] = A;
      s_B[threadIdx.x] = B;
      s_D[threadIdx.x] = D;
      s_S[threadIdx.x] = S;
    }
  }
  cluster_sync_all();
  const int cx = threadIdx.x % tx;
  const int ry = threadIdx.x / tx;
  const int cv = blockIdx.y * tx + cx;
  float sc[8], sh[8], B[8], D[8];
  {
    const float4 a0 = ld_dsmem_f32x4(s_A + cx * 8, 0), a1 = ld_dsmem_f32x4(s_A + cx * 8 + 4, 0);
    const float4 b0 = ld_dsmem_f32x4(s_B + cx * 8, 0), b1 = ld_dsmem_f32x4(s_B + cx * 8 + 4, 0);
    const float4 d0 = ld_dsmem_f32x4(s_D + cx * 8, 0), d1 = ld_dsmem_f32x4(s_D + cx * 8 + 4, 0);
    sc[0] = a0.x; sc[1] = a0.y; sc[2] = a0.z; sc[3] = a0.w; sc[4] = a1.x; sc[5] = a1.y; sc[6] = a1.z; sc[7] = a1.w;
    B[0] = b0.x; B[1] = b0.y; B[2] = b0.z; B[3] = b0.w; B[4] = b1.x; B[5] = b1.y; B[6] = b1.z; B[7] = b1.w;
    D[0] = d0.x; D[1] = d0.y; D[2] = d0.z; D[3] = d0.w; D[4] = d1.x; D[5] = d1.y; D[6] = d1.z; D[7] = d1.w;
    if constexpr (MODE == kRecompute) {
      const float4 s0 = ld_dsmem_f32x4(s_S + cx * 8, 0), s1 = ld_dsmem_f32x4(s_S + cx * 8 + 4, 0);
      sh[0] = s0.x; sh[1] = s0.y; sh[2] = s0.z; sh[3] = s0.w; sh[4] = s1.x; sh[5] = s1.y; sh[6] = s1.z; sh[7] = s1.w;
    }
  }
  cluster_sync_all();
  if (cv >= C / 8) return;
  const int64_t step = (int64_t)gridDim.x * ty;
  const int64_t col = (int64_t)cv * 8;
  const int c8 = C >> 3;
  const uint8_t* zb = reinterpret_cast<const uint8_t*>(zraw);
  const bf16* zt = reinterpret_cast<const bf16*>(zraw);
  auto one = [&](int64_t r, const bf16x8& gv, const bf16x8& xv, int mb, const bf16x8& zv) {
    float g[8], f[8], o[8];
    unpack8(gv, g);
    unpack8(xv, f);
    apply_mask<MODE>(g, f, sc, sh, mb, zv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(sc[j], g[j], fmaf(B[j], f[j], D[j]));
    st8(dx + r * C + col, pack8(o));
    if constexpr (kDres) st8(dres + r * C + col, pack8(g));
  };
  constexpr int U = MODE == kFromZ ? 2 : 3;
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + (U - 1) * step < rows; r += U * step) {
    bf16x8 gv[U], xv[U], zv[U];
    int mb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      gv[u] = ld8(dz + (r + u * step) * C + col);
      xv[u] = ld8(x + (r + u * step) * C + col);
      if constexpr (MODE == kBits) mb[u] = zb[(r + u * step) * c8 + cv];
      if constexpr (MODE == kFromZ) zv[u] = ld8(zt + (r + u * step) * C + col);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) one(r + u * step, gv[u], xv[u], mb[u], zv[u]);
  }
  for (; r < rows; r += step) {
    bf16x8 gv = ld8(dz + r * C + col), xv = ld8(x + r * C + col), zv;
    int mb = 0;
    if constexpr (MODE == kBits) mb = zb[r * c8 + cv];
    if constexpr (MODE == kFromZ) zv = ld8(zt + r * C + col);
    one(r, gv, xv, mb, zv);
  }
}

// ---- backward apply, specialised per mask source (the generic kernel above carries every variant's registers) ----
// dx = A*dy + B*x + D (+ addend) (masked by x > 0); dres = dy.  kExtra compiles the addend / mask_by_x handling in.
template <int MODE, bool kDres, bool kExtra>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_t_kernel(const bf16* __restrict__ dz, const bf16* __restrict__ x, const void* __restrict__ zraw,
                      const float* __restrict__ scale, const float* __restrict__ shift,
                      const float* __restrict__ mean, const float* __restrict__ invstd,
                      const float* __restrict__ dbeta, const float* __restrict__ dgamma, float inv_n,
                      bf16* __restrict__ dx, bf16* __restrict__ dres, int64_t rows, int C, int tx, int ty,
                      int64_t ld_dz, const bf16* __restrict__ addend, int mask_by_x) {
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
    sc[j] = scale[c];
    if constexpr (MODE == kRecompute) sh[j] = shift[c];
    const float k = sc[j] * invstd[c] * dgamma[c] * inv_n;
    B[j] = -k;
    D[j] = k * mean[c] - sc[j] * dbeta[c] * inv_n;
  }
  const int64_t step = (int64_t)gridDim.x * ty;
  const int64_t col = (int64_t)cv * 8;
  const int c8 = C >> 3;
  const uint8_t* zb = reinterpret_cast<const uint8_t*>(zraw);
  const bf16* zt = reinterpret_cast<const bf16*>(zraw);
  auto one = [&](int64_t r, const bf16x8& gv, const bf16x8& xv, int mb, const bf16x8& zv, const bf16x8& av) {
    float g[8], f[8], o[8];
    unpack8(gv, g);
    unpack8(xv, f);
    apply_mask<MODE>(g, f, sc, sh, mb, zv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaf(sc[j], g[j], fmaf(B[j], f[j], D[j]));
    if constexpr (kExtra) {
      if (addend) {
        float a[8];
        unpack8(av, a);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += a[j];
      }
      if (mask_by_x) {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = f[j] > 0.f ? o[j] : 0.f;
      }
    }
    st8(dx + r * C + col, pack8(o));
    if constexpr (kDres) st8(dres + r * C + col, pack8(g));
  };
  constexpr int U = 2;
  int64_t r = (int64_t)blockIdx.x * ty + ry;
  for (; r + (U - 1) * step < rows; r += U * step) {
    bf16x8 gv[U], xv[U], zv[U], av[U];
    int mb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      gv[u] = ld8(dz + (r + u * step) * ld_dz + col);
      xv[u] = ld8(x + (r + u * step) * C + col);
      if constexpr (MODE == kBits) mb[u] = zb[(r + u * step) * c8 + cv];
      if constexpr (MODE == kFromZ) zv[u] = ld8(zt + (r + u * step) * C + col);
      if constexpr (kExtra) {
        if (addend) av[u] = ld8(addend + (r + u * step) * C + col);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) one(r + u * step, gv[u], xv[u], mb[u], zv[u], av[u]);
  }
  for (; r < rows; r += step) {
    bf16x8 gv = ld8(dz + r * ld_dz + col), xv = ld8(x + r * C + col), zv, av;
    int mb = 0;
    if constexpr (MODE == kBits) mb = zb[r * c8 + cv];
    if constexpr (MODE == kFromZ) zv = ld8(zt + r * C + col);
    if constexpr (kExtra) {
      if (addend) av = ld8(addend + r * C + col);
    }
    one(r, gv, xv, mb, zv, av);
  }
}

// Launch as clusters of `cl` CTAs along x (cl divides grid.x), with the programmatic-launch attribute when enabled.
template <typename... KArgs, typename... Args>
static cudaError_t launch_clustered(void (*kernel)(KArgs...), dim3 grid, int cl, size_t smem, cudaStream_t st,
                                    Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = tdl::pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// grid of the fused apply kernels: the plain apply shape with gx rounded down to whole clusters
static ReduceShape fused_shape(int64_t rows, int C, int* cluster) {
  ReduceShape s = apply_shape(rows, C);
  static int max_cl = 0;  // TDL_BN_CLUSTER: cluster size of the fused apply kernels (1, 2, 4 or 8; default 8)
  if (max_cl == 0) {
    const char* e = getenv("TDL_BN_CLUSTER");
    max_cl = e ? atoi(e) : kMaxCluster;
    if (max_cl < 1 || max_cl > kMaxCluster || (max_cl & (max_cl - 1))) max_cl = kMaxCluster;
  }
  int cl = max_cl;
  while (cl > s.gx) cl >>= 1;
  if (cl < 1) cl = 1;
  s.gx = s.gx / cl * cl;
  *cluster = cl;
  return s;
}

static int mask_mode(int relu, const void* z, int z_is_bitmask) {
  if (!relu) return kNoRelu;
  if (!z) return kRecompute;
  return z_is_bitmask ? kBits : kFromZ;
}

static int g_bwd_partial_t_ctas[4] = {0, 0, 0, 0};

}  // namespace bn
}  // namespace tdl

// y = relu?(BN(x) (+ residual)) straight from the partial (sum, sum of squares) rows: statistics finalize, moving
// averages and the normalise pass in ONE launch.  partial: [nparts][2][C] fp32 covering `stat_rows` pixels (from
// tdl_conv2d_fprop_stats_bf16 / tdl_bn_partial_stats).  mean / invstd / scale / shift ([C] each) are outputs (the
// backward pass needs them).  mask_out as in tdl_bn_apply.
TDL_API int tdl_bn_apply_fused(const void* x, const float* partial, int nparts, int64_t stat_rows, const float* gamma,
                               const float* beta, float* running_mean, float* running_var, float decay, float eps,
                               float* mean, float* invstd, float* scale, float* shift, const void* residual, void* y,
                               void* mask_out, int64_t rows, int C, int relu, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(nparts > 0 && stat_rows > 0 && rows > 0, "bad arguments");
  TDL_REQUIRE(mean && invstd && scale && shift, "per-channel outputs required");
  int cl = 1;
  ReduceShape s = fused_shape(rows, C, &cl);
  dim3 grid(s.gx, s.gy);
  const double inv_n = 1.0 / (double)stat_rows;
  const double unbias = stat_rows > 1 ? (double)stat_rows / (double)(stat_rows - 1) : 1.0;
  cudaStream_t st = (cudaStream_t)stream;
  const bool res = residual != nullptr, mk = mask_out != nullptr && relu;
  cudaError_t e;
#define TDL_BN_FWD(R_, L_, M_)                                                                                     \
  e = launch_clustered(bn_apply_fin_kernel<R_, L_, M_>, grid, cl, 0, st, (const bf16*)x, partial, nparts, inv_n,   \
                       unbias, gamma, beta, running_mean, running_var, decay, eps, mean, invstd, scale, shift,     \
                       (const bf16*)residual, (bf16*)y, (uint8_t*)mask_out, rows, C, s.tx, s.ty)
  if (res && relu && mk) TDL_BN_FWD(true, true, true);
  else if (res && relu) TDL_BN_FWD(true, true, false);
  else if (res) TDL_BN_FWD(true, false, false);
  else if (relu && mk) TDL_BN_FWD(false, true, true);
  else if (relu) TDL_BN_FWD(false, true, false);
  else TDL_BN_FWD(false, false, false);
#undef TDL_BN_FWD
  if (e != cudaSuccess) return tdl::cuda_fail(e, "bn_apply_fin_kernel");
  TDL_LAUNCH_OK("bn_apply_fin_kernel");
  return 0;
}

// Partial (sum, sum of squares) rows of x[rows, C] into scratch ([*nparts][2][C]); the first half of tdl_bn_fwd_stats.
TDL_API int tdl_bn_partial_stats(const void* x, int64_t rows, int C, float* scratch, int64_t scratch_floats,
                                 int* nparts, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(rows > 0 && nparts, "bad arguments");
  ReduceShape s = reduce_shape(rows, C, resident_ctas(bn_partial_stats_kernel, &g_fwd_stats_ctas));
  TDL_REQUIRE(scratch_floats >= (int64_t)s.gx * 2 * C, "scratch too small (see tdl_bn_scratch_floats)");
  cudaError_t e = tdl::launch_pdl(bn_partial_stats_kernel, dim3(s.gx, s.gy), dim3(kThreads),
                                  kThreads * 16 * sizeof(float), (cudaStream_t)stream, (const bf16*)x, rows, C, s.tx,
                                  s.ty, scratch);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "bn_partial_stats_kernel");
  TDL_LAUNCH_OK("bn_partial_stats_kernel");
  *nparts = s.gx;
  return 0;
}

// Partial rows of (sum dy, sum dy * xhat), dy = dz with the ReLU mask applied: the reduction half of
// tdl_bn_bwd_reduce, specialised per mask source.  Feed scratch / *nparts to tdl_bn_bwd_apply_fused.
TDL_API int tdl_bn_bwd_partial_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                                  const float* shift, const float* mean, const float* invstd, int64_t rows, int C,
                                  int relu, int64_t ld_dz, float* scratch, int64_t scratch_floats, int* nparts,
                                  void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(rows > 0 && nparts, "bad arguments");
  TDL_REQUIRE(ld_dz >= C && ld_dz % 8 == 0, "ld_dz must be a multiple of 8, at least C");
  const int mode = mask_mode(relu, z, z_is_bitmask);
  void (*kern)(const bf16*, const bf16*, const void*, const float*, const float*, const float*, const float*, int64_t,
               int, int, int, float*, int64_t) =
      mode == kNoRelu ? bn_bwd_partial_t_kernel<kNoRelu>
      : mode == kRecompute ? bn_bwd_partial_t_kernel<kRecompute>
      : mode == kBits ? bn_bwd_partial_t_kernel<kBits> : bn_bwd_partial_t_kernel<kFromZ>;
  ReduceShape s = reduce_shape(rows, C, resident_ctas(kern, &g_bwd_partial_t_ctas[mode]));
  TDL_REQUIRE(scratch_floats >= (int64_t)s.gx * 2 * C, "scratch too small (see tdl_bn_scratch_floats)");
  cudaError_t e = tdl::launch_pdl(kern, dim3(s.gx, s.gy), dim3(kThreads), kThreads * 16 * sizeof(float),
                                  (cudaStream_t)stream, (const bf16*)dz, (const bf16*)x, z, scale, shift, mean, invstd,
                                  rows, C, s.tx, s.ty, scratch, ld_dz);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "bn_bwd_partial_t_kernel");
  TDL_LAUNCH_OK("bn_bwd_partial_t_kernel");
  *nparts = s.gx;
  return 0;
}
TDL_API int tdl_bn_bwd_partial(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                               const float* shift, const float* mean, const float* invstd, int64_t rows, int C,
                               int relu, float* scratch, int64_t scratch_floats, int* nparts, void* stream) {
  return tdl_bn_bwd_partial_ld(dz, x, z, z_is_bitmask, scale, shift, mean, invstd, rows, C, relu, C, scratch,
                               scratch_floats, nparts, stream);
}

// dgamma / dbeta from the partial rows AND dx (and dres) in one launch (replaces tdl_bn_bwd_finalize_partials +
// tdl_bn_bwd_apply).  dgamma / dbeta ([C]) are written; acc_* (may be NULL) are added to.
TDL_API int tdl_bn_bwd_apply_fused(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                                   const float* shift, const float* mean, const float* invstd, const float* partial,
                                   int nparts, float* dgamma, float* dbeta, float* acc_dgamma, float* acc_dbeta,
                                   void* dx, void* dres, int64_t rows, int C, int relu, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(nparts > 0 && rows > 0 && dgamma && dbeta, "bad arguments");
  int cl = 1;
  ReduceShape s = fused_shape(rows, C, &cl);
  dim3 grid(s.gx, s.gy);
  cudaStream_t st = (cudaStream_t)stream;
  const int mode = mask_mode(relu, z, z_is_bitmask);
  const float inv_n = 1.f / (float)rows;
  cudaError_t e;
#define TDL_BN_BWD(M_, D_)                                                                                          \
  e = launch_clustered(bn_bwd_apply_fin_kernel<M_, D_>, grid, cl, 0, st, (const bf16*)dz, (const bf16*)x, z, scale, \
                       shift, mean, invstd, partial, nparts, inv_n, dgamma, dbeta, acc_dgamma, acc_dbeta,           \
                       (bf16*)dx, (bf16*)dres, rows, C, s.tx, s.ty)
  if (dres) {
    if (mode == kNoRelu) TDL_BN_BWD(kNoRelu, true);
    else if (mode == kRecompute) TDL_BN_BWD(kRecompute, true);
    else if (mode == kBits) TDL_BN_BWD(kBits, true);
    else TDL_BN_BWD(kFromZ, true);
  } else {
    if (mode == kNoRelu) TDL_BN_BWD(kNoRelu, false);
    else if (mode == kRecompute) TDL_BN_BWD(kRecompute, false);
    else if (mode == kBits) TDL_BN_BWD(kBits, false);
    else TDL_BN_BWD(kFromZ, false);
  }
#undef TDL_BN_BWD
  if (e != cudaSuccess) return tdl::cuda_fail(e, "bn_bwd_apply_fin_kernel");
  TDL_LAUNCH_OK("bn_bwd_apply_fin_kernel");
  return 0;
}

// addend ([rows, C] bf16, may be NULL): added to dx (a second gradient of the BatchNorm's input, e.g. a residual unit's
// identity shortcut).  mask_by_x != 0: the input x is a ReLU output; dx is zeroed where x == 0 (that ReLU's backward).
TDL_API int tdl_bn_bwd_apply_ex(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                                const float* shift, const float* mean, const float* invstd, const float* dgamma,
                                const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu, int64_t ld_dz,
                                const void* addend, int mask_by_x, void* stream) {
  TDL_REQUIRE(C > 0 && C % 8 == 0, "C must be a positive multiple of 8");
  TDL_REQUIRE(ld_dz >= C && ld_dz % 8 == 0, "ld_dz must be a multiple of 8, at least C");
  if (rows == 0) return 0;
  const bool extra = addend != nullptr || mask_by_x != 0;
  ReduceShape s = apply_shape(rows, C, extra ? 2 : 4);  // the addend / mask variants need 68-94 registers
  dim3 grid(s.gx, s.gy);
  const int mode = mask_mode(relu, z, z_is_bitmask);
  const float inv_n = 1.f / (float)rows;
  cudaError_t e;
#define TDL_BN_BWD_T(M_, D_, X_)                                                                                     \
  e = tdl::launch_pdl(bn_bwd_apply_t_kernel<M_, D_, X_>, dim3(grid), dim3(kThreads), 0, (cudaStream_t)stream,        \
                      (const bf16*)dz, (const bf16*)x, z, scale, shift, mean, invstd, dbeta, dgamma, inv_n, (bf16*)dx, \
                      (bf16*)dres, rows, C, s.tx, s.ty, ld_dz, (const bf16*)addend, mask_by_x)
#define TDL_BN_BWD_M(M_)                                  \
  do {                                                    \
    if (dres) {                                           \
      if (extra) TDL_BN_BWD_T(M_, true, true);            \
      else TDL_BN_BWD_T(M_, true, false);                 \
    } else {                                              \
      if (extra) TDL_BN_BWD_T(M_, false, true);           \
      else TDL_BN_BWD_T(M_, false, false);                \
    }                                                     \
  } while (0)
  if (mode == kNoRelu) TDL_BN_BWD_M(kNoRelu);
  else if (mode == kRecompute) TDL_BN_BWD_M(kRecompute);
  else if (mode == kBits) TDL_BN_BWD_M(kBits);
  else TDL_BN_BWD_M(kFromZ);
#undef TDL_BN_BWD_M
#undef TDL_BN_BWD_T
  if (e != cudaSuccess) return tdl::cuda_fail(e, "bn_bwd_apply_t_kernel");
  TDL_LAUNCH_OK("bn_bwd_apply_t_kernel");
  return 0;
}
TDL_API int tdl_bn_bwd_apply_ld(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                                const float* shift, const float* mean, const float* invstd, const float* dgamma,
                                const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu, int64_t ld_dz,
                                void* stream) {
  return tdl_bn_bwd_apply_ex(dz, x, z, z_is_bitmask, scale, shift, mean, invstd, dgamma, dbeta, dx, dres, rows, C, relu,
                             ld_dz, nullptr, 0, stream);
}
TDL_API int tdl_bn_bwd_apply(const void* dz, const void* x, const void* z, int z_is_bitmask, const float* scale,
                             const float* shift, const float* mean, const float* invstd, const float* dgamma, const float* dbeta,
                             void* dx, void* dres, int64_t rows, int C, int relu, void* stream) {
  return tdl_bn_bwd_apply_ld(dz, x, z, z_is_bitmask, scale, shift, mean, invstd, dgamma, dbeta, dx, dres, rows, C, relu,
                             C, stream);
}

// ------------------------------------------------------------------------------------------------------------
// BatchNorm WITHOUT activation in front of a 1x1 convolution, folded into that convolution (Xception's
// depthwise -> BatchNorm -> pointwise, core/xception.py:184-210): conv(scale * y + shift) = conv'(y) + b' with
// W'[k][c] = W[k][c] * scale[c] and b'[k] = b[k] + sum_c W[k][c] * shift[c].  The normalise pass over the activation
// (one read + one write of the whole tensor) is replaced by a pass over the K x C weight matrix.
// In backward the data gradient uses the ORIGINAL weights (it is the gradient w.r.t. the BatchNorm output, which the
// BatchNorm backward kernels take as usual), and the weight gradient computed on the raw input y is corrected:
// dW[k][c] = scale[c] * sum_px dz[px][k] y[px][c] + shift[c] * sum_px dz[px][k].
// ------------------------------------------------------------------------------------------------------------
namespace tdl {
namespace bn {
__global__ void __launch_bounds__(128)
fold_affine_kernel(const float* __restrict__ w, const float* __restrict__ scale, const float* __restrict__ shift,
                   const float* __restrict__ bias_in, bf16* __restrict__ wf, float* __restrict__ bias_out, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int k = blockIdx.x;
  const float* wr = w + (size_t)k * C;
  bf16* out = wf + (size_t)k * C;
  float acc = 0.f;
  for (int c = threadIdx.x; c < C; c += 128) {
    const float v = wr[c];
    out[c] = __float2bfloat16(v * scale[c]);
    acc = fmaf(v, shift[c], acc);
  }
  __shared__ float red[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) bias_out[k] = (bias_in ? bias_in[k] : 0.f) + red[0] + red[1] + red[2] + red[3];
}

__global__ void __launch_bounds__(256)
fold_wgrad_fixup_kernel(float* __restrict__ dw, const float* __restrict__ scale, const float* __restrict__ shift,
                        const float* __restrict__ dzsum, int K, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)K * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int k = (int)(i / C);
    dw[i] = dzsum ? fmaf(scale[c], dw[i], shift[c] * dzsum[k]) : scale[c] * dw[i];
  }
}
}  // namespace bn
}  // namespace tdl

// w [K][C] fp32 master weights of a 1x1 convolution; wf [K][C] bf16 and bias_out [K] fp32 are written.
TDL_API int tdl_conv1x1_fold_affine(const float* w, const float* scale, const float* shift, const float* bias_in,
                                    void* wf, float* bias_out, int K, int C, void* stream) {
  TDL_REQUIRE(K > 0 && C > 0, "empty weight matrix");
  cudaError_t e = tdl::launch_pdl(tdl::bn::fold_affine_kernel, dim3(K), dim3(128), 0, (cudaStream_t)stream, w, scale,
                                  shift, bias_in, (tdl::bf16*)wf, bias_out, C);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "fold_affine_kernel");
  TDL_LAUNCH_OK("fold_affine_kernel");
  return 0;
}

// In place on dw [K][C] (this step's raw weight gradient, nothing else accumulated into it yet); dzsum [K] = column
// sums of the output gradient, or NULL when they are known to be zero (the convolution feeds a training-mode
// BatchNorm: that BatchNorm's input gradient has zero mean per channel).
TDL_API int tdl_conv1x1_fold_wgrad_fixup(float* dw, const float* scale, const float* shift, const float* dzsum, int K,
                                         int C, void* stream) {
  TDL_REQUIRE(K > 0 && C > 0, "empty weight matrix");
  int64_t blocks = tdl::cdiv((int64_t)K * C, 256);
  if (blocks > (int64_t)tdl::num_sms() * 8) blocks = (int64_t)tdl::num_sms() * 8;
  cudaError_t e = tdl::launch_pdl(tdl::bn::fold_wgrad_fixup_kernel, dim3((unsigned)blocks), dim3(256), 0,
                                  (cudaStream_t)stream, dw, scale, shift, dzsum, K, C);
  if (e != cudaSuccess) return tdl::cuda_fail(e, "fold_wgrad_fixup_kernel");
  TDL_LAUNCH_OK("fold_wgrad_fixup_kernel");
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
        invstd, scale, shift, (const float*)nullptr);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "bn_finalize_stats_kernel");
  }
  TDL_LAUNCH_OK("bn_finalize_stats_kernel");
  return 0;
}
