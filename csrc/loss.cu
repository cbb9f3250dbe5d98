// Losses and metrics on the GPU.
//
//   * Lovasz hinge, per image (reference: core/losses.py:5-92).  The reference pins this loss to
//     the CPU (model.py:391-394) and sorts with tf.nn.top_k; here one CTA owns one image: the
//     errors are sorted in shared memory (bitonic, ties broken by ascending index exactly like
//     top_k), scanned, turned into the Jaccard gradient and dotted with relu(errors) -- forward
//     value and d(loss)/d(logits) leave the kernel together, nothing crosses PCIe.
//   * thresholded mean-IoU and pixel accuracy per image (reference: core/metric.py:6-71),
//     with the reference's extra `score *` factor behind `reference_compat`.
//   * softmax cross-entropy fwd+bwd and in-top-k (not in the reference; BASELINE.json configs).
#include "tdl_common.cuh"

namespace tdl {
namespace loss {

// ------------------------------------------------------------------------------------------
// Lovasz hinge
// ------------------------------------------------------------------------------------------
constexpr int kLovaszThreads = 1024;
constexpr int kLovaszMaxP = 16384;

__device__ __forceinline__ bool lv_before(float ka, uint32_t va, float kb, uint32_t vb) {
  // descending error; equal errors keep ascending original index (tf.nn.top_k order)
  return ka > kb || (ka == kb && (va & 0x7fffffffu) < (vb & 0x7fffffffu));
}

// Sort key: one 64-bit integer per pixel whose ASCENDING order is "descending error, ties by ascending index"
// (tf.nn.top_k order): high word = ~orderable(error), low word = index * 2 + label.  A compare-exchange is then one
// 64-bit compare and two selects instead of a four-way float/index comparison.
__device__ __forceinline__ uint32_t lv_orderable(float f) {
  const uint32_t u = __float_as_uint(f);
  return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}
__device__ __forceinline__ float lv_from_orderable(uint32_t o) {
  return __uint_as_float(o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu));
}

// Each thread owns IPT consecutive positions in registers.  Network stages whose partner distance stays inside a
// thread run on registers, inside a warp on shuffles, and only the 15 (of 105 at 16,384 elements) longest ones go
// through shared memory (transposed [IPT][threads] so every access is conflict-free).  The first version kept
// everything in shared memory with a block barrier per stage: 351 us for 32 images of 101x101.
template <int IPT>
__global__ void __launch_bounds__(kLovaszThreads)
lovasz_hinge_kernel(const float* __restrict__ logits, const float* __restrict__ labels, int P, float grad_scale,
                    float* __restrict__ loss_out, float* __restrict__ dlogits) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  extern __shared__ uint8_t lv_smem[];
  constexpr int T = kLovaszThreads;
  unsigned long long* skey = reinterpret_cast<unsigned long long*>(lv_smem);  // [IPT][T]
  constexpr int NP = IPT * T;
  __shared__ float red[32];
  __shared__ int redi[32];
  const int img = blockIdx.x;
  const float* lg = logits + (size_t)img * P;
  const float* lb = labels + (size_t)img * P;
  const int tid = threadIdx.x;
  const int lane = tid & 31;

  unsigned long long key[IPT];
  int my_pos = 0;
#pragma unroll
  for (int e = 0; e < IPT; ++e) {
    const int i = tid * IPT + e;
    if (i < P) {
      const int y = lb[i] > 0.5f ? 1 : 0;
      const float er = 1.f - lg[i] * (y ? 1.f : -1.f);
      key[e] = ((unsigned long long)(~lv_orderable(er)) << 32) | (unsigned long long)((uint32_t)i * 2u + (uint32_t)y);
      my_pos += y;
    } else {
      key[e] = ~0ull;  // pads sort last
    }
  }
  {
    int v = my_pos;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) redi[tid >> 5] = v;
  }
  __syncthreads();
  int gts = 0;
  for (int w = 0; w < T / 32; ++w) gts += redi[w];
  __syncthreads();

  for (int k = 2; k <= NP; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= 32 * IPT) {
        // partner in another warp: exchange through shared memory
        const int tp = tid ^ (j / IPT);
#pragma unroll
        for (int e = 0; e < IPT; ++e) skey[e * T + tid] = key[e];
        __syncthreads();
        const bool lower = (tid & (j / IPT)) == 0;
#pragma unroll
        for (int e = 0; e < IPT; ++e) {
          const unsigned long long ko = skey[e * T + tp];
          const bool up = ((tid * IPT + e) & k) == 0;
          if ((ko < key[e]) == (lower == up)) key[e] = ko;  // the lower slot of an ascending pair keeps the minimum
        }
        __syncthreads();
      } else if (j >= IPT) {
        const int m = j / IPT;  // partner lane distance
        const bool lower = (lane & m) == 0;
#pragma unroll
        for (int e = 0; e < IPT; ++e) {
          const unsigned long long ko = __shfl_xor_sync(0xffffffffu, key[e], m);
          const bool up = ((tid * IPT + e) & k) == 0;
          if ((ko < key[e]) == (lower == up)) key[e] = ko;
        }
      } else {
        // both elements in this thread's registers (j < IPT): unrolled over every possible j so the indices are static
#pragma unroll
        for (int jj = IPT / 2; jj > 0; jj >>= 1) {
          if (jj != j) continue;
#pragma unroll
          for (int e = 0; e < IPT; ++e) {
            if (e & jj) continue;
            const int o = e | jj;
            const bool up = ((tid * IPT + e) & k) == 0;
            const unsigned long long a = key[e], b2 = key[o];
            const bool swap = (b2 < a) == up;
            key[e] = swap ? b2 : a;
            key[o] = swap ? a : b2;
          }
        }
      }
    }
  }

  float kerr[IPT];
  uint32_t val[IPT];  // as before the rewrite: index | label << 31
#pragma unroll
  for (int e = 0; e < IPT; ++e) {
    const uint32_t lo = (uint32_t)key[e];
    kerr[e] = lv_from_orderable(~(uint32_t)(key[e] >> 32));
    val[e] = (tid * IPT + e < P) ? ((lo >> 1) | ((lo & 1u) << 31)) : 0x7fffffffu;
  }

  // scan of positives over the sorted order: this thread's chunk is already in its registers
  int local = 0;
#pragma unroll
  for (int e = 0; e < IPT; ++e) local += (int)(val[e] >> 31);
  int incl = local;
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) redi[tid >> 5] = incl;
  __syncthreads();
  int warp_off = 0;
  for (int w = 0; w < (tid >> 5); ++w) warp_off += redi[w];
  int cum = warp_off + incl - local;  // positives strictly before my chunk

  float part = 0.f;
#pragma unroll
  for (int e = 0; e < IPT; ++e) {
    const int i = tid * IPT + e;
    if (i >= P) continue;
    const uint32_t v = val[e];
    const int y = (int)(v >> 31);
    const int idx = (int)(v & 0x7fffffffu);
    const float er = kerr[e];
    float prev = 0.f;
    if (i > 0) {
      const float inter_p = (float)(gts - cum);
      const float union_p = (float)(gts + (i - cum));
      prev = 1.f - inter_p / union_p;
    }
    cum += y;
    const float inter = (float)(gts - cum);
    const float uni = (float)(gts + (i + 1 - cum));
    const float jac = 1.f - inter / uni;
    const float gr = jac - prev;
    const bool on = er > 0.f;
    part += on ? er * gr : 0.f;
    if (dlogits) dlogits[(size_t)img * P + idx] = on ? -(y ? 1.f : -1.f) * gr * grad_scale : 0.f;
  }
  part = warp_sum(part);
  if (lane == 0) red[tid >> 5] = part;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int w = 0; w < T / 32; ++w) s += red[w];
    loss_out[img] = s;
  }
}

template <int IPT>
static int launch_lovasz(const float* logits, const float* labels, int B, int P, float grad_scale, float* loss_out,
                         float* dlogits, cudaStream_t st) {
  const int smem = IPT * kLovaszThreads * 8;
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(lovasz_hinge_kernel<IPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  {
    cudaError_t _le = tdl::launch_pdl(lovasz_hinge_kernel<IPT>, dim3(B), dim3(kLovaszThreads), smem, st, logits,
        labels, P, grad_scale, loss_out, dlogits);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "lovasz_hinge_kernel");
  }
  TDL_LAUNCH_OK("lovasz_hinge_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------
// IoU / accuracy per image
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
seg_metrics_kernel(const float* __restrict__ logits, const float* __restrict__ labels, int P, float logit_thr,
                   int reference_compat, float* __restrict__ iou_out, float* __restrict__ acc_out) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ int red[4][8];
  const int img = blockIdx.x;
  int tp = 0, fp = 0, fn = 0, ok = 0;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    const int pr = logits[(size_t)img * P + i] > logit_thr ? 1 : 0;
    const int y = labels[(size_t)img * P + i] > 0.5f ? 1 : 0;
    tp += pr & y;
    fp += pr & (1 - y);
    fn += (1 - pr) & y;
    ok += pr == y;
  }
  int v[4] = {tp, fp, fn, ok};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int t[4] = {0, 0, 0, 0};
    for (int k = 0; k < 4; ++k)
      for (int w = 0; w < (int)blockDim.x / 32; ++w) t[k] += red[k][w];
    const int den = t[0] + t[1] + t[2];
    const double score = den == 0 ? 1.0 : (double)t[0] / (double)den;
    double m = 0.0;
    for (int k = 0; k < 10; ++k) {
      const double thr = 0.5 + 0.05 * k;
      if (score > thr) m += reference_compat ? score : 1.0;
    }
    iou_out[img] = (float)(m / 10.0);
    acc_out[img] = (float)t[3] / (float)P;
  }
}

// ------------------------------------------------------------------------------------------
// softmax cross-entropy (+ gradient) and in-top-k, one warp per row
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
softmax_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, int rows, int classes,
                  float grad_scale, int topk, float* __restrict__ loss_rows, float* __restrict__ dlogits,
                  float* __restrict__ top1, float* __restrict__ topk_out) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* x = logits + (size_t)row * classes;
  const int label = (int)labels[row];
  float mx = -INFINITY;
  for (int j = lane; j < classes; j += 32) mx = fmaxf(mx, x[j]);
  mx = warp_max(mx);
  const float xl = x[label];
  float se = 0.f;
  int rank = 0;
  for (int j = lane; j < classes; j += 32) {
    const float v = x[j];
    se += __expf(v - mx);
    rank += (v > xl) || (v == xl && j < label);
  }
  se = warp_sum(se);
  for (int o = 16; o > 0; o >>= 1) rank += __shfl_xor_sync(0xffffffffu, rank, o);
  const float lse = mx + __logf(se);
  if (lane == 0) {
    loss_rows[row] = lse - xl;
    if (top1) top1[row] = rank < 1 ? 1.f : 0.f;
    if (topk_out) topk_out[row] = rank < topk ? 1.f : 0.f;
  }
  if (dlogits) {
    float* d = dlogits + (size_t)row * classes;
    for (int j = lane; j < classes; j += 32) {
      const float pj = __expf(x[j] - lse);
      d[j] = (pj - (j == label ? 1.f : 0.f)) * grad_scale;
    }
  }
}

}  // namespace loss
}  // namespace tdl

using namespace tdl;
using namespace tdl::loss;

// per-image Lovasz hinge: loss_out[B]; dlogits[B,P] = d(loss_b)/d(logit) * grad_scale (may be null)
TDL_API int tdl_lovasz_hinge(const float* logits, const float* labels, int B, int P, float grad_scale,
                             float* loss_out, float* dlogits, void* stream) {
  TDL_REQUIRE(P > 0 && P <= kLovaszMaxP, "pixels per image must be in (0, 16384]");
  if (B <= 0) return 0;
  int NP = kLovaszThreads;
  while (NP < P) NP <<= 1;
  cudaStream_t st = (cudaStream_t)stream;
  switch (NP / kLovaszThreads) {
    case 1: return launch_lovasz<1>(logits, labels, B, P, grad_scale, loss_out, dlogits, st);
    case 2: return launch_lovasz<2>(logits, labels, B, P, grad_scale, loss_out, dlogits, st);
    case 4: return launch_lovasz<4>(logits, labels, B, P, grad_scale, loss_out, dlogits, st);
    case 8: return launch_lovasz<8>(logits, labels, B, P, grad_scale, loss_out, dlogits, st);
    default: return launch_lovasz<16>(logits, labels, B, P, grad_scale, loss_out, dlogits, st);
  }
}

TDL_API int tdl_seg_metrics(const float* logits, const float* labels, int B, int P, float logit_threshold,
                            int reference_compat, float* iou_out, float* acc_out, void* stream) {
  if (B <= 0) return 0;
  {
    cudaError_t _le = tdl::launch_pdl(seg_metrics_kernel, dim3(B), dim3(256), 0, (cudaStream_t)stream, logits, labels,
        P, logit_threshold, reference_compat, iou_out, acc_out);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "seg_metrics_kernel");
  }
  TDL_LAUNCH_OK("seg_metrics_kernel");
  return 0;
}

TDL_API int tdl_softmax_ce(const float* logits, const void* labels_i64, int rows, int classes, float grad_scale,
                           int topk, float* loss_rows, float* dlogits, float* top1, float* topk_out, void* stream) {
  if (rows <= 0) return 0;
  const int warps = 8;
  {
    cudaError_t _le = tdl::launch_pdl(softmax_ce_kernel, dim3((rows + warps - 1) / warps), dim3(warps * 32), 0,
        (cudaStream_t)stream, logits, (const long long*)labels_i64, rows, classes, grad_scale, topk, loss_rows,
        dlogits, top1, topk_out);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "softmax_ce_kernel");
  }
  TDL_LAUNCH_OK("softmax_ce_kernel");
  return 0;
}
