// Training-time augmentation of a whole batch on the GPU (segmentation task).
//
// Replaces the per-sample CPU pipeline of the reference -- reflect pad 40, random transpose, flips, rotation,
// shift as one composed projective transform (tf.contrib.image.transform, BILINEAR for the image, NEAREST for the
// mask), central crop back to HxW, normalisation and the "laplace" second channel
// (reference: preprocessing/preprocessing.py:112-246 and :11-30).  The reference maps this single-threaded on the
// host (model.py:315); eight B200s cannot be fed that way, so every output pixel is one thread here and nothing
// but the raw decoded images crosses PCIe.  The CPU restatement in data/preprocessing.py is the oracle.
#include "tdl_common.cuh"

namespace tdl {
namespace augment {

struct Params {
  int H, W, pad;
  float mean, inv_std;
  int laplace_2d;  // 0 = the reference's 1-D [1,-6,1] along W, 1 = the intended 2-D kernel
  float brightness;  // per-sample additive delta on the normalised image (tf.image.random_brightness), set per pixel
};

// decoded pixel in [0, 1]: fp32 as is, uint8 / 255 with the correctly rounded division the CPU decode uses
template <typename T>
__device__ __forceinline__ float load_px(const T* __restrict__ img, int i);
template <>
__device__ __forceinline__ float load_px<float>(const float* __restrict__ img, int i) { return img[i]; }
template <>
__device__ __forceinline__ float load_px<uint8_t>(const uint8_t* __restrict__ img, int i) {
  return __fdiv_rn((float)img[i], 255.f);
}

__device__ __forceinline__ int reflect(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

// value of the (optionally transposed) reflect-padded, normalised image at padded coordinates (px, py); 0 outside
template <typename T>
__device__ __forceinline__ float padded_at(const T* __restrict__ img, const Params& p, int transposed, int px,
                                           int py, bool normalise) {
  const int Hp = p.H + 2 * p.pad, Wp = p.W + 2 * p.pad;
  if (px < 0 || py < 0 || px >= Wp || py >= Hp) return 0.f;
  if (transposed) { const int t = px; px = py; py = t; }
  const int y = reflect(py - p.pad, p.H), x = reflect(px - p.pad, p.W);
  const float v = load_px<T>(img, y * p.W + x);
  return normalise ? (v - p.mean) * p.inv_std + p.brightness : v;
}

template <typename T>
__device__ __forceinline__ float sample_bilinear(const T* __restrict__ img, const Params& p, const float* t,
                                                 int transposed, int xo, int yo) {
  const float X = (float)(xo + p.pad), Y = (float)(yo + p.pad);
  const float k = t[6] * X + t[7] * Y + 1.f;
  const float sx = (t[0] * X + t[1] * Y + t[2]) / k;
  const float sy = (t[3] * X + t[4] * Y + t[5]) / k;
  const float fx0 = floorf(sx), fy0 = floorf(sy);
  const int x0 = (int)fx0, y0 = (int)fy0;
  const float fx = sx - fx0, fy = sy - fy0;
  const float tl = padded_at(img, p, transposed, x0, y0, true), tr = padded_at(img, p, transposed, x0 + 1, y0, true);
  const float bl = padded_at(img, p, transposed, x0, y0 + 1, true),
              br = padded_at(img, p, transposed, x0 + 1, y0 + 1, true);
  const float top = tl * (1.f - fx) + tr * fx;
  const float bot = bl * (1.f - fx) + br * fx;
  return top * (1.f - fy) + bot * fy;
}

template <typename T>
__device__ __forceinline__ float sample_nearest(const T* __restrict__ img, const Params& p, const float* t,
                                                int transposed, int xo, int yo) {
  const float X = (float)(xo + p.pad), Y = (float)(yo + p.pad);
  const float k = t[6] * X + t[7] * Y + 1.f;
  const float sx = (t[0] * X + t[1] * Y + t[2]) / k;
  const float sy = (t[3] * X + t[4] * Y + t[5]) / k;
  return padded_at(img, p, transposed, (int)rintf(sx), (int)rintf(sy), false);
}

template <typename T>
__global__ void __launch_bounds__(256)
augment_seg_kernel(const T* __restrict__ images, const T* __restrict__ masks,
                   const float* __restrict__ transforms, const int* __restrict__ transposed,
                   const float* __restrict__ brightness, float* __restrict__ out, float* __restrict__ out_mask, int B,
                   Params p) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)B * p.H * p.W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % p.W);
    const int y = (int)((i / p.W) % p.H);
    const int b = (int)(i / ((int64_t)p.W * p.H));
    const T* img = images + (int64_t)b * p.H * p.W;
    p.brightness = brightness ? brightness[b] : 0.f;
    float t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) t[j] = transforms[b * 8 + j];
    const int tr = transposed[b];
    const float c = sample_bilinear(img, p, t, tr, x, y);
    const float l = x > 0 ? sample_bilinear(img, p, t, tr, x - 1, y) : 0.f;
    const float r = x < p.W - 1 ? sample_bilinear(img, p, t, tr, x + 1, y) : 0.f;
    float lap = l - 6.f * c + r;
    if (p.laplace_2d) {
      float acc = -6.f * c + l + r;
      for (int dy = -1; dy <= 1; dy += 2) {
        const int yy = y + dy;
        if (yy < 0 || yy >= p.H) continue;
        acc += sample_bilinear(img, p, t, tr, x, yy);
        if (x > 0) acc += 0.5f * sample_bilinear(img, p, t, tr, x - 1, yy);
        if (x < p.W - 1) acc += 0.5f * sample_bilinear(img, p, t, tr, x + 1, yy);
      }
      lap = acc;
    }
    out[i * 2] = c;
    out[i * 2 + 1] = lap;
    if (masks) out_mask[i] = sample_nearest(masks + (int64_t)b * p.H * p.W, p, t, tr, x, y);
  }
}

}  // namespace augment
}  // namespace tdl

// images/masks: [B, H, W] decoded, un-normalised: fp32 in [0,1], or (input_is_u8) raw uint8 that the kernel divides
// by 255; transforms: [B, 8] output->input projective transforms in the reflect-padded frame; transposed: [B] 0/1;
// brightness: [B] additive delta on the normalised image (tf.image.random_brightness,
// preprocessing/preprocessing.py:169-170) or NULL.  out: [B, H, W, 2] (normalised image, laplace), out_mask:
// [B, H, W, 1].  masks/out_mask may be NULL.
TDL_API int tdl_augment_seg_ex(const void* images, const void* masks, int input_is_u8, const float* transforms,
                               const int* transposed, const float* brightness, float* out, float* out_mask, int B,
                               int H, int W, int pad, float mean, float stddev, int laplace_2d, void* stream) {
  TDL_REQUIRE(H > pad && W > pad, "reflect padding must be smaller than the image");
  TDL_REQUIRE(stddev > 0.f, "stddev must be positive");
  if (B <= 0) return 0;
  tdl::augment::Params p = {H, W, pad, mean, 1.f / stddev, laplace_2d, 0.f};
  const int64_t total = (int64_t)B * H * W;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)tdl::num_sms() * 8;
  if (blocks > cap) blocks = cap;
  cudaError_t le;
  if (input_is_u8)
    le = tdl::launch_pdl(tdl::augment::augment_seg_kernel<uint8_t>, dim3((int)blocks), dim3(256), 0,
                         (cudaStream_t)stream, (const uint8_t*)images, (const uint8_t*)masks, transforms, transposed,
                         brightness, out, out_mask, B, p);
  else
    le = tdl::launch_pdl(tdl::augment::augment_seg_kernel<float>, dim3((int)blocks), dim3(256), 0,
                         (cudaStream_t)stream, (const float*)images, (const float*)masks, transforms, transposed,
                         brightness, out, out_mask, B, p);
  if (le != cudaSuccess) return tdl::cuda_fail(le, "tdl::augment::augment_seg_kernel");
  TDL_LAUNCH_OK("augment_seg_kernel");
  return 0;
}

TDL_API int tdl_augment_seg(const float* images, const float* masks, const float* transforms, const int* transposed,
                            float* out, float* out_mask, int B, int H, int W, int pad, float mean, float stddev,
                            int laplace_2d, void* stream) {
  return tdl_augment_seg_ex(images, masks, 0, transforms, transposed, nullptr, out, out_mask, B, H, W, pad, mean,
                            stddev, laplace_2d, stream);
}
