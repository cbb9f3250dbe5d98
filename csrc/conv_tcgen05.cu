// Convolution as tap-shifted GEMMs on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// Replaces the cuDNN convolutions that the reference reaches through slim.conv2d
// (reference: core/resnet.py:79-85,130-144,164-166,449-492; core/layers.py:43-49).
//
// Design (B200-first, not a cuDNN clone):
//   * Activations are NHWC bf16, weights are [Cout, R*S, Cin] bf16 ("OHWI").
//   * A CTA tile of 128 GEMM rows is a (TN x TH x TW) patch of output pixels.  For filter tap
//     (r, s) the A operand is the input patch shifted by (r*dil - pad, s*dil - pad); ONE 4-D
//     TMA box fetches it, and TMA's out-of-bounds zero fill *is* the convolution padding.
//     Stride comes from the tensor map's element strides, dilation from the coordinates.
//     No im2col buffer, no index math in the mainloop.
//   * K loop = taps x ceil(Cin/64); every stage is a 128x64 (A) and BLOCK_Nx64 (B) bf16 tile in
//     128B-swizzled shared memory feeding tcgen05.mma (M=128, N=BLOCK_N, K=16), fp32
//     accumulators in TMEM, double buffered so the epilogue of tile i overlaps tile i+1.
//   * Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
//     warps 2-9 = epilogue (tcgen05.ld -> bias / residual / ReLU -> bf16 -> smem -> TMA store), two per TMEM lane quarter.
//   * Persistent: grid = #SMs, static round-robin tile schedule with the Cout tile fastest so
//     CTAs running together share the activation patch in L2.
//   * wgrad is the transposed problem: both operands are "MN-major" (pixel index is the GEMM K
//     dimension), split over pixel patches across CTAs, fp32 `red.global.add` into dW.
#include "tdl_common.cuh"

#include <cudaTypedefs.h>

namespace tdl {
namespace conv {

// ------------------------------------------------------------------------------------------
// Host: tensor-map encoding through the driver entry point (no link-time libcuda dependency,
// so the library also loads on GPU-less build machines).
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
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

// dims/box/estr are innermost-first; strides are derived for a dense tensor.
static int encode_map(CUtensorMap* m, CUtensorMapDataType dtype, int elem_bytes, CUtensorMapSwizzle swizzle,
                      const void* base, int rank, const uint64_t* dims, const uint32_t* box, const uint32_t* estr,
                      const uint64_t* strides_bytes) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return tdl::fail(-3, "%s%s", "cuTensorMapEncodeTiled entry point unavailable", "");
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t b[5], e[5];
  uint64_t pitch = (uint64_t)elem_bytes;
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    b[i] = box[i];
    e[i] = estr[i];
    pitch *= dims[i];
    if (i + 1 < rank) gstr[i] = strides_bytes ? strides_bytes[i] : pitch;  // stride of dimension i+1, bytes
  }
  CUresult r = fn(m, dtype, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // A thread that has made no runtime call yet (autograd's backward thread whose first kernel is ours) has no
    // current context for the driver entry point: bind the primary context and encode again.
    cudaFree(nullptr);
    r = fn(m, dtype, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
           swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    snprintf(tdl::err_buf(), 512,
             "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]", (int)r, rank,
             (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
             (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
             rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return -4;
  }
  return 0;
}

static int encode_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint32_t* box,
                       const uint32_t* estr, const uint64_t* strides_bytes = nullptr) {
  return encode_map(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, CU_TENSOR_MAP_SWIZZLE_128B, base, rank, dims, box, estr,
                    strides_bytes);
}

// ------------------------------------------------------------------------------------------
// Problem description shared by fprop and wgrad.
// ------------------------------------------------------------------------------------------
struct Geometry {
  int N, H, W, C;          // input (as seen by the kernel; the flat 1x1 mode rewrites these)
  int P, Q, K;             // output spatial size, output channels
  int R, S;                // filter taps
  int stride, dil, pad_h, pad_w;
  int TN, TH, TW;          // pixel patch owned by one GEMM tile (TN*TH*TW <= rows per tile)
  int tiles_n, tiles_h, tiles_w;
  int num_kb;              // ceil(C / 64)
  int n_tiles;             // Cout tiles (fprop)
  int relu, has_bias, has_res;
  // first-layer ("stem") mode only: byte pitches of the haloed space-to-depth patch in shared memory
  int sp_plane, sp_row, sp_tx, sp_wbytes;
  // kBwdStats: the BatchNorm whose output gradient this (data-gradient) launch produces
  const void* bn_x;
  const float* bn_mean;
  const float* bn_invstd;
  const float* bn_scale;  // non-null: the BatchNorm ends in a ReLU whose mask is recomputed as x * scale + shift > 0
  const float* bn_shift;  //           (the stored gradient stays unmasked; only the sums see the mask)
  int res_mma;  // residual tile added on the tensor core (extra K iterations against an identity block)
  // halo mode (stride-1 RxS > 1x1): the haloed input patch of a K block is loaded ONCE and every tap reads it at a row
  // shift instead of re-fetching a shifted copy per tap (the tap-shifted form is bound by L2->smem bandwidth there)
  int halo, halo_pw, halo_pbytes, halo_ptx, halo_nb, halo_bo;
  // halo_res: the whole weight tensor of the (single) Cout tile stays in shared memory behind the two patch buffers,
  // so a tile costs one patch load per K block and its taps issue back to back with no barrier in between
  int halo_res;
  int has_mask; // res_mma only: zero outputs whose bit in a [pixels][K/8] byte mask is clear (fused ReLU backward)
  // which tap of the weight tensor filter position (r, s) reads: tap_base + r * tap_sr + s * tap_ss (default 0, S, 1).
  // A parity class of a strided data gradient walks a strided subset of the full flipped filter this way, so it needs
  // no re-laid-out copy of its taps.
  int tap_base, tap_sr, tap_ss;
  // wgrad, three taps per CTA ("row mode", 1x3 / 3x3 stride-1 filters on maps up to 62 wide): a stage is TH whole image
  // rows at a pitch of Q + 2 positions for BOTH operands (dY's two extra columns are out of range = zero), so the
  // three taps of a filter row read ONE haloed X tile at row shifts 0 / 1 / 2 instead of three shifted copies.
  // 1 = one N = 192 MMA whose 64-column blocks sit one pixel row (128 B) apart, 2 = three N = 64 MMAs.
  int wrow;
};

// Pick the (TN, TH, TW) pixel patch with the best row utilisation for `rows` GEMM rows.
static void choose_patch(int N, int P, int Q, int rows, int stride, int* tn, int* th, int* tw) {
  double best = -1.0;
  int bn = 1, bh = 1, bw = 1;
  int max_w = Q < rows ? Q : rows;
  if (max_w * stride > 256) max_w = 256 / stride;
  for (int w = 1; w <= max_w; ++w) {
    int max_h = rows / w;
    if (max_h > P) max_h = P;
    if (max_h * stride > 256) max_h = 256 / stride;
    for (int h = 1; h <= max_h; ++h) {
      int n = 1;
      if (w == Q && h == P) {
        n = rows / (w * h);
        if (n > N) n = N;
        if (n < 1) n = 1;
      }
      double tiles = (double)tdl::cdiv(N, n) * tdl::cdiv(P, h) * tdl::cdiv(Q, w);
      double eff = ((double)N * P * Q) / (tiles * rows);
      // prefer wider patches on ties: longer contiguous runs per TMA row
      if (eff > best + 1e-9 || (eff > best - 1e-9 && w > bw)) {
        best = eff;
        bn = n;
        bh = h;
        bw = w;
      }
    }
  }
  *tn = bn;
  *th = bh;
  *tw = bw;
}

// ------------------------------------------------------------------------------------------
// fprop kernel
// ------------------------------------------------------------------------------------------
constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                       // bf16 elements = 128 bytes = one swizzle row
constexpr int kABytes = kBlockM * kBlockK * 2;    // 16 KB
constexpr int kThreads = 192;                     // wgrad: 6 warps
constexpr int kFpropThreads = 320;                // fprop: TMA warp, MMA warp, 2 x 4 epilogue warps
constexpr int kEpiWarp0 = 2;

template <int BLOCK_N, bool kBwd = false>
struct FpropCfg {
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // accumulator ring in TMEM: narrow tiles finish in a few hundred ns, so four buffers keep the MMA warp from waiting
  // on the epilogue's hand-back latency every other tile
  static constexpr int kAccBufs = BLOCK_N <= 128 ? 4 : 2;
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;
  static constexpr int kOutBufs = 2;             // 128-row x 64-channel bf16 staging tile per epilogue group
  static constexpr int kOutBytes = kOutBufs * kBlockM * 128;
  // kBwd (BatchNorm-backward sums in the epilogue): one more 128-row x 64-channel tile per epilogue group, the
  // BatchNorm input at the output tile's position, prefetched by TMA while the accumulator is drained
  static constexpr int kXBytes = kBwd ? kOutBufs * kBlockM * 128 : 0;
  static constexpr int kStatBytes = 4 * 2 * 64 * 4;  // per-warp (sum, sumsq) partials of one 64-channel chunk
  static constexpr int kBarBytes = 256;    // (2*kStages + 4) mbarriers + the TMEM slot
  static constexpr int kAlignSlack = 384;  // the dynamic smem window is 1 KB aligned in practice; checked at run time
  static constexpr int kStages = (227 * 1024 - kOutBytes - kXBytes - kStatBytes - kAlignSlack - kBarBytes) / kStageBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + kXBytes + kStatBytes + kAlignSlack + kBarBytes;
  static_assert((2 * kStages + 2 * kAccBufs + 1 + 4 + 2) * 8 + 4 <= kBarBytes, "barrier region too small");
};

// kFused: bias / residual / ReLU present in the epilogue.  kStats: per-channel (sum, sum of squares) of the bf16
// output accumulated into this CTA's slot of `stats` ([gridDim.x][2][K] fp32) for the BatchNorm that follows.
//
// kStem: first-layer mode for a stride-2 convolution on <= 4 input channels (7x7/3x3 roots: core/resnet.py:219-224,
// :164, core/xception.py:352).  The image arrives space-to-depth packed as two 8-channel planes
// [N][H/2][plane][W/2][8] (tdl_conv_stem_pack), which turns the stride-2 RxS conv into a stride-1 (R/2+1)^2-tap conv on
// 16 channels.  A GEMM tile is ONE output row (<= 120 pixels): one TMA box brings the haloed patch
// [taps_h][plane][TW + taps_w - 1 pixels][8 ch] into shared memory un-swizzled, and because a pixel is exactly one
// 16-byte core-matrix row, every tap's A operand is just that patch at a shifted start address (K-major "interleave"
// descriptor: SBO = 8 pixels, LBO = plane pitch).  No im2col, no per-tap loads; the whole weight tensor
// [K][taps*16] stays resident in shared memory.
//
// kResMma: `y = conv(x) + residual` with the residual added ON THE TENSOR CORE: after the real K iterations the
// producer streams the residual tile (same pixel patch, 64 channels at a time) through the normal smem pipeline as an
// A operand and the MMA warp multiplies it by a 64x64 identity block into the matching 64 accumulator columns
// (bf16 x 1.0 accumulates exactly in fp32).  The epilogue stays the lean one, and the residual read is prefetched by
// the same mbarrier pipeline as everything else.  Used by the data gradient of a unit's first conv to absorb the
// identity-shortcut gradient (otherwise a separate 3-pass elementwise add per residual unit).
template <int BLOCK_N, bool kFused, bool kStats, bool kStem = false, bool kResMma = false, bool kBwdStats = false>
__global__ void __launch_bounds__(kFpropThreads, 1)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                  const __grid_constant__ CUtensorMap tm_y, const __grid_constant__ CUtensorMap tm_res,
                  const __grid_constant__ CUtensorMap tm_id, const __grid_constant__ CUtensorMap tm_bx,
                  const Geometry g, const float* __restrict__ bias, const bf16* __restrict__ residual,
                  float* __restrict__ stats) {
  ptx::pdl_trigger();
  using Cfg = FpropCfg<BLOCK_N, kBwdStats>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if (smem - smem_raw > Cfg::kAlignSlack) {
    if (threadIdx.x == 0) printf("tdl_b200: dynamic shared memory base is misaligned by %d bytes\n", (int)(smem - smem_raw));
    __trap();
  }
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint8_t* smem_out = smem + kStages * Cfg::kStageBytes;
  uint8_t* smem_x = smem_out + Cfg::kOutBytes;  // kBwdStats: BatchNorm-input tile per epilogue group
  float* smem_stat = reinterpret_cast<float*>(smem_x + Cfg::kXBytes);  // [4 warps][2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_x + Cfg::kXBytes + Cfg::kStatBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  constexpr int kAcc = Cfg::kAccBufs;
  uint64_t* tmem_empty = tmem_full + kAcc;
  uint64_t* w_bar = tmem_empty + kAcc;  // stem mode: resident weights have landed
  uint64_t* patch_full = w_bar + 1;     // halo mode: two input-patch buffers
  uint64_t* patch_empty = patch_full + 2;
  uint64_t* x_bar = patch_empty + 2;    // kBwdStats: the group's BatchNorm-input tile has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x_bar + 2);
  // halo mode: weight-tile ring behind the two patch buffers; resident form: all weights behind halo_nb patch buffers
  uint8_t* smem_bh = smem + (g.halo_res ? g.halo_nb : 2) * g.halo_pbytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = g.tiles_n * g.tiles_h * g.tiles_w;
  const int total_tiles = m_tiles * g.n_tiles;
  const int k_iters = kStem ? 1 : g.R * g.S * g.num_kb;
  const uint32_t a_tx = (uint32_t)(g.TN * g.TH * g.TW) * 128u;
  const uint32_t stage_tx = a_tx + (uint32_t)Cfg::kBBytes;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_x);
    ptx::prefetch_tmap(&tm_w);
    ptx::prefetch_tmap(&tm_y);
    if constexpr (kResMma) {
      ptx::prefetch_tmap(&tm_res);
      ptx::prefetch_tmap(&tm_id);
    }
    if constexpr (kBwdStats) ptx::prefetch_tmap(&tm_bx);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        ptx::mbar_init(&full_bar[i], 1);
        ptx::mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < kAcc; ++i) {
        ptx::mbar_init(&tmem_full[i], 1);
        ptx::mbar_init(&tmem_empty[i], 4);  // one arrival per warp of the epilogue group that drains the tile
      }
      ptx::mbar_init(w_bar, 1);
      for (int i = 0; i < 2; ++i) {
        ptx::mbar_init(&patch_full[i], 1);
        ptx::mbar_init(&patch_empty[i], 1);
        ptx::mbar_init(&x_bar[i], 1);
      }
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  ptx::pdl_wait();  // set-up above overlapped the previous kernel's tail; global memory is touched from here on
  if constexpr (kStats || kBwdStats) {  // this CTA's statistics slot starts at zero
    float* slot = stats + (size_t)blockIdx.x * 2 * g.K;
    for (int i = threadIdx.x; i < 2 * g.K; i += kFpropThreads) slot[i] = 0.f;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The whole warp walks the (warp-uniform) schedule and only the issue instructions run on one elected lane:
    // with the loop itself inside `if (lane == 0)` ptxas cannot treat the descriptors / coordinates as uniform and
    // wraps every TMA and MMA issue in an elect-and-broadcast loop (tens of instructions per issue, which made the
    // issuing thread -- not the tensor core -- the limit of narrow tiles).  Elected once: elect.sync is a warp
    // synchronisation, re-electing per instruction cost more than it saved.
    const bool leader = ptx::elect_one();
    {
      int stage = 0;
      uint32_t phase = 0;
      int pb = 0;           // halo mode: patch buffer index / parity
      uint32_t pphase = 0;
      if constexpr (kStem) {
        if (leader) ptx::mbar_expect_tx(w_bar, (uint32_t)g.sp_wbytes);
        for (int a = 0; a * 8192 < g.sp_wbytes; ++a)
          if (leader) ptx::tma_load_2d(smem_b + a * 8192, &tm_w, w_bar, a * 64, 0);
      }
      if (!kStem && !kResMma && g.halo_res) {  // resident weights: [K block][tap] tiles, one barrier for all of them
        const int taps = g.R * g.S;
        if (leader) ptx::mbar_expect_tx(w_bar, (uint32_t)(taps * g.num_kb * Cfg::kBBytes));
        for (int kb2 = 0; kb2 < g.num_kb; ++kb2)
          for (int t = 0; t < taps; ++t)
            if (leader) ptx::tma_load_3d(smem_bh + (kb2 * taps + t) * Cfg::kBBytes, &tm_w, w_bar, kb2 * kBlockK, t, 0);
      }
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % g.n_tiles;
        const int m_tile = tile / g.n_tiles;
        const int tw_i = m_tile % g.tiles_w;
        const int th_i = (m_tile / g.tiles_w) % g.tiles_h;
        const int tn_i = m_tile / (g.tiles_w * g.tiles_h);
        if constexpr (kStem) {
          // coordinates: (8-byte units along the plane line, plane, packed row, image); out-of-range = zero padding
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) {
            ptx::mbar_expect_tx(&full_bar[stage], (uint32_t)g.sp_tx);
            ptx::tma_load_4d(smem_a + stage * kABytes, &tm_x, &full_bar[stage], (tw_i * g.TW - g.pad_w) * 2, 0,
                             th_i - g.pad_h, tn_i);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          continue;
        }
        const int w_in = tw_i * g.TW * g.stride - g.pad_w;
        const int h_in = th_i * g.TH * g.stride - g.pad_h;
        const int n_in = tn_i * g.TN;
        if (!kStem && !kResMma && g.halo) {
          const int taps = g.R * g.S;
          if (g.halo_res) {  // patches ride the stage barriers: halo_nb buffers, nothing else to load
            for (int kb2 = 0; kb2 < g.num_kb; ++kb2) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              if (leader) ptx::mbar_expect_tx(&full_bar[stage], (uint32_t)g.halo_ptx);
              if (leader) ptx::tma_load_4d(smem + stage * g.halo_pbytes, &tm_x, &full_bar[stage], kb2 * kBlockK, w_in, h_in, n_in);
              if (++stage == g.halo_nb) {
                stage = 0;
                phase ^= 1;
              }
            }
            continue;
          }
          for (int kb2 = 0; kb2 < g.num_kb; ++kb2) {
            ptx::mbar_wait(&patch_empty[pb], pphase ^ 1);
            if (leader) ptx::mbar_expect_tx(&patch_full[pb], (uint32_t)g.halo_ptx);
            if (leader) ptx::tma_load_4d(smem + pb * g.halo_pbytes, &tm_x, &patch_full[pb], kb2 * kBlockK, w_in, h_in, n_in);
            if ((pb ^= 1) == 0) pphase ^= 1;
            for (int t = 0; t < taps; ++t) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              if (leader) ptx::mbar_expect_tx(&full_bar[stage], (uint32_t)Cfg::kBBytes);
              if (leader) ptx::tma_load_3d(smem_bh + stage * Cfg::kBBytes, &tm_w, &full_bar[stage], kb2 * kBlockK, t,
                               n_tile * BLOCK_N);
              if (++stage == g.halo_nb) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
          continue;
        }
        int tap = g.tap_base, kb = 0, r = 0, s = 0;  // advanced incrementally: no divisions on the TMA-issue path
        for (int it = 0; it < k_iters; ++it) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) {
            ptx::mbar_expect_tx(&full_bar[stage], stage_tx);
            ptx::tma_load_4d(smem_a + stage * kABytes, &tm_x, &full_bar[stage], kb * kBlockK, w_in + s * g.dil,
                             h_in + r * g.dil, n_in);
            ptx::tma_load_3d(smem_b + stage * Cfg::kBBytes, &tm_w, &full_bar[stage], kb * kBlockK, tap,
                             n_tile * BLOCK_N);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          if (++kb == g.num_kb) {
            kb = 0;
            tap += g.tap_ss;
            if (++s == g.S) {
              s = 0;
              ++r;
              tap = g.tap_base + r * g.tap_sr;
            }
          }
        }
        if constexpr (kResMma) {
          const int res_chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
          for (int j = 0; j < res_chunks; ++j) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            if (leader) {
              ptx::mbar_expect_tx(&full_bar[stage], a_tx + 8192u);
              ptx::tma_load_4d(smem_a + stage * kABytes, &tm_res, &full_bar[stage], n_tile * BLOCK_N + j * 64,
                               tw_i * g.TW, th_i * g.TH, tn_i * g.TN);
              ptx::tma_load_2d(smem_b + stage * Cfg::kBBytes, &tm_id, &full_bar[stage], 0, 0);
            }
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp walks the schedule, one elected lane issues) =====================
    const bool leader = ptx::elect_one();
    {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(kBlockM, BLOCK_N, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int hpb = 0;  // halo mode: patch buffer index / parity
      uint32_t hpphase = 0;
      uint64_t stem_a0 = 0, stem_b0 = 0;
      if constexpr (kStem) {
        stem_a0 = ptx::umma_smem_desc_nosw(ptx::smem_u32(smem_a), (uint32_t)g.sp_plane, 128);
        stem_b0 = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b), 16, 1024);
        if (blockIdx.x < total_tiles) ptx::mbar_wait(w_bar, 0);
      }
      if (!kStem && !kResMma && g.halo_res) {
        if (blockIdx.x < total_tiles) ptx::mbar_wait(w_bar, 0);
      }
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BLOCK_N);
        if (!kStem && !kResMma && g.halo) {
          const uint32_t sbo = (uint32_t)g.halo_pw * 128u;  // one tile row (8 pixels) per 8-row group of the M tile
          bool first = true;
          if (g.halo_res) {  // weights resident: the taps of a K block issue back to back, one barrier per patch
            uint32_t b_tile = ptx::smem_u32(smem_bh);
            for (int kb2 = 0; kb2 < g.num_kb; ++kb2) {
              ptx::mbar_wait(&full_bar[stage], phase);
              ptx::tc_fence_after();
              const uint32_t a_patch = ptx::smem_u32(smem + stage * g.halo_pbytes);
              uint32_t row_off = 0;  // r * dil * patch width, in pixels
              for (int r = 0; r < g.R; ++r, row_off += (uint32_t)(g.dil * g.halo_pw)) {
                for (int sx = 0; sx < g.S; ++sx, b_tile += (uint32_t)Cfg::kBBytes) {
                  const uint32_t shift = (uint32_t)(sx * g.dil);
                  const uint64_t a_desc = ptx::umma_smem_desc_sw128_bo(a_patch + (row_off + shift) * 128u, 16, sbo, 0u);
                  const uint64_t b_desc = ptx::umma_smem_desc_sw128(b_tile, 16, 1024);
#pragma unroll
                  for (int k = 0; k < kBlockK / 16; ++k) {
                    if (leader) ptx::umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc,
                                   (first && k == 0) ? 0u : 1u);
                  }
                  first = false;
                }
              }
              if (leader) ptx::umma_commit(&empty_bar[stage]);  // the patch may be overwritten once these MMAs retire
              if (++stage == g.halo_nb) {
                stage = 0;
                phase ^= 1;
              }
            }
            if (leader) ptx::umma_commit(&tmem_full[acc]);
            if (++acc == kAcc) {
              acc = 0;
              acc_phase ^= 1;
            }
            continue;
          }
          for (int kb2 = 0; kb2 < g.num_kb; ++kb2) {
            ptx::mbar_wait(&patch_full[hpb], hpphase);
            ptx::tc_fence_after();
            const uint32_t a_patch = ptx::smem_u32(smem + hpb * g.halo_pbytes);
            uint32_t row_off = 0;  // r * dil * patch width, in pixels
            for (int r = 0; r < g.R; ++r, row_off += (uint32_t)(g.dil * g.halo_pw)) {
              for (int sx = 0; sx < g.S; ++sx) {
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tc_fence_after();
                const uint32_t shift = (uint32_t)(sx * g.dil);
                const uint64_t a_desc = ptx::umma_smem_desc_sw128_bo(a_patch + (row_off + shift) * 128u, 16, sbo,
                                                                     g.halo_bo ? shift : 0u);
                const uint64_t b_desc =
                    ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_bh + stage * Cfg::kBBytes), 16, 1024);
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k) {
                  if (leader) ptx::umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc,
                                 (first && k == 0) ? 0u : 1u);
                }
                first = false;
                if (leader) ptx::umma_commit(&empty_bar[stage]);
                if (++stage == g.halo_nb) {
                  stage = 0;
                  phase ^= 1;
                }
              }
            }
            if (leader) ptx::umma_commit(&patch_empty[hpb]);  // the patch may be overwritten once these MMAs retire
            if ((hpb ^= 1) == 0) hpphase ^= 1;
          }
          if (leader) ptx::umma_commit(&tmem_full[acc]);
          if (++acc == kAcc) {
            acc = 0;
            acc_phase ^= 1;
          }
          continue;
        }
        for (int it = 0; it < k_iters; ++it) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          if constexpr (kStem) {
            // one MMA per packed tap
            // (descriptor offsets advance with counters: this thread's own instruction stream is the critical
            // path, a division per tap cost more than the MMA itself)
            const uint64_t a0 = stem_a0 + (uint64_t)((uint32_t)(stage * kABytes) >> 4);
            const int taps = g.R * g.S;
            if (leader) {  // one divergent region per stage, not one per instruction
              uint32_t a_row = 0, a_col = 0;  // in 16-byte units
              for (int t = 0; t < taps; ++t) {
                const uint32_t b_off = (uint32_t)((t >> 2) * 512 + (t & 3) * 2);
                ptx::umma_bf16(d_tmem, a0 + (uint64_t)(a_row + a_col), stem_b0 + (uint64_t)b_off, idesc, t > 0 ? 1u : 0u);
                if (++a_col == (uint32_t)g.S) {
                  a_col = 0;
                  a_row += (uint32_t)g.sp_row >> 4;
                }
              }
              ptx::umma_commit(&empty_bar[stage]);
            }
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
            continue;
          }
          const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * kABytes), 16, 1024);
          const uint64_t b_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBBytes), 16, 1024);
          if (leader) {
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k) {
              // advance 16 bf16 (32 bytes) along K inside the 128B swizzle row: +2 in 16-byte units
              ptx::umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc,
                             (it > 0 || k > 0) ? 1u : 0u);
            }
            ptx::umma_commit(&empty_bar[stage]);  // frees the smem stage when these MMAs retire
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if constexpr (kResMma) {
          constexpr uint32_t idesc64 = ptx::umma_idesc_bf16(kBlockM, 64, 0, 0);
          const int n_tile = tile % g.n_tiles;
          const int res_chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
          for (int j = 0; j < res_chunks; ++j) {
            ptx::mbar_wait(&full_bar[stage], phase);
            ptx::tc_fence_after();
            const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * kABytes), 16, 1024);
            const uint64_t b_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBBytes), 16, 1024);
            if (leader) {
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k)
                ptx::umma_bf16(d_tmem + (uint32_t)(j * 64), a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k),
                               idesc64, 1u);
              ptx::umma_commit(&empty_bar[stage]);
            }
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        if (leader) ptx::umma_commit(&tmem_full[acc]);  // accumulator complete -> epilogue
        if (++acc == kAcc) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue warps =====================
    // TMEM -> registers -> (bias / residual / ReLU) -> bf16 -> 128B-swizzled smem tile -> TMA store.
    // The 4-D store box is the pixel patch, so TMA clips everything outside the image / channel range.
    //
    // Two independent groups of four warps (a warp may touch TMEM lanes 32*(warp%4)..+31, so each group covers all
    // 128 lanes).  Group 0 drains the even tiles of this CTA, group 1 the odd ones, each with its own staging
    // buffer, named barrier and TMA-store issuer, so the dependent chain of one tile (accumulator wait, tcgen05.ld,
    // convert, stage, store, statistics) overlaps the chain of the next tile instead of adding to it.
    const int ewarp = warp - kEpiWarp0;  // 0..7
    const int group = ewarp >> 2;
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const int patch = g.TH * g.TW;
    const int dn = row / patch;
    const int rem = row - dn * patch;
    const int dh = rem / g.TW;
    const int dw = rem - dh * g.TW;
    const bool row_in_patch = row < g.TN * patch;
    const bool issuer = ((ewarp & 3) == 0 && lane == 0);
    const int bar_id = 1 + group;
    const uint32_t row_off = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const uint32_t tmem_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
    uint8_t* out_tile = smem_out + group * (kBlockM * 128);
    uint8_t* x_tile = smem_x + group * (kBlockM * 128);
    uint32_t x_phase = 0;
    // ---- fused BatchNorm statistics (kStats) ----
    // The bf16 tile staged for the TMA store doubles as the transpose: after the store is issued every thread reads
    // ONE channel pair (4 bytes) of its warp's 32 rows back from shared memory -- a warp reads a whole 128-byte row
    // per instruction, conflict-free under the swizzle -- and keeps (sum, sum of squares) of its two channels in
    // registers across all tiles of the current Cout tile.  The four row groups are folded through the staging
    // buffer only when the Cout tile changes and at the end (`flush_stats`).  (The first version reduced the
    // accumulator registers across lanes with shuffle butterflies: ~500 of the ~1000 epilogue instructions per
    // tile, which made the stats variant 2.8x slower than the plain kernel on output-heavy layers.)
    constexpr int kChunks = BLOCK_N / 64;
    float sacc[kChunks][4];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) sacc[c][0] = sacc[c][1] = sacc[c][2] = sacc[c][3] = 0.f;
    int stat_n_tile = -1;
    auto flush_stats = [&](int nt) {
      if (issuer) ptx::tma_store_wait_read<0>();  // nobody is reading the staging buffer any more
      ptx::named_bar_sync(bar_id, 128);
      float* st = reinterpret_cast<float*>(out_tile);  // [4 row groups][kChunks][4][32 lanes]
#pragma unroll
      for (int c = 0; c < kChunks; ++c)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          st[((quarter * kChunks + c) * 4 + v) * 32 + lane] = sacc[c][v];
          sacc[c][v] = 0.f;
        }
      ptx::named_bar_sync(bar_id, 128);
      float* slot = stats + (size_t)blockIdx.x * 2 * g.K;
      for (int idx = quarter * 32 + lane; idx < kChunks * 128; idx += 128) {
        const int c = idx >> 7, v = (idx >> 5) & 3, ln = idx & 31;
        float tot = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) tot += st[((e * kChunks + c) * 4 + v) * 32 + ln];
        const int ch = nt * BLOCK_N + c * 64 + 2 * ln + (v >> 1);  // v: 0 = sum ch0, 1 = sumsq ch0, 2 = sum ch1, 3 = sumsq ch1
        if (ch < g.K) {
          if constexpr (kBwdStats) {
            if (v & 1) tot *= g.bn_invstd[ch];  // sum dz * (x - mean)  ->  sum dz * xhat
          }
          atomicAdd(&slot[(size_t)(v & 1) * g.K + ch], tot);  // both groups add into the CTA's slot
        }
      }
      ptx::named_bar_sync(bar_id, 128);  // the staging buffer is free again
    };
    uint32_t tile_count = (uint32_t)group;
    for (int tile = blockIdx.x + group * gridDim.x; tile < total_tiles; tile += 2 * gridDim.x, tile_count += 2) {
      const int acc = (int)(tile_count % kAcc);
      const uint32_t acc_phase = (tile_count / kAcc) & 1u;
      const int n_tile = tile % g.n_tiles;
      const int m_tile = tile / g.n_tiles;
      const int tw_i = m_tile % g.tiles_w;
      const int th_i = (m_tile / g.tiles_w) % g.tiles_h;
      const int tn_i = m_tile / (g.tiles_w * g.tiles_h);
      bool valid = true;
      size_t pix = 0;
      if constexpr (kFused || kStats || kResMma || kBwdStats) {
        const int n = tn_i * g.TN + dn;
        const int p = th_i * g.TH + dh;
        const int q = tw_i * g.TW + dw;
        valid = row_in_patch && n < g.N && p < g.P && q < g.Q;
        pix = ((size_t)n * g.P + p) * g.Q + q;
      }
      const int chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
      uint32_t vm = 0xffffffffu;  // which of this warp's 32 rows are real output pixels
      if constexpr (kStats || kBwdStats) {
        if (n_tile != stat_n_tile) {
          if (stat_n_tile >= 0) flush_stats(stat_n_tile);
          stat_n_tile = n_tile;
        }
        vm = __ballot_sync(0xffffffffu, valid);
      }
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
#pragma unroll
      for (int cc = 0; cc < kChunks; ++cc) {
        if (cc >= chunks) break;
        const int co_chunk = n_tile * BLOCK_N + cc * 64;
        if (issuer) ptx::tma_store_wait_read<0>();  // the previous store of this group has drained the buffer
        ptx::named_bar_sync(bar_id, 128);
        if constexpr (kBwdStats) {
          // every thread of the group is past the previous chunk's read-back: fetch the BatchNorm input at this
          // chunk's position now, it lands while the accumulator is drained, converted and stored
          if (issuer) {
            ptx::mbar_expect_tx(&x_bar[group], a_tx);
            ptx::tma_load_4d(x_tile, &tm_bx, &x_bar[group], co_chunk, tw_i * g.TW, th_i * g.TH, tn_i * g.TN);
          }
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem_lane + (uint32_t)(acc * BLOCK_N + cc * 64 + half * 32), v);
          ptx::tmem_ld_wait();
          if (half == 1 && cc == chunks - 1) {
            // every tcgen05.ld of this accumulator has completed: hand the TMEM buffer back to the MMA warp
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty[acc]);
          }
          const int co0 = co_chunk + half * 32;
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if constexpr (kFused) {
#pragma unroll
            for (int j8 = 0; j8 < 4; ++j8) {
              const int co = co0 + j8 * 8;
              if (co < g.K) {
                if (g.has_bias) {
                  const float4 b0 = *reinterpret_cast<const float4*>(bias + co);
                  const float4 b1 = *reinterpret_cast<const float4*>(bias + co + 4);
                  f[j8 * 8 + 0] += b0.x; f[j8 * 8 + 1] += b0.y; f[j8 * 8 + 2] += b0.z; f[j8 * 8 + 3] += b0.w;
                  f[j8 * 8 + 4] += b1.x; f[j8 * 8 + 5] += b1.y; f[j8 * 8 + 6] += b1.z; f[j8 * 8 + 7] += b1.w;
                }
                if (g.has_res && valid) {
                  float rr[8];
                  unpack8(ld8(residual + pix * g.K + co), rr);
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j8 * 8 + j] += rr[j];
                }
                if (g.relu) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j8 * 8 + j] = fmaxf(f[j8 * 8 + j], 0.f);
                }
              }
            }
          }
          if constexpr (kResMma || kBwdStats) {
            // fused ReLU backward of the tensor this gradient belongs to: bit j of byte (pixel, c/8) = "was > 0"
            if (g.has_mask && valid && co0 < g.K) {
              const uint8_t* mrow = reinterpret_cast<const uint8_t*>(residual) + pix * (size_t)(g.K >> 3) + (co0 >> 3);
              const uint32_t mbits = *reinterpret_cast<const uint32_t*>(mrow);
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = ((mbits >> j) & 1u) ? f[j] : 0.f;
            }
          }
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
            packed[j] = *reinterpret_cast<const uint32_t*>(&h2);
          }
#pragma unroll
          for (int j8 = 0; j8 < 4; ++j8) {
            const uint32_t chunk16 = (uint32_t)(half * 4 + j8);
            *reinterpret_cast<uint4*>(out_tile + row_off + ((chunk16 ^ sw) << 4)) =
                make_uint4(packed[j8 * 4], packed[j8 * 4 + 1], packed[j8 * 4 + 2], packed[j8 * 4 + 3]);
          }
        }
        ptx::fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA (async proxy)
        ptx::named_bar_sync(bar_id, 128);
        if (issuer) {
          ptx::tma_store_4d(&tm_y, out_tile, co_chunk, tw_i * g.TW, th_i * g.TH, tn_i * g.TN);
          ptx::tma_store_commit();
        }
        if constexpr (kBwdStats) {
          // BatchNorm-backward reductions of the gradient as stored (masked, bf16): per channel sum(dz) and
          // sum(dz * (x - mean)), x = the BatchNorm's input at the same pixels (the TMA tile fetched at chunk start,
          // same swizzled layout as the staged tile).  Replaces the separate pass over (dz, x) (tdl_bn_bwd_reduce).
          const uint32_t roff = (uint32_t)(quarter * 32) * 128u + ((uint32_t)(lane & 3) << 2);
          const uint8_t* rows = out_tile + roff;
          const uint8_t* xrows = x_tile + roff;
          const uint32_t c16 = (uint32_t)(lane >> 2);
          // per-channel constants of this thread's two channels: fetched per chunk (L1 hits) rather than held in
          // registers for the whole kernel -- the epilogue is at the register ceiling of a 320-thread CTA
          const int ch0 = co_chunk + 2 * lane;
          const bool relu_rc = g.bn_scale != nullptr;
          float m0 = 0.f, m1 = 0.f, a0 = 0.f, b0 = 0.f, a1 = 0.f, b1 = 0.f;
          if (ch0 < g.K) {
            const float2 mm = __ldg(reinterpret_cast<const float2*>(g.bn_mean + ch0));
            m0 = mm.x;
            m1 = mm.y;
            if (relu_rc) {
              const float2 aa = __ldg(reinterpret_cast<const float2*>(g.bn_scale + ch0));
              const float2 bb = __ldg(reinterpret_cast<const float2*>(g.bn_shift + ch0));
              a0 = aa.x; a1 = aa.y; b0 = bb.x; b1 = bb.y;
            }
          }
          ptx::mbar_wait(&x_bar[group], x_phase);
          x_phase ^= 1u;
          float s0 = 0.f, q0 = 0.f, s1 = 0.f, q1 = 0.f;
#pragma unroll 8
          for (int i = 0; i < 32; ++i) {
            const uint32_t so = (uint32_t)i * 128u + ((c16 ^ (uint32_t)(i & 7)) << 4);
            const uint32_t wv = *reinterpret_cast<const uint32_t*>(rows + so);
            const bool ok = (vm >> i) & 1u;
            // rows past the patch are never written by the TMA load: whatever shared memory held must not reach the sums
            const uint32_t xw = ok ? *reinterpret_cast<const uint32_t*>(xrows + so) : 0u;
            const float x0 = __uint_as_float(xw << 16), x1 = __uint_as_float(xw & 0xffff0000u);
            // recomputed mask: fmaf(x, scale, shift) > 0, the same expression bn_apply evaluated in the forward pass
            const bool k0 = ok && (!relu_rc || fmaf(x0, a0, b0) > 0.f);
            const bool k1 = ok && (!relu_rc || fmaf(x1, a1, b1) > 0.f);
            const float f0 = k0 ? __uint_as_float(wv << 16) : 0.f;
            const float f1 = k1 ? __uint_as_float(wv & 0xffff0000u) : 0.f;
            s0 += f0;
            q0 = fmaf(f0, x0 - m0, q0);
            s1 += f1;
            q1 = fmaf(f1, x1 - m1, q1);
          }
          sacc[cc][0] += s0;
          sacc[cc][1] += q0;
          sacc[cc][2] += s1;
          sacc[cc][3] += q1;
        }
        if constexpr (kStats) {
          // statistics of the values as stored (bf16) over this warp's 32 rows, rows outside the image masked out
          const uint8_t* rows = out_tile + (uint32_t)(quarter * 32) * 128u + ((uint32_t)(lane & 3) << 2);
          const uint32_t c16 = (uint32_t)(lane >> 2);
          float s0 = 0.f, q0 = 0.f, s1 = 0.f, q1 = 0.f;
#pragma unroll
          for (int i0 = 0; i0 < 32; i0 += 16) {
            uint32_t wv[16];
#pragma unroll
            for (int i = 0; i < 16; ++i)
              wv[i] = *reinterpret_cast<const uint32_t*>(rows + (i0 + i) * 128 + ((c16 ^ (uint32_t)(i & 7)) << 4));
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const bool ok = (vm >> (i0 + i)) & 1u;
              const float f0 = ok ? __uint_as_float(wv[i] << 16) : 0.f;
              const float f1 = ok ? __uint_as_float(wv[i] & 0xffff0000u) : 0.f;
              s0 += f0;
              q0 = fmaf(f0, f0, q0);
              s1 += f1;
              q1 = fmaf(f1, f1, q1);
            }
          }
          sacc[cc][0] += s0;
          sacc[cc][1] += q0;
          sacc[cc][2] += s1;
          sacc[cc][3] += q1;
        }
      }
    }
    if constexpr (kStats || kBwdStats) {
      if (stat_n_tile >= 0) flush_stats(stat_n_tile);
    }
    if (issuer) ptx::tma_store_wait_all<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------
// fprop on CTA pairs (cta_group::2)
// ------------------------------------------------------------------------------------------
// The single-CTA kernel above moves (128 + BLOCK_N) x 64 operand elements from L2 into shared memory for every
// 128 x BLOCK_N x 64 MMA block.  With BLOCK_N = 256 that is 48 KB per 512 tensor-core cycles = 96 B/clk per SM, and all
// 148 SMs together can pull only ~43 B/clk each out of L2: layers with enough arithmetic to be tensor-bound run at
// ~45 % of the tensor peak because of operand traffic, not because of the MMA rate (this is also where a big cuBLAS
// GEMM saturates on this chip).  Here two CTAs of a cluster (one TPC) share ONE M = 256 MMA: each CTA owns its own
// 128-pixel tile (A operand, accumulator rows, epilogue) and loads only HALF of the weight tile; the tensor core reads
// the two halves out of both CTAs' shared memory.  Operand bytes per CTA and stage drop from 48 KB to 32 KB
// (BLOCK_N = 256) or 32 KB to 24 KB (BLOCK_N = 128), and the freed shared memory deepens the pipeline.
//   * the even CTA of the pair issues every MMA (tcgen05.mma.cta_group::2) and owns the "full" barriers: both CTAs'
//     TMA loads count their bytes on the leader's barrier (cp.async.bulk.tensor ... .cta_group::2);
//   * tcgen05.commit ... .multicast::cluster frees a pipeline stage / publishes an accumulator in BOTH CTAs;
//   * the epilogue warps of both CTAs hand their accumulator back by arriving on the leader's barrier.
// Epilogue variants (bias / residual / ReLU, BatchNorm statistics, residual on the tensor core + mask) are the
// single-CTA kernel's, per CTA.  Not here: the first-layer mode, halo mode, fused BatchNorm-backward sums.
template <int BLOCK_N>
struct Fprop2Cfg {
  static constexpr int kBHalfBytes = (BLOCK_N / 2) * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBHalfBytes;
  static constexpr int kAccBufs = BLOCK_N <= 128 ? 4 : 2;
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;
  static constexpr int kOutBytes = 2 * kBlockM * 128;
  static constexpr int kBarBytes = 256;
  static constexpr int kAlignSlack = 384;
  static constexpr int kStagesRaw = (227 * 1024 - kOutBytes - kAlignSlack - kBarBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + kAlignSlack + kBarBytes;
  static_assert((2 * kStages + 2 * kAccBufs) * 8 + 4 <= kBarBytes, "barrier region too small");
};

template <int BLOCK_N, bool kFused, bool kStats, bool kResMma>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFpropThreads, 1)
conv_fprop2_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                   const __grid_constant__ CUtensorMap tm_y, const __grid_constant__ CUtensorMap tm_res,
                   const __grid_constant__ CUtensorMap tm_id, const Geometry g, const float* __restrict__ bias,
                   const bf16* __restrict__ residual, float* __restrict__ stats) {
  ptx::pdl_trigger();
  using Cfg = Fprop2Cfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kAcc = Cfg::kAccBufs;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if (smem - smem_raw > Cfg::kAlignSlack) {
    if (threadIdx.x == 0) printf("tdl_b200: dynamic shared memory base is misaligned by %d bytes\n", (int)(smem - smem_raw));
    __trap();
  }
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  uint8_t* smem_out = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_out + Cfg::kOutBytes);
  uint64_t* full_bar = bars;                 // used in the leader only: bytes of BOTH CTAs land on it
  uint64_t* empty_bar = bars + kStages;      // per CTA: the leader's commit frees the stage in both
  uint64_t* tmem_full = bars + 2 * kStages;  // per CTA
  uint64_t* tmem_empty = tmem_full + kAcc;   // used in the leader only: 2 CTAs x 4 epilogue warps arrive
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + kAcc);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();  // 0 = leader
  const int pair = (int)blockIdx.x >> 1;
  const int num_pairs = (int)gridDim.x >> 1;

  const int m_tiles = g.tiles_n * g.tiles_h * g.tiles_w;
  const int pair_tiles = ((m_tiles + 1) >> 1) * g.n_tiles;  // an odd last M tile pairs with an all-padding one
  const int k_iters = g.R * g.S * g.num_kb;
  const uint32_t a_tx = (uint32_t)(g.TN * g.TH * g.TW) * 128u;
  const uint32_t stage_tx = 2u * (a_tx + (uint32_t)Cfg::kBHalfBytes);  // both CTAs' loads of one stage

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_x);
    ptx::prefetch_tmap(&tm_w);
    ptx::prefetch_tmap(&tm_y);
    if constexpr (kResMma) {
      ptx::prefetch_tmap(&tm_res);
      ptx::prefetch_tmap(&tm_id);
    }
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        ptx::mbar_init(&full_bar[i], 1);
        ptx::mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < kAcc; ++i) {
        ptx::mbar_init(&tmem_full[i], 1);
        ptx::mbar_init(&tmem_empty[i], 8);  // one arrival per epilogue warp of the draining group, from both CTAs
      }
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
  }
  ptx::pdl_wait();
  if constexpr (kStats) {
    float* slot = stats + (size_t)blockIdx.x * 2 * g.K;
    for (int i = threadIdx.x; i < 2 * g.K; i += kFpropThreads) slot[i] = 0.f;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // both CTAs' barriers are initialised and both TMEM allocations are done
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs: own A tile, own half of the weight tile) =====================
    const bool leader_lane = ptx::elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int pt = pair; pt < pair_tiles; pt += num_pairs) {
      const int n_tile = pt % g.n_tiles;
      const int m_tile = 2 * (pt / g.n_tiles) + (int)rank;  // may be == m_tiles: every coordinate out of range = zeros
      const int tw_i = m_tile % g.tiles_w;
      const int th_i = (m_tile / g.tiles_w) % g.tiles_h;
      const int tn_i = m_tile / (g.tiles_w * g.tiles_h);
      const int w_in = tw_i * g.TW * g.stride - g.pad_w;
      const int h_in = th_i * g.TH * g.stride - g.pad_h;
      const int n_in = tn_i * g.TN;
      const int co_half = n_tile * BLOCK_N + (int)rank * (BLOCK_N / 2);
      int tap = g.tap_base, kb = 0, r = 0, s = 0;
      for (int it = 0; it < k_iters; ++it) {
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (leader_lane) {
          if (rank == 0) ptx::mbar_expect_tx(&full_bar[stage], stage_tx);
          ptx::tma_load_4d_pair(smem_a + stage * kABytes, &tm_x, &full_bar[stage], kb * kBlockK, w_in + s * g.dil,
                                h_in + r * g.dil, n_in);
          ptx::tma_load_3d_pair(smem_b + stage * Cfg::kBHalfBytes, &tm_w, &full_bar[stage], kb * kBlockK, tap, co_half);
        }
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
        if (++kb == g.num_kb) {
          kb = 0;
          tap += g.tap_ss;
          if (++s == g.S) {
            s = 0;
            ++r;
            tap = g.tap_base + r * g.tap_sr;
          }
        }
      }
      if constexpr (kResMma) {
        const int res_chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
        for (int j = 0; j < res_chunks; ++j) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader_lane) {
            if (rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2u * (a_tx + 4096u));
            ptx::tma_load_4d_pair(smem_a + stage * kABytes, &tm_res, &full_bar[stage], n_tile * BLOCK_N + j * 64,
                                  tw_i * g.TW, th_i * g.TH, tn_i * g.TN);
            // this CTA's 32 rows of the 64 x 64 identity block
            ptx::tma_load_2d_pair(smem_b + stage * Cfg::kBHalfBytes, &tm_id, &full_bar[stage], 0, (int)rank * 32);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: the leader CTA only =====================
    if (rank == 0) {
      const bool leader_lane = ptx::elect_one();
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(2 * kBlockM, BLOCK_N, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int pt = pair; pt < pair_tiles; pt += num_pairs) {
        ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BLOCK_N);
        for (int it = 0; it < k_iters; ++it) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * kABytes), 16, 1024);
          const uint64_t b_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBHalfBytes), 16, 1024);
          if (leader_lane) {
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k)
              ptx::umma_bf16_pair(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc,
                                  (it > 0 || k > 0) ? 1u : 0u);
            ptx::umma_commit_pair(&empty_bar[stage]);  // frees the stage in both CTAs
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if constexpr (kResMma) {
          constexpr uint32_t idesc64 = ptx::umma_idesc_bf16(2 * kBlockM, 64, 0, 0);
          const int n_tile = pt % g.n_tiles;
          const int res_chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
          for (int j = 0; j < res_chunks; ++j) {
            ptx::mbar_wait(&full_bar[stage], phase);
            ptx::tc_fence_after();
            const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * kABytes), 16, 1024);
            const uint64_t b_desc =
                ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBHalfBytes), 16, 1024);
            if (leader_lane) {
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k)
                ptx::umma_bf16_pair(d_tmem + (uint32_t)(j * 64), a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k),
                                    idesc64, 1u);
              ptx::umma_commit_pair(&empty_bar[stage]);
            }
            if (++stage == kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        if (leader_lane) ptx::umma_commit_pair(&tmem_full[acc]);  // accumulator complete in both CTAs
        if (++acc == kAcc) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue warps (per CTA, as in the single-CTA kernel) =====================
    const int ewarp = warp - kEpiWarp0;
    const int group = ewarp >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int patch = g.TH * g.TW;
    const int dn = row / patch;
    const int rem = row - dn * patch;
    const int dh = rem / g.TW;
    const int dw = rem - dh * g.TW;
    const bool row_in_patch = row < g.TN * patch;
    const bool issuer = ((ewarp & 3) == 0 && lane == 0);
    const int bar_id = 1 + group;
    const uint32_t row_off = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const uint32_t tmem_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
    uint8_t* out_tile = smem_out + group * (kBlockM * 128);
    constexpr int kChunks = BLOCK_N / 64;
    float sacc[kChunks][4];
#pragma unroll
    for (int c = 0; c < kChunks; ++c) sacc[c][0] = sacc[c][1] = sacc[c][2] = sacc[c][3] = 0.f;
    int stat_n_tile = -1;
    auto flush_stats = [&](int nt) {
      if (issuer) ptx::tma_store_wait_read<0>();
      ptx::named_bar_sync(bar_id, 128);
      float* st = reinterpret_cast<float*>(out_tile);
#pragma unroll
      for (int c = 0; c < kChunks; ++c)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          st[((quarter * kChunks + c) * 4 + v) * 32 + lane] = sacc[c][v];
          sacc[c][v] = 0.f;
        }
      ptx::named_bar_sync(bar_id, 128);
      float* slot = stats + (size_t)blockIdx.x * 2 * g.K;
      for (int idx = quarter * 32 + lane; idx < kChunks * 128; idx += 128) {
        const int c = idx >> 7, v = (idx >> 5) & 3, ln = idx & 31;
        float tot = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) tot += st[((e * kChunks + c) * 4 + v) * 32 + ln];
        const int ch = nt * BLOCK_N + c * 64 + 2 * ln + (v >> 1);
        if (ch < g.K) atomicAdd(&slot[(size_t)(v & 1) * g.K + ch], tot);
      }
      ptx::named_bar_sync(bar_id, 128);
    };
    uint32_t tile_count = (uint32_t)group;
    for (int pt = pair + group * num_pairs; pt < pair_tiles; pt += 2 * num_pairs, tile_count += 2) {
      const int acc = (int)(tile_count % kAcc);
      const uint32_t acc_phase = (tile_count / kAcc) & 1u;
      const int n_tile = pt % g.n_tiles;
      const int m_tile = 2 * (pt / g.n_tiles) + (int)rank;
      const bool real_tile = m_tile < m_tiles;
      const int tw_i = m_tile % g.tiles_w;
      const int th_i = (m_tile / g.tiles_w) % g.tiles_h;
      const int tn_i = m_tile / (g.tiles_w * g.tiles_h);
      bool valid = true;
      size_t pix = 0;
      if constexpr (kFused || kStats || kResMma) {
        const int n = tn_i * g.TN + dn;
        const int p = th_i * g.TH + dh;
        const int q = tw_i * g.TW + dw;
        valid = real_tile && row_in_patch && n < g.N && p < g.P && q < g.Q;
        pix = ((size_t)n * g.P + p) * g.Q + q;
      }
      const int chunks = min(BLOCK_N / 64, (g.K - n_tile * BLOCK_N + 63) / 64);
      uint32_t vm = 0xffffffffu;
      if constexpr (kStats) {
        if (n_tile != stat_n_tile) {
          if (stat_n_tile >= 0) flush_stats(stat_n_tile);
          stat_n_tile = n_tile;
        }
        vm = __ballot_sync(0xffffffffu, valid);
      }
      ptx::mbar_wait(&tmem_full[acc], acc_phase);
      ptx::tc_fence_after();
#pragma unroll
      for (int cc = 0; cc < kChunks; ++cc) {
        if (cc >= chunks) break;
        const int co_chunk = n_tile * BLOCK_N + cc * 64;
        if (issuer) ptx::tma_store_wait_read<0>();
        ptx::named_bar_sync(bar_id, 128);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem_lane + (uint32_t)(acc * BLOCK_N + cc * 64 + half * 32), v);
          ptx::tmem_ld_wait();
          if (half == 1 && cc == chunks - 1) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_leader(&tmem_empty[acc]);  // the MMA warp lives in the leader CTA
          }
          const int co0 = co_chunk + half * 32;
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if constexpr (kFused) {
#pragma unroll
            for (int j8 = 0; j8 < 4; ++j8) {
              const int co = co0 + j8 * 8;
              if (co < g.K) {
                if (g.has_bias) {
                  const float4 b0 = *reinterpret_cast<const float4*>(bias + co);
                  const float4 b1 = *reinterpret_cast<const float4*>(bias + co + 4);
                  f[j8 * 8 + 0] += b0.x; f[j8 * 8 + 1] += b0.y; f[j8 * 8 + 2] += b0.z; f[j8 * 8 + 3] += b0.w;
                  f[j8 * 8 + 4] += b1.x; f[j8 * 8 + 5] += b1.y; f[j8 * 8 + 6] += b1.z; f[j8 * 8 + 7] += b1.w;
                }
                if (g.has_res && valid) {
                  float rr[8];
                  unpack8(ld8(residual + pix * g.K + co), rr);
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j8 * 8 + j] += rr[j];
                }
                if (g.relu) {
#pragma unroll
                  for (int j = 0; j < 8; ++j) f[j8 * 8 + j] = fmaxf(f[j8 * 8 + j], 0.f);
                }
              }
            }
          }
          if constexpr (kResMma) {
            if (g.has_mask && valid && co0 < g.K) {
              const uint8_t* mrow = reinterpret_cast<const uint8_t*>(residual) + pix * (size_t)(g.K >> 3) + (co0 >> 3);
              const uint32_t mbits = *reinterpret_cast<const uint32_t*>(mrow);
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = ((mbits >> j) & 1u) ? f[j] : 0.f;
            }
          }
          uint32_t packed[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
            packed[j] = *reinterpret_cast<const uint32_t*>(&h2);
          }
#pragma unroll
          for (int j8 = 0; j8 < 4; ++j8) {
            const uint32_t chunk16 = (uint32_t)(half * 4 + j8);
            *reinterpret_cast<uint4*>(out_tile + row_off + ((chunk16 ^ sw) << 4)) =
                make_uint4(packed[j8 * 4], packed[j8 * 4 + 1], packed[j8 * 4 + 2], packed[j8 * 4 + 3]);
          }
        }
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(bar_id, 128);
        if (issuer && real_tile) {
          ptx::tma_store_4d(&tm_y, out_tile, co_chunk, tw_i * g.TW, th_i * g.TH, tn_i * g.TN);
          ptx::tma_store_commit();
        }
        if constexpr (kStats) {
          const uint8_t* rows = out_tile + (uint32_t)(quarter * 32) * 128u + ((uint32_t)(lane & 3) << 2);
          const uint32_t c16 = (uint32_t)(lane >> 2);
          float s0 = 0.f, q0 = 0.f, s1 = 0.f, q1 = 0.f;
#pragma unroll
          for (int i0 = 0; i0 < 32; i0 += 16) {
            uint32_t wv[16];
#pragma unroll
            for (int i = 0; i < 16; ++i)
              wv[i] = *reinterpret_cast<const uint32_t*>(rows + (i0 + i) * 128 + ((c16 ^ (uint32_t)(i & 7)) << 4));
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const bool ok = (vm >> (i0 + i)) & 1u;
              const float f0 = ok ? __uint_as_float(wv[i] << 16) : 0.f;
              const float f1 = ok ? __uint_as_float(wv[i] & 0xffff0000u) : 0.f;
              s0 += f0;
              q0 = fmaf(f0, f0, q0);
              s1 += f1;
              q1 = fmaf(f1, f1, q1);
            }
          }
          sacc[cc][0] += s0;
          sacc[cc][1] += q0;
          sacc[cc][2] += s1;
          sacc[cc][3] += q1;
        }
      }
    }
    if constexpr (kStats) {
      if (stat_n_tile >= 0) flush_stats(stat_n_tile);
    }
    if (issuer) ptx::tma_store_wait_all<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // neither CTA may release TMEM / shared memory while its partner can still touch it
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------
// wgrad kernel:  dW[co, tap, ci] += sum_pixels dY[pixel, co] * X[pixel shifted by tap, ci]
// GEMM M = Cout (128), N = Cin (BLOCK_N), K = pixels (64 per stage).  Both operands are stored
// with the GEMM-K (pixel) index as the *row* of a 128B-swizzled tile, i.e. MN-major.
// ------------------------------------------------------------------------------------------
// A CTA owns MT x 128 output channels x BLOCK_N input channels of one tap, i.e. up to a 256x256 tile held in all
// 512 TMEM columns: the dY patch is fetched once for BLOCK_N input channels and the X patch once for MT*128 output
// channels, which halves the L2->SMEM bytes per FLOP compared with a 128x128 tile.
// TP > 1 ("taps per CTA", narrow layers only): the CTA computes TP filter taps at once.  The dY tile is fetched once
// and the B operand is TP differently shifted X tiles side by side, so ONE MMA of N = TP*BLOCK_N covers them all:
// 3x3 64->64 goes from nine N=64 problems that each re-read dY to three N=192 problems (wider, fewer MMAs per byte).
template <int MT, int BLOCK_N, int TP = 1>
struct WgradCfg {
  static_assert(TP == 1 || MT == 1, "taps-per-CTA packing is for single-M-tile layers");
  static_assert(BLOCK_N * TP <= 256, "one MMA covers at most 256 columns");
  static constexpr int kABytes = MT * 128 * 64 * 2;       // MT * 2 blocks of [64 px][64 co]
  static constexpr int kBBytes = TP * BLOCK_N * 64 * 2;   // TP * BLOCK_N/64 blocks of [64 px][64 ci]
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (192 * 1024) / kStageBytes;
  static constexpr int kTmemColsRaw = MT * BLOCK_N * TP;
  static constexpr int kTmemCols = kTmemColsRaw <= 32 ? 32 : kTmemColsRaw <= 64 ? 64 : kTmemColsRaw <= 128 ? 128
                                   : kTmemColsRaw <= 256 ? 256 : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

template <int MT, int BLOCK_N, int TP = 1>
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tm_dy, const __grid_constant__ CUtensorMap tm_x,
                  const Geometry g, float* __restrict__ dw, int m_tiles_co, int n_tiles_ci, int bulk_epi) {
  ptx::pdl_trigger();
  using Cfg = WgradCfg<MT, BLOCK_N, TP>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // work item
  int wi = blockIdx.y;
  const int ci_tile = wi % n_tiles_ci;
  wi /= n_tiles_ci;
  const int co_tile = wi % m_tiles_co;
  const int tap = (wi / m_tiles_co) * TP;  // first tap of this CTA's group
  const int ntaps = min(TP, g.R * g.S - tap);
  const int r = tap / g.S;
  const int s = tap - r * g.S;

  const int patches = g.tiles_n * g.tiles_h * g.tiles_w;
  const int my_iters = (patches - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (my_iters <= 0) return;  // uniform across the CTA

  // Pixel rows a TMA box does not cover must read as zero in BOTH operands for every stage.
  if (g.TN * g.TH * g.TW < 64 || g.wrow) {
    uint4 z = make_uint4(0, 0, 0, 0);
    uint4* p4 = reinterpret_cast<uint4*>(smem);
    const int n16 = kStages * Cfg::kStageBytes / 16;
    for (int i = threadIdx.x; i < n16; i += kThreads) p4[i] = z;
    ptx::fence_proxy_async_smem();
  }

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_dy);
    ptx::prefetch_tmap(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        ptx::mbar_init(&full_bar[i], 1);
        ptx::mbar_init(&empty_bar[i], 1);
      }
      ptx::mbar_init(tmem_full, 1);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  }
  ptx::pdl_wait();  // set-up above overlapped the previous kernel's tail
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const uint32_t px_bytes = (uint32_t)(g.TN * g.TH * g.TW) * 128u;  // one 64-channel box
  const bool wrow = TP == 3 && g.wrow != 0;
  // row mode loads only the 64-channel dY blocks that exist (the rest of the A tile stays at its initial zero)
  const int a_blocks = wrow ? min(2 * MT, (g.K - co_tile * (128 * MT) + 63) / 64) : 2 * MT;
  const uint32_t stage_tx = wrow ? px_bytes * (uint32_t)(a_blocks + 1)
                                 : px_bytes * (uint32_t)(2 * MT + ntaps * (BLOCK_N / 64));

  if (warp == 0) {
    if (lane == 0) {  // (the elected-lane form of conv_fprop_kernel measured slower here: few, long stages)
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_iters; ++i) {
        const int patch = (int)blockIdx.x + i * (int)gridDim.x;
        const int tw_i = patch % g.tiles_w;
        const int th_i = (patch / g.tiles_w) % g.tiles_h;
        const int tn_i = patch / (g.tiles_w * g.tiles_h);
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        ptx::mbar_expect_tx(&full_bar[stage], stage_tx);
        uint8_t* a = smem_a + stage * Cfg::kABytes;
        uint8_t* b = smem_b + stage * Cfg::kBBytes;
        if (wrow) {
          for (int h = 0; h < a_blocks; ++h)
            ptx::tma_load_4d(a + h * 8192, &tm_dy, &full_bar[stage], co_tile * (128 * MT) + h * 64, 0, th_i * g.TH, tn_i);
          ptx::tma_load_4d(b, &tm_x, &full_bar[stage], ci_tile * BLOCK_N, -g.pad_w, th_i * g.TH + r * g.dil - g.pad_h,
                           tn_i);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          continue;
        }
#pragma unroll
        for (int h = 0; h < 2 * MT; ++h)
          ptx::tma_load_4d(a + h * 8192, &tm_dy, &full_bar[stage], co_tile * (128 * MT) + h * 64, tw_i * g.TW,
                           th_i * g.TH, tn_i * g.TN);
        if (TP == 1) {
#pragma unroll
          for (int h = 0; h < BLOCK_N / 64; ++h)
            ptx::tma_load_4d(b + h * 8192, &tm_x, &full_bar[stage], ci_tile * BLOCK_N + h * 64,
                             tw_i * g.TW * g.stride + s * g.dil - g.pad_w,
                             th_i * g.TH * g.stride + r * g.dil - g.pad_h, tn_i * g.TN);
        } else {
          int rr = r, ss = s;
          for (int t = 0; t < ntaps; ++t) {  // one shifted X tile per tap, side by side along N
#pragma unroll
            for (int h = 0; h < BLOCK_N / 64; ++h)
              ptx::tma_load_4d(b + (t * (BLOCK_N / 64) + h) * 8192, &tm_x, &full_bar[stage], ci_tile * BLOCK_N + h * 64,
                               tw_i * g.TW * g.stride + ss * g.dil - g.pad_w,
                               th_i * g.TH * g.stride + rr * g.dil - g.pad_h, tn_i * g.TN);
            if (++ss == g.S) {
              ss = 0;
              ++rr;
            }
          }
        }
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, BLOCK_N * TP, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_iters; ++i) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        if (TP == 3 && g.wrow) {
          // one haloed X tile; tap s is the same tile one pixel row (128 B) further on
          const uint32_t b_addr = ptx::smem_u32(smem_b + stage * Cfg::kBBytes);
          const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * Cfg::kABytes), 8192, 1024);
          if (g.wrow == 1) {
            const uint64_t b_desc = ptx::umma_smem_desc_sw128(b_addr, 128, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_bf16(tmem_base, a_desc + (uint64_t)(128 * k), b_desc + (uint64_t)(128 * k), idesc,
                             (i > 0 || k > 0) ? 1u : 0u);
          } else {
            constexpr uint32_t idesc64 = ptx::umma_idesc_bf16(128, 64, 1, 1);
#pragma unroll
            for (int t = 0; t < 3; ++t) {
              const uint64_t b_desc = ptx::umma_smem_desc_sw128(b_addr + (uint32_t)t * 128u, 8192, 1024);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                ptx::umma_bf16(tmem_base + (uint32_t)(t * 64), a_desc + (uint64_t)(128 * k),
                               b_desc + (uint64_t)(128 * k), idesc64, (i > 0 || k > 0) ? 1u : 0u);
            }
          }
          ptx::umma_commit(&empty_bar[stage]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          continue;
        }
        // MN-major SW128: 64-element MN blocks 8192 B apart (LBO), 8-pixel-row groups 1024 B apart (SBO)
        const uint64_t b_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBBytes), 8192, 1024);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          const uint64_t a_desc =
              ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * Cfg::kABytes + mt * 16384), 8192, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // 16 pixel rows = 2048 bytes = +128 in 16-byte units
            ptx::umma_bf16(tmem_base + (uint32_t)(mt * BLOCK_N), a_desc + (uint64_t)(128 * k),
                           b_desc + (uint64_t)(128 * k), idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        ptx::umma_commit(&empty_bar[stage]);
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      ptx::umma_commit(tmem_full);
    }
  } else {
    const int quarter = warp & 3;
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    const int taps = g.R * g.S;
    if (bulk_epi) {
      // Every MMA has retired, so the pipeline buffers are free: each thread stages ITS accumulator row in shared
      // memory (row pitch 260 floats: conflict-free 16-byte stores) and hands whole (row, tap) segments -- contiguous
      // in dW -- to the copy engine as bulk reductions.  A per-thread red.global.add.v4 touches 32 different rows per
      // warp instruction, half a sector each; the bulk form adds full lines.
      constexpr int kPitch = 260;
      float* stg = reinterpret_cast<float*>(smem) + (size_t)(quarter * 32 + lane) * kPitch;
#pragma unroll 1
      for (int mt = 0; mt < MT; ++mt) {
        const int co = co_tile * (128 * MT) + mt * 128 + quarter * 32 + lane;
        if (mt > 0) ptx::tma_store_wait_read<0>();  // this thread's previous segments have left its staging row
#pragma unroll 1
        for (int c = 0; c < ntaps * (BLOCK_N / 32); ++c) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(mt * BLOCK_N * TP + c * 32), v);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<uint4*>(stg + c * 32 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        ptx::fence_proxy_async_smem();
        const int ci0 = ci_tile * BLOCK_N;
        const int ncols = min(BLOCK_N, g.C - ci0);
        if (co < g.K && ncols > 0) {
          for (int t = 0; t < ntaps; ++t)
            ptx::bulk_reduce_add_f32(dw + ((size_t)co * taps + tap + t) * g.C + ci0, stg + t * BLOCK_N,
                                     (uint32_t)ncols * 4u);
        }
        ptx::tma_store_commit();
      }
      ptx::tma_store_wait_all<0>();
    } else
#pragma unroll 1
    for (int mt = 0; mt < MT; ++mt) {
      const int co = co_tile * (128 * MT) + mt * 128 + quarter * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < ntaps * (BLOCK_N / 32); ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(mt * BLOCK_N * TP + c * 32), v);
        ptx::tmem_ld_wait();
        const int t_local = c / (BLOCK_N / 32);  // accumulator columns are [tap in group][input channel]
        const int ci0 = ci_tile * BLOCK_N + (c - t_local * (BLOCK_N / 32)) * 32;
        if (co < g.K) {
          float* dst = dw + ((size_t)co * taps + tap + t_local) * g.C + ci0;
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            if (ci0 + j < g.C)  // C is a multiple of 8, so a 4-wide group is all-in or all-out
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(v[j])),
                           "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])),
                           "f"(__uint_as_float(v[j + 3]))
                           : "memory");
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------
// wgrad on a CTA pair (cta_group::2): one 256 (Cout) x BLOCK_N (Cin) tile of one tap per PAIR.  Each CTA loads its own
// 128 output channels of dY and HALF of the X tile; the leader issues M = 256 MMAs that read both CTAs' shared memory.
// A single-CTA MMA is fed from shared memory at about (M + N) / 2 cycles per K = 16 step (measured: N = 64 -> 93 clk
// against a 32 clk floor, N = 256 -> ~190 against 128), so the 128 x 256 tiles of conv_wgrad_kernel<2, 256> cannot pass
// ~2/3 of the tensor rate; as a pair each SM fetches 128 + 128 rows per step, which is the floor.  It also halves the
// L2 -> shared-memory bytes of the X operand.  The tensor maps cover this GPU's memory only; nothing is multicast.
// ------------------------------------------------------------------------------------------
template <int BLOCK_N>
struct Wgrad2Cfg {
  static_assert(BLOCK_N == 128 || BLOCK_N == 256, "pair tiles are 256 x 128 or 256 x 256");
  static constexpr int kABytes = 128 * 64 * 2;              // this CTA's 2 blocks of [64 px][64 co]
  static constexpr int kBBytes = (BLOCK_N / 2) * 64 * 2;    // this CTA's BLOCK_N/128 blocks of [64 px][64 ci]
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = (192 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = BLOCK_N;                 // 128 lanes x BLOCK_N fp32 columns per CTA
  static constexpr int kDataBytes = kStages * kStageBytes < 136 * 1024 ? 136 * 1024 : kStages * kStageBytes;
  static constexpr int kSmemBytes = kDataBytes + 1024 + 256;  // (the bulk epilogue stages 128 rows x 1040 B)
};

template <int BLOCK_N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_wgrad2_kernel(const __grid_constant__ CUtensorMap tm_dy, const __grid_constant__ CUtensorMap tm_x,
                   const Geometry g, float* __restrict__ dw, int m_tiles_co, int n_tiles_ci, int bulk_epi) {
  ptx::pdl_trigger();
  using Cfg = Wgrad2Cfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * Cfg::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kDataBytes);
  uint64_t* full_bar = bars;             // used in the leader only: bytes of BOTH CTAs land on it
  uint64_t* empty_bar = bars + kStages;  // per CTA: the leader's commit frees the stage in both
  uint64_t* tmem_full = bars + 2 * kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();  // 0 = leader
  const int split = (int)blockIdx.x >> 1;
  const int nsplits = (int)gridDim.x >> 1;

  int wi = blockIdx.y;
  const int ci_tile = wi % n_tiles_ci;
  wi /= n_tiles_ci;
  const int co_tile = wi % m_tiles_co;
  const int tap = wi / m_tiles_co;
  const int r = tap / g.S;
  const int s = tap - r * g.S;

  const int patches = g.tiles_n * g.tiles_h * g.tiles_w;
  const int my_iters = (patches - split + nsplits - 1) / nsplits;  // the same in both CTAs of the pair (>= 1: host)

  if (g.TN * g.TH * g.TW < 64) {  // pixel rows a TMA box does not cover must read as zero in both operands
    uint4 z = make_uint4(0, 0, 0, 0);
    uint4* p4 = reinterpret_cast<uint4*>(smem);
    const int n16 = kStages * Cfg::kStageBytes / 16;
    for (int i = threadIdx.x; i < n16; i += kThreads) p4[i] = z;
    ptx::fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_dy);
    ptx::prefetch_tmap(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) {
        ptx::mbar_init(&full_bar[i], 1);
        ptx::mbar_init(&empty_bar[i], 1);
      }
      ptx::mbar_init(tmem_full, 1);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc_pair<Cfg::kTmemCols>(tmem_slot);
  }
  ptx::pdl_wait();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // both CTAs' barriers are initialised, both TMEM allocations are done, smem is zeroed
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const uint32_t px_bytes = (uint32_t)(g.TN * g.TH * g.TW) * 128u;  // one 64-channel box
  const uint32_t stage_tx = 2u * px_bytes * (uint32_t)(2 + BLOCK_N / 128);
  const int co_base = co_tile * 256 + (int)rank * 128;
  const int ci_base = ci_tile * BLOCK_N + (int)rank * (BLOCK_N / 2);

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_iters; ++i) {
        const int patch = split + i * nsplits;
        const int tw_i = patch % g.tiles_w;
        const int th_i = (patch / g.tiles_w) % g.tiles_h;
        const int tn_i = patch / (g.tiles_w * g.tiles_h);
        ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (rank == 0) ptx::mbar_expect_tx(&full_bar[stage], stage_tx);
        uint8_t* a = smem_a + stage * Cfg::kABytes;
        uint8_t* b = smem_b + stage * Cfg::kBBytes;
#pragma unroll
        for (int h = 0; h < 2; ++h)
          ptx::tma_load_4d_pair(a + h * 8192, &tm_dy, &full_bar[stage], co_base + h * 64, tw_i * g.TW, th_i * g.TH,
                                tn_i * g.TN);
#pragma unroll
        for (int h = 0; h < BLOCK_N / 128; ++h)
          ptx::tma_load_4d_pair(b + h * 8192, &tm_x, &full_bar[stage], ci_base + h * 64,
                                tw_i * g.TW * g.stride + s * g.dil - g.pad_w,
                                th_i * g.TH * g.stride + r * g.dil - g.pad_h, tn_i * g.TN);
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(256, BLOCK_N, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_iters; ++i) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        const uint64_t a_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_a + stage * Cfg::kABytes), 8192, 1024);
        const uint64_t b_desc = ptx::umma_smem_desc_sw128(ptx::smem_u32(smem_b + stage * Cfg::kBBytes), 8192, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16_pair(tmem_base, a_desc + (uint64_t)(128 * k), b_desc + (uint64_t)(128 * k), idesc,
                              (i > 0 || k > 0) ? 1u : 0u);
        ptx::umma_commit_pair(&empty_bar[stage]);  // frees the stage in both CTAs
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      ptx::umma_commit_pair(tmem_full);  // accumulators complete in both CTAs
    }
  } else {
    const int quarter = warp & 3;
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    const int taps = g.R * g.S;
    const int co = co_base + quarter * 32 + lane;
    const int ci0 = ci_tile * BLOCK_N;
    const int ncols = min(BLOCK_N, g.C - ci0);
    if (bulk_epi) {  // as in conv_wgrad_kernel: stage this thread's row, hand it to the copy engine as one reduction
      constexpr int kPitch = 260;
      float* stg = reinterpret_cast<float*>(smem) + (size_t)(quarter * 32 + lane) * kPitch;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(c * 32), v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<uint4*>(stg + c * 32 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
      ptx::fence_proxy_async_smem();
      if (co < g.K && ncols > 0)
        ptx::bulk_reduce_add_f32(dw + ((size_t)co * taps + tap) * g.C + ci0, stg, (uint32_t)ncols * 4u);
      ptx::tma_store_commit();
      ptx::tma_store_wait_all<0>();
    } else {
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(c * 32), v);
        ptx::tmem_ld_wait();
        if (co < g.K) {
          float* dst = dw + ((size_t)co * taps + tap) * g.C + ci0 + c * 32;
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            if (ci0 + c * 32 + j < g.C)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(v[j])),
                           "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])),
                           "f"(__uint_as_float(v[j + 3]))
                           : "memory");
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync_all();  // neither CTA may release TMEM / shared memory while its partner's MMAs can still touch it
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------
// weight re-layout:  fp32 master [K, R*S, C]  ->  bf16 fprop copy [K, R*S, C]
//                                              and bf16 dgrad copy [C, R*S (flipped), K]
// ------------------------------------------------------------------------------------------
__global__ void weight_prep_kernel(const float* __restrict__ w, bf16* __restrict__ w_fprop,
                                   bf16* __restrict__ w_dgrad, int K, int taps, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const size_t total = (size_t)K * taps * C;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int t = (int)((i / C) % taps);
    const int k = (int)(i / ((size_t)C * taps));
    const bf16 v = __float2bfloat16(w[i]);
    if (w_fprop) w_fprop[i] = v;
    if (w_dgrad) w_dgrad[((size_t)c * taps + (taps - 1 - t)) * K + k] = v;
  }
}

// Same result through a 32x32 shared-memory transpose per (tap, k-tile, c-tile): reads are coalesced along C and the
// flipped/transposed dgrad copy is written coalesced along K (the plain kernel scatters 2-byte stores K*taps apart,
// which made this 6 % of the reference network's step).  blockDim = (32, 8), grid = (c-tiles, k-tiles, taps).
__global__ void weight_prep_tiled_kernel(const float* __restrict__ w, bf16* __restrict__ w_fprop,
                                         bf16* __restrict__ w_dgrad, int K, int taps, int C) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  __shared__ float tile[32][33];
  const int t = blockIdx.z;
  const int c0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int k = k0 + j, c = c0 + threadIdx.x;
    if (k < K && c < C) {
      const size_t i = ((size_t)k * taps + t) * C + c;
      const float v = w[i];
      tile[j][threadIdx.x] = v;
      if (w_fprop) w_fprop[i] = __float2bfloat16(v);
    }
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, k = k0 + threadIdx.x;
    if (k < K && c < C) w_dgrad[((size_t)c * taps + (taps - 1 - t)) * K + k] = __float2bfloat16(tile[threadIdx.x][j]);
  }
}

// All layers of a network in ONE launch: block b serves tile (b - block_start[l]) of layer l (binary search over the
// prefix sums).  A training step otherwise spends one ~5 us launch per convolution on this re-layout (60 launches =
// 4.5 % of the reference network's step); the table is built once, the weights live at fixed addresses.
struct WPrepDesc {
  const float* w;
  bf16* wf;
  bf16* wd;
  int K, taps, C, ctiles, ktiles;
  int pad_;
};

__global__ void __launch_bounds__(256)
weight_prep_batched_kernel(const WPrepDesc* __restrict__ descs, const int* __restrict__ block_start, int nlayers) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  // One 64 (Cout) x 64 (Cin) tile of one tap per block, two elements per access on both sides: 8-byte loads of the
  // fp32 master, 4-byte bf16x2 stores of the forward copy, and after the transpose through shared memory 4-byte stores
  // of the data-gradient copy (128 contiguous bytes per warp instruction everywhere).  The 32 x 32 scalar version
  // moved 2-byte elements: 92 us for ResNet-50's 100 MB of weights against ~30 us of traffic.
  __shared__ __nv_bfloat16 tile[64][66];
  int lo = 0, hi = nlayers - 1;
  const int b = blockIdx.x;
  while (lo < hi) {  // last layer whose first block is <= b
    const int mid = (lo + hi + 1) >> 1;
    if (block_start[mid] <= b) lo = mid;
    else hi = mid - 1;
  }
  const WPrepDesc d = descs[lo];
  int local = b - block_start[lo];
  const int ct = local % d.ctiles;
  local /= d.ctiles;
  const int kt = local % d.ktiles;
  const int t = local / d.ktiles;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = ct * 64, k0 = kt * 64;
  for (int j = ty; j < 64; j += 8) {
    const int k = k0 + j, c = c0 + 2 * tx;
    if (k < d.K && c < d.C) {  // C and K are even: a pair is all-in or all-out
      const size_t i = ((size_t)k * d.taps + t) * d.C + c;
      const float2 v = *reinterpret_cast<const float2*>(d.w + i);
      const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
      *reinterpret_cast<__nv_bfloat162*>(&tile[j][2 * tx]) = h;
      *reinterpret_cast<__nv_bfloat162*>(d.wf + i) = h;
    }
  }
  if (d.wd == nullptr) return;  // uniform across the block
  __syncthreads();
  for (int j = ty; j < 64; j += 8) {
    const int c = c0 + j, k = k0 + 2 * tx;
    if (k < d.K && c < d.C) {
      __nv_bfloat162 h;
      h.x = tile[2 * tx][j];
      h.y = tile[2 * tx + 1][j];
      *reinterpret_cast<__nv_bfloat162*>(d.wd + ((size_t)c * d.taps + (d.taps - 1 - t)) * d.K + k) = h;
    }
  }
}

// ------------------------------------------------------------------------------------------
// first-layer ("stem") helpers
// ------------------------------------------------------------------------------------------
// x [N,H,W,C] bf16 (C <= 4)  ->  planes [N][H2][2][W2][8] bf16, packed channel cp = (dy*2 + dx)*C + c holds
// x[2*h2 + dy][2*w2 + dx][c]; lanes >= 4*C and pixels beyond an odd H/W are zero.
__global__ void stem_pack_kernel(const bf16* __restrict__ x, bf16* __restrict__ planes, int N, int H, int W, int C,
                                 int H2, int W2) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)N * H2 * 2 * W2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int w2 = (int)(i % W2);
    const int plane = (int)((i / W2) & 1);
    const int64_t nh = i / (2 * W2);
    const int h2 = (int)(nh % H2);
    const int n = (int)(nh / H2);
    bf16x8 out;
    bf16* o = reinterpret_cast<bf16*>(&out);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int cp = plane * 8 + j;
      bf16 v = __float2bfloat16(0.f);
      if (cp < 4 * C) {
        const int blk = cp / C;
        const int c = cp - blk * C;
        const int h = 2 * h2 + (blk >> 1);
        const int w = 2 * w2 + (blk & 1);
        if (h < H && w < W) v = x[(((int64_t)n * H + h) * W + w) * C + c];
      }
      o[j] = v;
    }
    st8(planes + i * 8, out);
  }
}

// RGB images (C = 3, even H and W): one thread builds BOTH planes of a 2x2 pixel block from two 12-byte reads (three
// aligned 32-bit words per image row) and writes two 16-byte vectors.  The generic kernel above does eight scalar loads
// with a division each per vector: 110 us for 256 x 224 x 224 against ~30 us of traffic.
__global__ void __launch_bounds__(256)
stem_pack_c3_kernel(const bf16* __restrict__ x, bf16* __restrict__ planes, int N, int H, int W, int H2, int W2) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int64_t total = (int64_t)N * H2 * W2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int w2 = (int)(i % W2);
    const int64_t nh = i / W2;  // n * H2 + h2
    const int h2 = (int)(nh % H2);
    const int n = (int)(nh / H2);
    // pixels (2h2, 2w2), (2h2, 2w2+1) = 6 bf16 = words a0..a2; the row below = b0..b2
    const uint32_t* r0 = reinterpret_cast<const uint32_t*>(x + (((int64_t)n * H + 2 * h2) * W + 2 * w2) * 3);
    const uint32_t* r1 = reinterpret_cast<const uint32_t*>(x + (((int64_t)n * H + 2 * h2 + 1) * W + 2 * w2) * 3);
    const uint32_t a0 = __ldg(r0), a1 = __ldg(r0 + 1), a2 = __ldg(r0 + 2);
    const uint32_t b0 = __ldg(r1), b1 = __ldg(r1 + 1), b2 = __ldg(r1 + 2);
    // packed channel cp = block * 3 + c, block = dy * 2 + dx: plane 0 = cp 0..7, plane 1 = cp 8..11 then zeros
    bf16x8 p0, p1;
    p0.raw = make_uint4(a0, a1, a2, b0);
    p1.raw = make_uint4(b1, b2, 0u, 0u);
    bf16* o = planes + (nh * 2 * W2 + w2) * 8;
    st8(o, p0);
    st8(o + (int64_t)W2 * 8, p1);
  }
}

// fp32 master [K, R, S, C]  ->  bf16 [K][Th*Tw][16] for the space-to-depth problem: packed tap (r2, s2), channel
// cp = (dy*2+dx)*C + c  <-  original tap r = 2*(r2 - pad2_h) + dy + pad_h (same for s), zero when outside the filter.
__global__ void stem_weight_prep_kernel(const float* __restrict__ w, bf16* __restrict__ out, int K, int R, int S, int C,
                                        int Th, int Tw, int pad_h, int pad_w, int pad2_h, int pad2_w) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int total = K * Th * Tw * 16;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int cp = i & 15;
    const int t = (i >> 4) % (Th * Tw);
    const int k = (i >> 4) / (Th * Tw);
    float v = 0.f;
    if (cp < 4 * C) {
      const int blk = cp / C;
      const int c = cp - blk * C;
      const int r = 2 * (t / Tw - pad2_h) + (blk >> 1) + pad_h;
      const int sx = 2 * (t % Tw - pad2_w) + (blk & 1) + pad_w;
      if (r >= 0 && r < R && sx >= 0 && sx < S) v = w[(((size_t)k * R + r) * S + sx) * C + c];
    }
    out[i] = __float2bfloat16(v);
  }
}

template <bool kStats>
static int launch_stem(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& ty, const Geometry& g,
                       float* stats, int* nparts, cudaStream_t st) {
  using Cfg = FpropCfg<64>;
  static_assert(Cfg::kStages * Cfg::kBBytes >= 4 * 8192, "resident stem weights need 32 KB of the B region");
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(conv_fprop_kernel<64, false, kStats, true>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  const int total = g.tiles_n * g.tiles_h * g.tiles_w;
  const int grid = total < tdl::num_sms() ? total : tdl::num_sms();
  if (nparts) *nparts = grid;
  {
    cudaError_t _le = tdl::launch_pdl(conv_fprop_kernel<64, false, kStats, true>, dim3(grid), dim3(kFpropThreads),
        Cfg::kSmemBytes, st, tx, tw, ty, ty, ty, ty, g, nullptr, nullptr, stats);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "conv_fprop_kernel");
  }
  TDL_LAUNCH_OK("conv_fprop_kernel(stem)");
  return 0;
}

// ------------------------------------------------------------------------------------------
// first-layer weight gradient on the packed image:
//   dWp[co][tap (r2, s2)][cp] = sum over (n, p, q) dY[n, p, q, co] * planes[n, p - pad2_h + r2, cp / 8, q - pad2_w + s2, cp % 8]
// GEMM M = Cout (128 rows, the upper 64 a resident zero block), K = the pixels of ONE output row (16 per MMA),
// N = packed taps x channels.  The dY row is a 128B-swizzled MN-major A tile exactly as in conv_wgrad_kernel.  The
// haloed image patch [taps_h][plane][TWB + taps_w - 1 pixels][8 ch] sits un-swizzled in shared memory, one pixel per
// 16-byte core-matrix row, so as an MN-major "interleave" B operand its 8-channel groups may start at ANY pixel:
// with a group stride (SBO) of ONE pixel the taps_w column taps of a (row tap, plane) pair become one MMA of
// N = 8 * taps_w whose column groups overlap in shared memory.  No im2col, each byte of dY and of the image is
// fetched once; the packed gradient stays in TMEM for the whole kernel and leaves through red.global.add.
//   variant 0: M = 128 (upper half zero)      variant 1: M = 64
// ------------------------------------------------------------------------------------------
struct StemWgradGeo {
  int N, P, Q, K, th, tw, pad2_h, pad2_w;
  int TWB, tiles_w, sp_plane, sp_row;
  int a_block, b_off, stage_bytes, stages, tx_bytes;
  int ncols, nblocks, variant;
};

template <int NBLK>  // NBLK = 2 * row taps
__global__ void __launch_bounds__(kThreads, 1)
stem_wgrad_kernel(const __grid_constant__ CUtensorMap tm_dy, const __grid_constant__ CUtensorMap tm_x,
                  const StemWgradGeo g, float* __restrict__ dwp) {
  ptx::pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + g.stages * g.stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + g.stages;
  uint64_t* tmem_full = bars + 2 * g.stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = g.N * g.P * g.tiles_w;
  const int my_iters = (total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (my_iters <= 0) return;

  {  // the upper 64 rows of A and the slack behind each patch must read as zero for every stage
    uint4 z = make_uint4(0, 0, 0, 0);
    uint4* p4 = reinterpret_cast<uint4*>(smem);
    const int n16 = g.stages * g.stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kThreads) p4[i] = z;
    ptx::fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_dy);
    ptx::prefetch_tmap(&tm_x);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < g.stages; ++i) {
        ptx::mbar_init(&full_bar[i], 1);
        ptx::mbar_init(&empty_bar[i], 1);
      }
      ptx::mbar_init(tmem_full, 1);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc<256>(tmem_slot);
  }
  ptx::pdl_wait();  // set-up above overlapped the previous kernel's tail
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Both issuing warps walk their schedule as a whole warp and one elected lane issues (see conv_fprop_kernel).
  if (warp == 0) {
    const bool leader = ptx::elect_one();
    int stage = 0;
    uint32_t phase = 0;
    for (int i = 0; i < my_iters; ++i) {
      const int tile = (int)blockIdx.x + i * (int)gridDim.x;
      const int tw_i = tile % g.tiles_w;
      const int p = (tile / g.tiles_w) % g.P;
      const int n = tile / (g.tiles_w * g.P);
      ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* a = smem + stage * g.stage_bytes;
      if (leader) {
        ptx::mbar_expect_tx(&full_bar[stage], (uint32_t)g.tx_bytes);
        ptx::tma_load_4d(a, &tm_dy, &full_bar[stage], 0, tw_i * g.TWB, p, n);
        ptx::tma_load_4d(a + g.b_off, &tm_x, &full_bar[stage], (tw_i * g.TWB - g.pad2_w) * 2, 0, p - g.pad2_h, n);
      }
      if (++stage == g.stages) {
        stage = 0;
        phase ^= 1;
      }
    }
  } else if (warp == 1) {
    const bool leader = ptx::elect_one();
    {
      // variant 1: M = 64 (no zero upper half to read; rows land in lanes 0..15 of each TMEM quarter)
      const uint32_t idesc = ptx::umma_idesc_bf16(g.variant == 1 ? 64 : 128, g.ncols, 1, 1);
      const int ksteps = g.TWB >> 4;
      // Everything an MMA needs is a 64-bit add away: block offsets (16-byte units) and TMEM columns are hoisted, so
      // the single issuing thread spends a handful of instructions per tcgen05.mma, not a descriptor rebuild.
      uint32_t boff[NBLK];
#pragma unroll
      for (int blk = 0; blk < NBLK; ++blk) boff[blk] = (uint32_t)((blk >> 1) * g.sp_row + (blk & 1) * g.sp_plane) >> 4;
      const uint64_t b_desc0 = ptx::umma_smem_desc_nosw(0, 128, 16);  // K groups 128 B apart, column-tap groups 16 B
      const uint64_t a_desc0 = ptx::umma_smem_desc_sw128(0, (uint32_t)g.a_block, 1024);
      const uint32_t ncols = (uint32_t)g.ncols;
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < my_iters; ++i) {
        ptx::mbar_wait(&full_bar[stage], phase);
        ptx::tc_fence_after();
        const uint32_t a_addr = ptx::smem_u32(smem + stage * g.stage_bytes);
        uint64_t a_k = a_desc0 + (uint64_t)(a_addr >> 4);
        uint64_t b_k = b_desc0 + (uint64_t)((a_addr + (uint32_t)g.b_off) >> 4);
        uint32_t acc = i > 0 ? 1u : 0u;
        for (int ks = 0; ks < ksteps; ++ks) {
          if (leader) {
#pragma unroll
            for (int blk = 0; blk < NBLK; ++blk)  // blk = r2 * 2 + plane
              ptx::umma_bf16(tmem_base + (uint32_t)blk * ncols, a_k, b_k + boff[blk], idesc, acc);
          }
          a_k += 128;  // 16 pixel rows of 128 bytes
          b_k += 16;   // 16 pixels of 16 bytes
          acc = 1u;
        }
        if (leader) ptx::umma_commit(&empty_bar[stage]);
        if (++stage == g.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (leader) ptx::umma_commit(tmem_full);
    }
  } else {
    const int quarter = warp & 3;
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    const int co = g.variant == 1 ? (lane < 16 ? quarter * 16 + lane : g.K) : quarter * 32 + lane;
    const int T = g.th * g.tw;
    const int total_cols = g.nblocks * g.ncols;
#pragma unroll 1
    for (int c = 0; c < total_cols; c += 32) {
      uint32_t v[32];
      ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c, v);
      ptx::tmem_ld_wait();
      if (co < g.K) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int col = c + j;
          const int blk = col / g.ncols, within = col - blk * g.ncols;
          const int s2 = within >> 3;
          if (s2 >= g.tw) continue;  // columns added to round N up to a multiple of 16
          const int t = (blk >> 1) * g.tw + s2;
          const int cp = (blk & 1) * 8 + (within & 7);
          float* dst = dwp + ((size_t)co * T + t) * 16 + cp;
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(__uint_as_float(v[j])),
                       "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                       : "memory");
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<256>(tmem_base);
  }
}

// dWp [K][Th*Tw][16] (packed problem)  ->  dW [K, R, S, C] fp32 (inverse of stem_weight_prep_kernel's index map)
__global__ void stem_wgrad_unpack_kernel(const float* __restrict__ dwp, float* __restrict__ dw, int K, int R, int S,
                                         int C, int Th, int Tw, int pad_h, int pad_w, int pad2_h, int pad2_w,
                                         int accumulate) {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  const int total = K * R * S * C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i % C;
    const int s = (i / C) % S;
    const int r = (i / (C * S)) % R;
    const int k = i / (C * S * R);
    const int rr = r - pad_h, ss = s - pad_w;
    const int rq = rr >= 0 ? rr / 2 : -((1 - rr) / 2), sq = ss >= 0 ? ss / 2 : -((1 - ss) / 2);  // floor division
    const int cp = ((rr - 2 * rq) * 2 + (ss - 2 * sq)) * C + c;
    const float v = dwp[((size_t)k * Th * Tw + (rq + pad2_h) * Tw + (sq + pad2_w)) * 16 + cp];
    dw[i] = accumulate ? dw[i] + v : v;
  }
}

template <int BLOCK_N, bool kFused, bool kStats, bool kResMma = false, bool kBwdStats = false>
static int launch_fprop_t(
const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& ty, const CUtensorMap& tres,
                          const CUtensorMap& tid, const CUtensorMap& tbx, const Geometry& g, const float* bias, const bf16* res, float* stats,
                          int* nparts, cudaStream_t st) {
  using Cfg = FpropCfg<BLOCK_N, kBwdStats>;
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(conv_fprop_kernel<BLOCK_N, kFused, kStats, false, kResMma, kBwdStats>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  const int total = g.tiles_n * g.tiles_h * g.tiles_w * g.n_tiles;
  const int grid = total < tdl::num_sms() ? total : tdl::num_sms();
  if (nparts) *nparts = grid;
  {
    cudaError_t _le = tdl::launch_pdl(conv_fprop_kernel<BLOCK_N, kFused, kStats, false, kResMma, kBwdStats>,
        dim3(grid), dim3(kFpropThreads), Cfg::kSmemBytes, st, tx, tw, ty, tres, tid, tbx, g, bias, res, stats);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "conv_fprop_kernel");
  }
  TDL_LAUNCH_OK("conv_fprop_kernel");
  return 0;
}

template <int BLOCK_N>
static int launch_fprop(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& ty, const CUtensorMap& tres,
                        const CUtensorMap& tid, const CUtensorMap& tbx, const Geometry& g, const float* bias, const bf16* res, float* stats,
                        int* nparts, cudaStream_t st) {
  if (g.bn_x) {  // data gradient that also reduces the BatchNorm-backward sums of the tensor it produces
    if (g.res_mma)
      return launch_fprop_t<BLOCK_N, false, false, true, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
    return launch_fprop_t<BLOCK_N, false, false, false, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
  }
  if (g.res_mma) {
    if (stats)  // + BatchNorm statistics of the stored (bias / residual / ReLU applied) output
      return launch_fprop_t<BLOCK_N, true, true, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
    if (g.has_bias || g.relu)
      return launch_fprop_t<BLOCK_N, true, false, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
    return launch_fprop_t<BLOCK_N, false, false, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
  }
  const bool fused = g.has_bias || g.has_res || g.relu;
  if (stats) {
    if (fused) return launch_fprop_t<BLOCK_N, true, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
    return launch_fprop_t<BLOCK_N, false, true>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
  }
  if (fused) return launch_fprop_t<BLOCK_N, true, false>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
  return launch_fprop_t<BLOCK_N, false, false>(tx, tw, ty, tres, tid, tbx, g, bias, res, stats, nparts, st);
}

template <int BLOCK_N, bool kFused, bool kStats, bool kResMma>
static int launch_fprop2_t(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& ty, const CUtensorMap& tres,
                           const CUtensorMap& tid, const Geometry& g, const float* bias, const bf16* res, float* stats,
                           int* nparts, cudaStream_t st) {
  using Cfg = Fprop2Cfg<BLOCK_N>;
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(conv_fprop2_kernel<BLOCK_N, kFused, kStats, kResMma>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  const int m_tiles = g.tiles_n * g.tiles_h * g.tiles_w;
  const int pair_tiles = ((m_tiles + 1) / 2) * g.n_tiles;
  const int max_pairs = tdl::num_sms() / 2;
  const int pairs = pair_tiles < max_pairs ? pair_tiles : max_pairs;
  const int grid = 2 * pairs;  // clusters of two CTAs (compile-time __cluster_dims__)
  if (nparts) *nparts = grid;
  {
    cudaError_t _le = tdl::launch_pdl(conv_fprop2_kernel<BLOCK_N, kFused, kStats, kResMma>, dim3(grid),
                                      dim3(kFpropThreads), Cfg::kSmemBytes, st, tx, tw, ty, tres, tid, g, bias, res,
                                      stats);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "conv_fprop2_kernel");
  }
  TDL_LAUNCH_OK("conv_fprop2_kernel");
  return 0;
}

// CTA-pair variants (see conv_fprop2_kernel); same dispatch as launch_fprop without the BatchNorm-backward form
template <int BLOCK_N>
static int launch_fprop2(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& ty, const CUtensorMap& tres,
                         const CUtensorMap& tid, const Geometry& g, const float* bias, const bf16* res, float* stats,
                         int* nparts, cudaStream_t st) {
  if (g.res_mma) {
    if (stats) return launch_fprop2_t<BLOCK_N, true, true, true>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
    if (g.has_bias || g.relu)
      return launch_fprop2_t<BLOCK_N, true, false, true>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
    return launch_fprop2_t<BLOCK_N, false, false, true>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
  }
  const bool fused = g.has_bias || g.has_res || g.relu;
  if (stats) {
    if (fused) return launch_fprop2_t<BLOCK_N, true, true, false>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
    return launch_fprop2_t<BLOCK_N, false, true, false>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
  }
  if (fused) return launch_fprop2_t<BLOCK_N, true, false, false>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
  return launch_fprop2_t<BLOCK_N, false, false, false>(tx, tw, ty, tres, tid, g, bias, res, stats, nparts, st);
}

// 64x64 bf16 identity block in global memory (the B operand of the residual-on-the-tensor-core iterations)
__device__ __align__(1024) bf16 g_identity64[64 * 64];
__global__ void identity64_init_kernel() {
  ptx::pdl_trigger();
  ptx::pdl_wait();
  for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) g_identity64[i] = __float2bfloat16((i >> 6) == (i & 63) ? 1.f : 0.f);
}
static int identity64_address(const void** out, cudaStream_t st) {
  constexpr int kMaxDev = 64;
  static void* addr[kMaxDev] = {nullptr};
  int dev = 0;
  TDL_CUDA_OK(cudaGetDevice(&dev));
  TDL_REQUIRE(dev >= 0 && dev < kMaxDev, "device ordinal out of range");
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    TDL_CUDA_OK(cudaGetSymbolAddress(&p, g_identity64));
    {
      cudaError_t _le = tdl::launch_pdl(identity64_init_kernel, dim3(1), dim3(256), 0, st);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "identity64_init_kernel");
    }  // ordered before the first consumer on the same stream
    TDL_LAUNCH_OK("identity64_init_kernel");
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}

// wgrad epilogue: 1 = bulk reductions from a shared-memory staging tile, 0 = per-thread red.global.add.v4.f32
static int g_wgrad_bulk = -1;
static int g_wgrad_snap = -1;         // whole-wave grids (see launch_wgrad)
static float g_wgrad_model_k = -1.f;  // constant of the split-count cost model (see launch_wgrad)

template <int MT, int BLOCK_N, int TP = 1>
static int launch_wgrad(const CUtensorMap& tdy, const CUtensorMap& tx, const Geometry& g, float* dw, cudaStream_t st) {
  using Cfg = WgradCfg<MT, BLOCK_N, TP>;
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(conv_wgrad_kernel<MT, BLOCK_N, TP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     Cfg::kSmemBytes));
    attr_set = true;
  }
  if (g_wgrad_bulk < 0) {
    const char* e = getenv("TDL_WGRAD_BULK");
    g_wgrad_bulk = e ? atoi(e) : 1;
  }
  const int m_tiles = (int)tdl::cdiv(g.K, 128 * MT);
  const int n_tiles = (int)tdl::cdiv(g.C, BLOCK_N);
  const int items = m_tiles * n_tiles * (int)tdl::cdiv(g.R * g.S, TP);
  const int patches = g.tiles_n * g.tiles_h * g.tiles_w;
  // split the pixel patches so the grid is a whole number of waves (1 CTA per SM) with >= 8 patches per CTA
  const int sms = tdl::num_sms();
  // Every CTA ends with a full-tile fp32 reduction into dW, so few long CTAs beat many short ones: aim for one
  // wave, take a second only while each CTA still keeps >= 32 pixel patches.
  int splits = sms / items;
  if (splits < 1) splits = 1;
  if ((2 * sms) / items > splits && patches / ((2 * sms) / items) >= 32) splits = (2 * sms) / items;
  if (splits > patches) splits = patches;
  {
    // Split count from a two-term cost model (measured on B200, profiles/README.md): a CTA spends
    // patches/splits x t_patch in its main loop and then adds its whole fp32 tile into dW with red.global.add, and
    // those adds (all CTAs finish together) drain at ~0.55 TB/s.  Both terms scale with the tile size, so the
    // optimum depends only on patches / items: splits* = sqrt(1.1 * patches / items).  More splits than that just
    // moves time from the tensor core into atomics (13x13 / 7x7 layers: 9 -> 2-4 splits, -10 % step time on the
    // reference network); fewer, longer CTAs also leave SMs to the main stream the wgrad overlaps with.
    float& model_k = g_wgrad_model_k;  // TDL_WGRAD_SPLIT_MODEL: the constant of the model (0 = off)
    if (model_k < 0.f) {
      const char* e = getenv("TDL_WGRAD_SPLIT_MODEL");
      model_k = e ? (float)atof(e) : 1.1f;
    }
    if (model_k > 0.f) {
      int s_model = (int)(sqrtf(model_k * (float)patches / (float)items) + 0.5f);
      if (s_model < 1) s_model = 1;
      if (splits > s_model) splits = s_model;
    }
  }
  if (splits < 1) splits = 1;
  {
    // One CTA per SM (the pipeline takes the whole shared memory): a grid of 1.5 waves costs two.  Snap a grid that
    // ends in a partial wave down to whole waves (3x3 64 -> 64 @56x56: 219 CTAs = 1.48 waves -> 147).
    int& snap = g_wgrad_snap;  // TDL_WGRAD_WAVE_SNAP=0 turns it off
    if (snap < 0) {
      const char* e = getenv("TDL_WGRAD_WAVE_SNAP");
      snap = e ? atoi(e) : 1;
    }
    const int ctas = splits * items;
    if (snap && ctas > sms && ctas % sms != 0) {
      const int whole = (ctas / sms) * sms / items;
      if (whole >= 1) splits = whole;
    }
  }
  dim3 grid(splits, items);
  {
    cudaError_t _le = tdl::launch_pdl(conv_wgrad_kernel<MT, BLOCK_N, TP>, dim3(grid), dim3(kThreads), Cfg::kSmemBytes,
        st, tdy, tx, g, dw, m_tiles, n_tiles,
        (g_wgrad_bulk != 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0) ? 1 : 0);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "conv_wgrad_kernel");
  }
  TDL_LAUNCH_OK("conv_wgrad_kernel");
  return 0;
}

// CTA-pair weight gradient (conv_wgrad2_kernel): -1 = read TDL_WGRAD_2CTA on first use, 0 = off, 1 = on
static int g_wgrad_pair = -1;

template <int BLOCK_N>
static int launch_wgrad2(const CUtensorMap& tdy, const CUtensorMap& tx, const Geometry& g, float* dw, cudaStream_t st) {
  using Cfg = Wgrad2Cfg<BLOCK_N>;
  static bool attr_set = false;
  if (!attr_set) {
    TDL_CUDA_OK(cudaFuncSetAttribute(conv_wgrad2_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     Cfg::kSmemBytes));
    attr_set = true;
  }
  if (g_wgrad_bulk < 0) {
    const char* e = getenv("TDL_WGRAD_BULK");
    g_wgrad_bulk = e ? atoi(e) : 1;
  }
  if (g_wgrad_model_k < 0.f) {
    const char* e = getenv("TDL_WGRAD_SPLIT_MODEL");
    g_wgrad_model_k = e ? (float)atof(e) : 1.1f;
  }
  const int m_tiles = (int)tdl::cdiv(g.K, 256);
  const int n_tiles = (int)tdl::cdiv(g.C, BLOCK_N);
  const int items = m_tiles * n_tiles * g.R * g.S;
  const int patches = g.tiles_n * g.tiles_h * g.tiles_w;
  const int pairs = tdl::num_sms() / 2;  // one CTA per SM, two SMs per tile
  int splits = pairs / items;
  if (splits < 1) splits = 1;
  if ((2 * pairs) / items > splits && patches / ((2 * pairs) / items) >= 32) splits = (2 * pairs) / items;
  if (g_wgrad_model_k > 0.f) {  // same two-term model as launch_wgrad (main loop vs the tile's reduction into dW)
    int s_model = (int)(sqrtf(g_wgrad_model_k * (float)patches / (float)items) + 0.5f);
    if (s_model < 1) s_model = 1;
    if (splits > s_model) splits = s_model;
  }
  if (splits > patches) splits = patches;
  if (splits * items > pairs && (splits * items) % pairs != 0) {  // whole waves of pairs
    const int whole = ((splits * items) / pairs) * pairs / items;
    if (whole >= 1) splits = whole;
  }
  if (splits < 1) splits = 1;
  dim3 grid(2 * splits, items);  // clusters of two CTAs along x (compile-time __cluster_dims__)
  {
    cudaError_t _le = tdl::launch_pdl(conv_wgrad2_kernel<BLOCK_N>, dim3(grid), dim3(kThreads), Cfg::kSmemBytes, st,
                                      tdy, tx, g, dw, m_tiles, n_tiles,
                                      (g_wgrad_bulk != 0 && (reinterpret_cast<uintptr_t>(dw) & 15) == 0) ? 1 : 0);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "conv_wgrad2_kernel");
  }
  TDL_LAUNCH_OK("conv_wgrad2_kernel");
  return 0;
}

}  // namespace conv
}  // namespace tdl

using namespace tdl;
using namespace tdl::conv;

// Where the output pixel (n, p, q) lives: element ((n*full_h + p*step + off_h)*full_w + q*step + off_w)*K of `y`.
// The plain convolution uses {P, Q, 1, 0, 0}; the strided data gradient writes one input-parity class at a time.
struct OutView {
  int full_h, full_w, step, off_h, off_w;
};

// The BatchNorm a data-gradient launch reduces for (kBwdStats): its input tensor and batch statistics.
struct BnBwd {
  const void* x;
  const float* mean;
  const float* invstd;
  const float* scale;  // both null, or both set = "recompute the ReLU mask from x for the sums"
  const float* shift;
};

// A strided subset of the taps of a larger filter: `w` holds w_taps taps per output channel and filter position
// (r, s) of this launch reads tap base + r * sr + s * ss of it.
struct TapSel {
  int w_taps, base, sr, ss;
};

// CTA-pair (cta_group::2) kernel for Cout tiles of 128 / 256: -1 = read TDL_CONV_2CTA on first use, 0 = off, 1 = on
static int g_pair_mode = -1;

static int g_halo_mode = -1;  // -1 = read TDL_HALO on first use, 0 = off (default), 1 = on, 2 = on with base offsets
static int g_halo_res_mode = -1;  // halo form with resident weights where they fit: -1 = read TDL_HALO_RES, default 1
static int g_wrow_mode = -1;      // wgrad row mode for 64-channel 3-wide filters: -1 = read TDL_WGRAD_ROW, default 1

static int fprop_impl(const void* x, const void* w, void* y, const float* bias, const void* residual, int N, int H,
                      int W, int C, int K, int R, int S, int stride, int dil, int pad_h, int pad_w, int P, int Q,
                      int relu, const OutView* view, float* stats, int* nparts, void* stream,
                      const void* out_mask = nullptr, const BnBwd* bnb = nullptr, const TapSel* taps = nullptr) {
  TDL_REQUIRE(C % 8 == 0 && K % 8 == 0, "channel counts must be multiples of 8 (16-byte TMA strides)");
  TDL_REQUIRE(stride >= 1 && stride <= 8 && dil >= 1, "bad stride/dilation");
  TDL_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(y) & 15) == 0,
              "tensors must be 16-byte aligned");
  TDL_REQUIRE(!(view && residual), "residual fusion is not available with a strided output view");
  Geometry g;
  memset(&g, 0, sizeof(g));
  const bool flat = !view && (R == 1 && S == 1 && stride == 1 && pad_h == 0 && pad_w == 0);
  if (flat) {  // pure GEMM: all pixels form one long row of width N*H*W
    g.N = 1; g.H = 1; g.W = N * H * W; g.P = 1; g.Q = N * P * Q;
  } else {
    g.N = N; g.H = H; g.W = W; g.P = P; g.Q = Q;
  }
  g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil = dil; g.pad_h = pad_h; g.pad_w = pad_w;
  g.tap_base = taps ? taps->base : 0; g.tap_sr = taps ? taps->sr : S; g.tap_ss = taps ? taps->ss : 1;
  if (taps) {
    const int last = taps->base + (R - 1) * taps->sr + (S - 1) * taps->ss;
    TDL_REQUIRE(taps->base >= 0 && last >= 0 && taps->base < taps->w_taps && last < taps->w_taps,
                "tap selection walks outside the weight tensor");
  }
  choose_patch(g.N, g.P, g.Q, kBlockM, stride, &g.TN, &g.TH, &g.TW);
  g.tiles_n = (int)cdiv(g.N, g.TN);
  g.tiles_h = (int)cdiv(g.P, g.TH);
  g.tiles_w = (int)cdiv(g.Q, g.TW);
  g.num_kb = (int)cdiv(C, kBlockK);
  const int block_n_pre = K > 128 ? 256 : (K > 64 ? 128 : 64);
  {
    // halo mode: stride-1 filters with taps, maps at least 12 pixels high and wide, dilation <= 2.  Tile = TH rows x 8
    // columns so that every 8-row group of the M tile is one tile row and the groups sit SBO = patch-row bytes apart.
    // Measured on B200: (1) the unit derives the 128B-swizzle XOR from the absolute shared-memory address of every
    // row, exactly like the TMA write did, so a start address shifted by whole 128-byte rows needs NO "matrix base
    // offset" (setting it to (addr >> 7) & 7 breaks the result; g_halo_mode 2 keeps that interpretation testable);
    // (2) the halo form is numerically right but SLOWER than the tap-shifted form (3x3 64->64 @56x56: 130 us vs
    // 101 us, 128->128 @28x28: 79 vs 62 us): those layers are bound by per-stage issue overhead of narrow tiles, not
    // by L2->smem bytes.  It therefore stays off unless tdl_conv2d_set_halo_mode(1) / TDL_HALO=1 asks for it.
    if (g_halo_mode < 0) {
      const char* e = getenv("TDL_HALO");
      g_halo_mode = (e != nullptr) ? atoi(e) : 0;
    }
    if (g_halo_res_mode < 0) {
      const char* e = getenv("TDL_HALO_RES");
      g_halo_res_mode = (e != nullptr) ? atoi(e) : 1;
    }
    const int enabled = g_halo_mode != 0, use_bo = g_halo_mode == 2;
    if (!taps && !flat && !view && residual == nullptr && bnb == nullptr && R * S > 1 && stride == 1 && dil <= 2 && P >= 12 &&
        Q >= 12) {
      const int th_tiles = (int)cdiv(P, 16);
      const int th = (int)cdiv(P, th_tiles);
      const int pw = (int)cdiv(8 + (S - 1) * dil, 8) * 8;
      const int ph = th + (R - 1) * dil;
      const int pbytes = (int)cdiv((int64_t)pw * ph * 128, 1024) * 1024;
      const int bbytes = block_n_pre * kBlockK * 2;
      const int ring = 192 * 1024 - 2 * pbytes;  // every BLOCK_N configuration has 192 KB of pipeline stages
      int nb = ring / bbytes;
      const int max_nb = block_n_pre == 256 ? 4 : (block_n_pre == 128 ? 6 : 8);
      if (nb > max_nb) nb = max_nb;
      // Resident weights (halo_res): with ONE Cout tile and the whole [K blocks][taps] weight tensor fitting behind the
      // patch buffers (3x3 64 -> 64: 72 KB), the per-tap weight loads and their barrier round trips -- what made the
      // plain halo form lose -- disappear: a tile is one patch load and R*S*4 MMAs issued back to back.
      // The resident form needs no 8-pixel rounding of the patch width (the swizzle follows absolute addresses, so any
      // 128-byte row pitch works) and spends the rest of the ring on patch buffers: a patch takes ~2 us to arrive and
      // a tile's MMAs ~0.6 us, so two buffers would leave the tensor core waiting.
      const int pw_r = 8 + (S - 1) * dil;
      const int pbytes_r = (int)cdiv((int64_t)pw_r * ph * 128, 1024) * 1024;
      const int wbytes_r = R * S * g.num_kb * bbytes;
      int nb_r = (192 * 1024 - wbytes_r) / pbytes_r;
      if (nb_r > (block_n_pre == 64 ? 8 : 6)) nb_r = block_n_pre == 64 ? 8 : 6;  // the stage barriers there are
      const bool res = g_halo_res_mode != 0 && K <= block_n_pre && wbytes_r < 192 * 1024 && nb_r >= 3;
      if (res) {
        g.halo = 1; g.halo_pw = pw_r; g.halo_pbytes = pbytes_r; g.halo_ptx = pw_r * ph * 128; g.halo_nb = nb_r;
        g.halo_bo = 0; g.halo_res = 1;
        g.TN = 1; g.TH = th; g.TW = 8;
        g.tiles_n = g.N; g.tiles_h = (int)cdiv(g.P, g.TH); g.tiles_w = (int)cdiv(g.Q, 8);
      } else if (enabled && nb >= 2) {
        g.halo = 1; g.halo_pw = pw; g.halo_pbytes = pbytes; g.halo_ptx = pw * ph * 128; g.halo_nb = nb;
        g.halo_bo = use_bo;
        g.TN = 1; g.TH = th; g.TW = 8;
        g.tiles_n = g.N; g.tiles_h = (int)cdiv(g.P, g.TH); g.tiles_w = (int)cdiv(g.Q, 8);
      }
    }
  }
  g.relu = relu; g.has_bias = bias != nullptr; g.has_res = residual != nullptr;
  // "+ residual" without statistics / strided view: add it on the tensor core (an extra A operand x identity block);
  // the epilogue then never waits on a global load of the residual -- bias and ReLU, if any, stay in the epilogue
  g.res_mma = (residual != nullptr && view == nullptr && (!bnb || (!bias && !relu))) ? 1 : 0;
  if (g.res_mma) g.has_res = 0;
  if (bnb) {
    TDL_REQUIRE(stats && nparts && !view && bias == nullptr && !relu && stride == 1,
                "BatchNorm-backward sums: plain stride-1 data gradient only");
    TDL_REQUIRE(bnb->x && bnb->mean && bnb->invstd, "BatchNorm tensors required");
    TDL_REQUIRE((int64_t)N * P * Q < ((int64_t)1 << 31), "pixel index exceeds 32 bits");
    TDL_REQUIRE((bnb->scale == nullptr) == (bnb->shift == nullptr), "scale and shift come together");
    TDL_REQUIRE(!(bnb->scale && out_mask), "either a stored mask or a recomputed one");
    g.bn_x = bnb->x; g.bn_mean = bnb->mean; g.bn_invstd = bnb->invstd; g.bn_scale = bnb->scale; g.bn_shift = bnb->shift;
  }
  if (out_mask) {
    TDL_REQUIRE(g.res_mma || bnb, "the output mask needs a plain residual or the BatchNorm-backward sums");
    TDL_REQUIRE(K % 32 == 0, "the output mask needs K % 32 == 0");
    TDL_REQUIRE(bias == nullptr && !relu, "the output mask excludes bias / ReLU");
    g.has_mask = 1;
  }
  const int block_n = K > 128 ? 256 : (K > 64 ? 128 : 64);
  g.n_tiles = (int)cdiv(K, block_n);
  if (g_pair_mode < 0) {
    const char* e = getenv("TDL_CONV_2CTA");
    g_pair_mode = e ? atoi(e) : 1;
  }
  // CTA pairs: M = 256 MMAs, each CTA loads half of the weight tile (conv_fprop2_kernel).  Needs at least two M tiles,
  // and a K loop of at least four stages: with fewer (e.g. 1x1 128 -> 512) the per-tile hand-over between two CTAs
  // costs more than the halved weight traffic saves (measured 0.93x there, 1.05-1.12x on the deep layers,
  // profiles/conv_pair_timing.txt).
  const bool pair = g_pair_mode != 0 && !g.halo && !bnb && block_n >= 128 &&
                    g.tiles_n * g.tiles_h * g.tiles_w >= 2 && R * S * g.num_kb >= 4;

  CUtensorMap tx, tw;
  {
    uint64_t dims[4] = {(uint64_t)C, (uint64_t)g.W, (uint64_t)g.H, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)(g.TW * stride), (uint32_t)(g.TH * stride), (uint32_t)g.TN};
    uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
    if (g.halo) {  // the whole haloed patch of one 64-channel block
      box[1] = (uint32_t)g.halo_pw;
      box[2] = (uint32_t)(g.TH + (R - 1) * dil);
      box[3] = 1;
    }
    int rc = encode_bf16(&tx, x, 4, dims, box, es);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)C, (uint64_t)(taps ? taps->w_taps : R * S), (uint64_t)K};
    uint32_t box[3] = {64, 1, (uint32_t)(pair ? block_n / 2 : block_n)};
    uint32_t es[3] = {1, 1, 1};
    int rc = encode_bf16(&tw, w, 3, dims, box, es);
    if (rc) return rc;
  }
  CUtensorMap ty;
  {
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)g.Q, (uint64_t)g.P, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc;
    if (view) {
      const uint64_t kb = (uint64_t)K * 2;
      uint64_t strides[3] = {kb * view->step, kb * view->full_w * view->step, kb * view->full_w * view->full_h};
      const char* base = (const char*)y + ((size_t)view->off_h * view->full_w + view->off_w) * kb;
      rc = encode_bf16(&ty, base, 4, dims, box, es, strides);
    } else {
      rc = encode_bf16(&ty, y, 4, dims, box, es);
    }
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap tres = ty, tid = ty;
  if (g.res_mma) {
    TDL_REQUIRE((reinterpret_cast<uintptr_t>(residual) & 15) == 0, "residual must be 16-byte aligned");
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)g.Q, (uint64_t)g.P, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_bf16(&tres, residual, 4, dims, box, es);
    if (rc) return rc;
    const void* ident = nullptr;
    rc = identity64_address(&ident, st);
    if (rc) return rc;
    uint64_t idims[2] = {64, 64};
    uint32_t ibox[2] = {64, (uint32_t)(pair ? 32 : 64)};
    rc = encode_bf16(&tid, ident, 2, idims, ibox, es);
    if (rc) return rc;
  }
  // in res_mma mode the residual arrives through its tensor map; the pointer slot carries the optional output mask
  const bf16* aux = (g.res_mma || bnb) ? (const bf16*)out_mask : (const bf16*)residual;
  CUtensorMap tbx = ty;
  if (bnb) {  // the BatchNorm input, tiled exactly like the output
    TDL_REQUIRE((reinterpret_cast<uintptr_t>(bnb->x) & 15) == 0, "bn_x must be 16-byte aligned");
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)g.Q, (uint64_t)g.P, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_bf16(&tbx, bnb->x, 4, dims, box, es);
    if (rc) return rc;
  }
  if (pair) {
    if (block_n == 256) return launch_fprop2<256>(tx, tw, ty, tres, tid, g, bias, aux, stats, nparts, st);
    return launch_fprop2<128>(tx, tw, ty, tres, tid, g, bias, aux, stats, nparts, st);
  }
  if (block_n == 256) return launch_fprop<256>(tx, tw, ty, tres, tid, tbx, g, bias, aux, stats, nparts, st);
  if (block_n == 128) return launch_fprop<128>(tx, tw, ty, tres, tid, tbx, g, bias, aux, stats, nparts, st);
  return launch_fprop<64>(tx, tw, ty, tres, tid, tbx, g, bias, aux, stats, nparts, st);
}

// y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R*S,C]) (+bias) (+residual) (ReLU)      bf16 in/out, fp32 accumulate
TDL_API int tdl_conv2d_fprop_bf16(const void* x, const void* w, void* y, const float* bias, const void* residual,
                                  int N, int H, int W, int C, int K, int R, int S, int stride, int dil, int pad_h,
                                  int pad_w, int P, int Q, int relu, void* stream) {
  return fprop_impl(x, w, y, bias, residual, N, H, W, C, K, R, S, stride, dil, pad_h, pad_w, P, Q, relu, nullptr,
                    nullptr, nullptr, stream);
}

// y = (conv(x, w) + residual) with outputs zeroed where the bit mask says so: out_mask is [N*P*Q][K/8] bytes, bit j of
// byte (pixel, c/8) set = keep channel 8*(c/8)+j.  This is the data gradient of a residual unit's first conv with the
// identity-shortcut gradient (residual) and the ReLU backward of the unit's input (mask, as written by tdl_bn_apply)
// folded in; the residual is added on the tensor core.  K % 32 == 0.
TDL_API int tdl_conv2d_fprop_res_mask_bf16(const void* x, const void* w, void* y, const void* residual,
                                           const void* out_mask, int N, int H, int W, int C, int K, int R, int S,
                                           int stride, int dil, int pad_h, int pad_w, int P, int Q, void* stream) {
  TDL_REQUIRE(residual != nullptr, "residual required");
  return fprop_impl(x, w, y, nullptr, residual, N, H, W, C, K, R, S, stride, dil, pad_h, pad_w, P, Q, 0, nullptr,
                    nullptr, nullptr, stream, out_mask);
}

// Data gradient (a stride-1 convolution of dy with the flipped weights) that ALSO reduces, for the BatchNorm whose
// output gradient it produces, the per-channel sums BatchNorm backward needs: stats_partial [tdl_num_sms()][2][K]
// = per-CTA (sum dz, sum dz * xhat) with dz the stored (optionally masked / residual-added) output and
// xhat = (bn_x - mean) * invstd.  bn_scale / bn_shift (optional, not with out_mask): the BatchNorm ends in a ReLU
// whose mask fmaf(x, scale, shift) > 0 is applied to the SUMS only (the stored gradient stays unmasked, as
// tdl_bn_bwd_apply re-derives the same mask).  Feed to tdl_bn_bwd_finalize_partials; tdl_bn_bwd_reduce's pass over (dz, x) is gone.
TDL_API int tdl_conv2d_dgrad_bnstats_bf16(const void* x, const void* w, void* y, const void* residual,
                                          const void* out_mask, const void* bn_x, const float* bn_mean,
                                          const float* bn_invstd, const float* bn_scale, const float* bn_shift, int N, int H, int W, int C, int K, int R, int S,
                                          int dil, int pad_h, int pad_w, int P, int Q, float* stats_partial,
                                          int* nparts_host, void* stream) {
  BnBwd b = {bn_x, bn_mean, bn_invstd, bn_scale, bn_shift};
  return fprop_impl(x, w, y, nullptr, residual, N, H, W, C, K, R, S, 1, dil, pad_h, pad_w, P, Q, 0, nullptr,
                    stats_partial, nparts_host, stream, out_mask, &b);
}

// Same convolution, additionally producing per-CTA partial (sum, sum of squares) of the bf16 output per channel:
// stats_partial is [tdl_num_sms()][2][K] fp32 (written, not accumulated); *nparts_host receives how many leading
// slots are valid.  Feed both to tdl_bn_finalize_partials: the BatchNorm statistics pass costs no extra read.
TDL_API int tdl_conv2d_fprop_stats_bf16(const void* x, const void* w, void* y, const float* bias, int N, int H, int W,
                                        int C, int K, int R, int S, int stride, int dil, int pad_h, int pad_w, int P,
                                        int Q, float* stats_partial, int* nparts_host, void* stream) {
  TDL_REQUIRE(stats_partial != nullptr && nparts_host != nullptr, "stats buffers required");
  return fprop_impl(x, w, y, bias, nullptr, N, H, W, C, K, R, S, stride, dil, pad_h, pad_w, P, Q, 0, nullptr,
                    stats_partial, nparts_host, stream);
}

// y = relu?(conv(x, w) + bias? + residual) with the residual added on the tensor core AND the per-CTA partial
// (sum, sum of squares) of the stored output, for the BatchNorm that consumes y next (the reference unit's
// relu(shortcut + conv3(.)) feeding the next unit's pre-activation BatchNorm, core/resnet.py:113,148-150).
TDL_API int tdl_conv2d_fprop_res_stats_bf16(const void* x, const void* w, void* y, const float* bias,
                                            const void* residual, int N, int H, int W, int C, int K, int R, int S,
                                            int stride, int dil, int pad_h, int pad_w, int P, int Q, int relu,
                                            float* stats_partial, int* nparts_host, void* stream) {
  TDL_REQUIRE(residual != nullptr && stats_partial != nullptr && nparts_host != nullptr, "residual and stats required");
  return fprop_impl(x, w, y, bias, residual, N, H, W, C, K, R, S, stride, dil, pad_h, pad_w, P, Q, relu, nullptr,
                    stats_partial, nparts_host, stream);
}

// Stride-1 convolution whose output pixel (p, q) is stored at (p*y_step + y_off_h, q*y_step + y_off_w) of a
// [N, y_full_h, y_full_w, K] tensor.  One call computes one input-parity class of a strided convolution's data
// gradient (x = dy, w = the taps of that class, flipped and transposed; see ops/functional.py).
TDL_API int tdl_conv2d_fprop_strided_out_bf16(const void* x, const void* w, void* y, int N, int H, int W, int C, int K,
                                              int R, int S, int dil, int pad_h, int pad_w, int P, int Q, int y_full_h,
                                              int y_full_w, int y_step, int y_off_h, int y_off_w, void* stream) {
  OutView v = {y_full_h, y_full_w, y_step, y_off_h, y_off_w};
  TDL_REQUIRE(y_step >= 1 && y_off_h >= 0 && y_off_w >= 0, "bad output view");
  TDL_REQUIRE((P - 1) * y_step + y_off_h < y_full_h && (Q - 1) * y_step + y_off_w < y_full_w, "view exceeds tensor");
  return fprop_impl(x, w, y, nullptr, nullptr, N, H, W, C, K, R, S, 1, dil, pad_h, pad_w, P, Q, 0, &v, nullptr, nullptr,
                    stream);
}

// Same, reading its R x S taps out of a LARGER filter: `w` is [K, w_taps, C] and filter position (r, s) uses tap
// tap_base + r * tap_sr + s * tap_ss.  With w = the flipped / transposed data-gradient copy of a strided convolution's
// weights (tdl_conv2d_weight_prep) one call per input-parity class needs no per-class weight copy.
TDL_API int tdl_conv2d_fprop_strided_out_taps_bf16(const void* x, const void* w, void* y, int N, int H, int W, int C,
                                                   int K, int R, int S, int w_taps, int tap_base, int tap_sr, int tap_ss,
                                                   int dil, int pad_h, int pad_w, int P, int Q, int y_full_h,
                                                   int y_full_w, int y_step, int y_off_h, int y_off_w, void* stream) {
  OutView v = {y_full_h, y_full_w, y_step, y_off_h, y_off_w};
  TapSel t = {w_taps, tap_base, tap_sr, tap_ss};
  TDL_REQUIRE(y_step >= 1 && y_off_h >= 0 && y_off_w >= 0, "bad output view");
  TDL_REQUIRE((P - 1) * y_step + y_off_h < y_full_h && (Q - 1) * y_step + y_off_w < y_full_w, "view exceeds tensor");
  return fprop_impl(x, w, y, nullptr, nullptr, N, H, W, C, K, R, S, 1, dil, pad_h, pad_w, P, Q, 0, &v, nullptr, nullptr,
                    stream, nullptr, nullptr, &t);
}

// dw[K,R*S,C] (fp32) += sum over pixels of dy[N,P,Q,K]^T * shifted x[N,H,W,C].  Caller zeroes dw.
TDL_API int tdl_conv2d_wgrad_bf16(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int K, int R,
                                  int S, int stride, int dil, int pad_h, int pad_w, int P, int Q, void* stream) {
  TDL_REQUIRE(C % 8 == 0 && K % 8 == 0, "channel counts must be multiples of 8 (16-byte TMA strides)");
  Geometry g;
  memset(&g, 0, sizeof(g));
  const bool flat = (R == 1 && S == 1 && stride == 1 && pad_h == 0 && pad_w == 0);
  if (flat) {
    g.N = 1; g.H = 1; g.W = N * H * W; g.P = 1; g.Q = N * P * Q;
  } else {
    g.N = N; g.H = H; g.W = W; g.P = P; g.Q = Q;
  }
  g.C = C; g.K = K; g.R = R; g.S = S; g.stride = stride; g.dil = dil; g.pad_h = pad_h; g.pad_w = pad_w;
  choose_patch(g.N, g.P, g.Q, 64, stride, &g.TN, &g.TH, &g.TW);
  g.tiles_n = (int)cdiv(g.N, g.TN);
  g.tiles_h = (int)cdiv(g.P, g.TH);
  g.tiles_w = (int)cdiv(g.Q, g.TW);
  const int block_n = C >= 256 ? 256 : (C > 64 ? 128 : 64);
  const int mt = K >= 256 ? 2 : 1;
  if (g_wrow_mode < 0) {
    const char* e = getenv("TDL_WGRAD_ROW");
    g_wrow_mode = e ? atoi(e) : 1;
  }
  // row mode (see Geometry::wrow): the three taps of a filter row share one haloed X tile
  if (g_wrow_mode != 0 && !flat && mt == 1 && block_n == 64 && S == 3 && stride == 1 && dil == 1 && pad_w == 1 &&
      Q == W && Q + 2 <= 64) {
    g.wrow = g_wrow_mode;
    g.TN = 1; g.TW = Q + 2; g.TH = 64 / (Q + 2);
    if (g.TH > P) g.TH = P;
    g.tiles_n = N; g.tiles_h = (int)cdiv(P, g.TH); g.tiles_w = 1;
  }
  CUtensorMap tdy, tx;
  {
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)g.Q, (uint64_t)g.P, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_bf16(&tdy, dy, 4, dims, box, es);
    if (rc) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)C, (uint64_t)g.W, (uint64_t)g.H, (uint64_t)g.N};
    uint32_t box[4] = {64, (uint32_t)(g.TW * stride), (uint32_t)(g.TH * stride), (uint32_t)g.TN};
    uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
    int rc = encode_bf16(&tx, x, 4, dims, box, es);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  // CTA pairs for 256-wide Cout tiles (conv_wgrad2_kernel).  Alone the pair kernel is 1.3-1.5x faster on the deep 1x1
  // layers (1024 -> 256 @14x14: 57 -> 39 us, 2048 -> 512 @7x7: 51 -> 35 us; profiles/wgrad_epilogue_and_grid.txt), but
  // inside the training step the weight gradients run on a side stream next to the data-gradient chain, are not on
  // the critical path, and two-CTA clusters wait for a free SM pair: ResNet-50 unchanged in cycles, Xception-41 /
  // the reference network 0.6 % / 2 % slower.  Hence opt-in: TDL_WGRAD_2CTA=1 / tdl_conv2d_set_wgrad_pair_mode(1).
  if (g_wgrad_pair < 0) {
    const char* e = getenv("TDL_WGRAD_2CTA");
    g_wgrad_pair = e ? atoi(e) : 0;
  }
  if (mt == 2 && g_wgrad_pair != 0 && block_n >= 128) {
    if (block_n == 256) return launch_wgrad2<256>(tdy, tx, g, dw, st);
    return launch_wgrad2<128>(tdy, tx, g, dw, st);
  }
  if (mt == 2) {
    if (block_n == 256) return launch_wgrad<2, 256>(tdy, tx, g, dw, st);
    if (block_n == 128) return launch_wgrad<2, 128>(tdy, tx, g, dw, st);
    return launch_wgrad<2, 64>(tdy, tx, g, dw, st);
  }
  if (block_n == 256) return launch_wgrad<1, 256>(tdy, tx, g, dw, st);
  // narrow filters with taps: several taps per CTA share the dY tile and one wide MMA (see WgradCfg)
  const bool pack = R * S >= 2 && C <= 128;
  if (block_n == 128) return pack ? launch_wgrad<1, 128, 2>(tdy, tx, g, dw, st) : launch_wgrad<1, 128>(tdy, tx, g, dw, st);
  return pack ? launch_wgrad<1, 64, 3>(tdy, tx, g, dw, st) : launch_wgrad<1, 64>(tdy, tx, g, dw, st);
}

// fp32 master weights -> bf16 compute copies (fprop layout and flipped/transposed dgrad layout)
TDL_API int tdl_conv2d_weight_prep(const float* w, void* w_fprop, void* w_dgrad, int K, int taps, int C,
                                   void* stream) {
  const size_t total = (size_t)K * taps * C;
  if (total == 0) return 0;
  if (w_dgrad != nullptr && taps <= 65535 && cdiv(K, 32) <= 65535) {
    dim3 grid((unsigned)cdiv(C, 32), (unsigned)cdiv(K, 32), (unsigned)taps);
    {
      cudaError_t _le = tdl::launch_pdl(weight_prep_tiled_kernel, dim3(grid), dim3(dim3(32, 8)), 0,
          (cudaStream_t)stream, w, (bf16*)w_fprop, (bf16*)w_dgrad, K, taps, C);
      if (_le != cudaSuccess) return tdl::cuda_fail(_le, "weight_prep_tiled_kernel");
    }
    TDL_LAUNCH_OK("weight_prep_tiled_kernel");
    return 0;
  }
  int blocks = (int)cdiv((int64_t)total, 256);
  if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
  {
    cudaError_t _le = tdl::launch_pdl(weight_prep_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w,
        (bf16*)w_fprop, (bf16*)w_dgrad, K, taps, C);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "weight_prep_kernel");
  }
  TDL_LAUNCH_OK("weight_prep_kernel");
  return 0;
}

// Batched form: `descs` is a device array of nlayers 48-byte records {const float* w; bf16* wf; bf16* wd (may be
// null); int K, taps, C, ctiles = ceil(C/64), ktiles = ceil(K/64), 0}; block_start[l] = sum over earlier layers of
// ctiles * ktiles * taps (device, nlayers ints); total_blocks = the full sum.
TDL_API int tdl_conv2d_weight_prep_batched(const void* descs, const int* block_start, int nlayers, int total_blocks,
                                           void* stream) {
  static_assert(sizeof(WPrepDesc) == 48, "record layout is part of the ABI");
  if (nlayers <= 0 || total_blocks <= 0) return 0;
  cudaError_t _le = tdl::launch_pdl(weight_prep_batched_kernel, dim3(total_blocks), dim3(256), 0, (cudaStream_t)stream,
                                    (const WPrepDesc*)descs, block_start, nlayers);
  if (_le != cudaSuccess) return tdl::cuda_fail(_le, "weight_prep_batched_kernel");
  TDL_LAUNCH_OK("weight_prep_batched_kernel");
  return 0;
}

// ---- first layer (stride-2 conv on <= 4 channels) -------------------------------------------------------------
static void stem_taps(int R, int pad, int* taps, int* pad2) {
  *pad2 = (pad + 1) / 2;
  const int hi = (R - 1 - pad) >= 0 ? (R - 1 - pad) / 2 : -((pad - R + 2) / 2);
  *taps = hi + *pad2 + 1;
}

// x [N,H,W,C] bf16 -> planes [N][ceil(H/2)][2][ceil(W/2)][8] bf16 (space-to-depth, two 8-channel planes)
TDL_API int tdl_conv_stem_pack(const void* x, void* planes, int N, int H, int W, int C, void* stream) {
  TDL_REQUIRE(C >= 1 && C <= 4, "stem pack handles 1..4 input channels");
  const int H2 = (H + 1) / 2, W2 = (W + 1) / 2;
  const int64_t total = (int64_t)N * H2 * 2 * W2;
  if (total == 0) return 0;
  int64_t blocks = cdiv(total, 256);
  if (blocks > (int64_t)num_sms() * 16) blocks = (int64_t)num_sms() * 16;
  if (C == 3 && H % 2 == 0 && W % 2 == 0 && (reinterpret_cast<uintptr_t>(x) & 3) == 0) {
    int64_t b3 = cdiv(total / 2, 256);
    if (b3 > (int64_t)num_sms() * 16) b3 = (int64_t)num_sms() * 16;
    cudaError_t _le = tdl::launch_pdl(stem_pack_c3_kernel, dim3((int)b3), dim3(256), 0, (cudaStream_t)stream,
                                      (const bf16*)x, (bf16*)planes, N, H, W, H2, W2);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "stem_pack_c3_kernel");
    TDL_LAUNCH_OK("stem_pack_c3_kernel");
    return 0;
  }
  {
    cudaError_t _le = tdl::launch_pdl(stem_pack_kernel, dim3((int)blocks), dim3(256), 0, (cudaStream_t)stream,
        (const bf16*)x, (bf16*)planes, N, H, W, C, H2, W2);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "stem_pack_kernel");
  }
  TDL_LAUNCH_OK("stem_pack_kernel");
  return 0;
}

// Number of bf16 elements of the packed weight tensor [K][Th*Tw][16] for this filter / padding.
TDL_API int tdl_conv_stem_weight_elems(int K, int R, int S, int pad_h, int pad_w) {
  int th, tw, p2h, p2w;
  stem_taps(R, pad_h, &th, &p2h);
  stem_taps(S, pad_w, &tw, &p2w);
  return K * th * tw * 16;
}

TDL_API int tdl_conv_stem_weight_prep(const float* w, void* w_packed, int K, int R, int S, int C, int pad_h, int pad_w,
                                      void* stream) {
  TDL_REQUIRE(C >= 1 && C <= 4, "stem handles 1..4 input channels");
  int th, tw, p2h, p2w;
  stem_taps(R, pad_h, &th, &p2h);
  stem_taps(S, pad_w, &tw, &p2w);
  const int total = K * th * tw * 16;
  {
    cudaError_t _le = tdl::launch_pdl(stem_weight_prep_kernel, dim3((int)cdiv(total, 256)), dim3(256), 0,
        (cudaStream_t)stream, w, (bf16*)w_packed, K, R, S, C, th, tw, pad_h, pad_w, p2h, p2w);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "stem_weight_prep_kernel");
  }
  TDL_LAUNCH_OK("stem_weight_prep_kernel");
  return 0;
}

// y[N,P,Q,K] = stride-2 conv of the packed image with the packed weights.  (H, W) are the ORIGINAL image size.
// stats_partial / nparts_host as in tdl_conv2d_fprop_stats_bf16 (both null = no statistics).
TDL_API int tdl_conv_stem_fprop(const void* planes, const void* w_packed, void* y, int N, int H, int W, int K, int R,
                                int S, int pad_h, int pad_w, int P, int Q, float* stats_partial, int* nparts_host,
                                void* stream) {
  TDL_REQUIRE(K % 8 == 0 && K >= 8 && K <= 64, "stem kernel handles 8..64 output channels (multiple of 8)");
  TDL_REQUIRE((stats_partial == nullptr) == (nparts_host == nullptr), "stats buffers come in pairs");
  int th, tw, p2h, p2w;
  stem_taps(R, pad_h, &th, &p2h);
  stem_taps(S, pad_w, &tw, &p2w);
  TDL_REQUIRE(th >= 1 && tw >= 1 && th * tw <= 16, "stem kernel handles at most 16 packed taps (filters up to 7x7)");
  const int H2 = (H + 1) / 2, W2 = (W + 1) / 2;
  Geometry g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.H = H2; g.W = W2; g.C = 16; g.P = P; g.Q = Q; g.K = K; g.R = th; g.S = tw; g.stride = 1; g.dil = 1;
  g.pad_h = p2h; g.pad_w = p2w;
  g.TN = 1; g.TH = 1;
  const int w_tiles = (int)cdiv(Q, 120);
  g.TW = (int)cdiv(Q, w_tiles);
  g.tiles_n = N; g.tiles_h = P; g.tiles_w = (int)cdiv(Q, g.TW);
  g.num_kb = 1; g.n_tiles = 1;
  const int box_px = g.TW + tw - 1;
  g.sp_plane = box_px * 16;
  g.sp_row = 2 * g.sp_plane;
  g.sp_tx = th * g.sp_row;
  g.sp_wbytes = (int)cdiv(th * tw * 16, 64) * 8192;
  TDL_REQUIRE(box_px * 2 <= 256, "stem tile too wide for one TMA box");
  TDL_REQUIRE((th - 1) * g.sp_row + g.sp_plane + (tw - 1) * 16 + 2048 <= kABytes, "stem patch exceeds its smem stage");
  CUtensorMap tx, tmw, ty;
  {
    // the plane line of W2 pixels x 16 bytes is addressed in 8-byte units so one box row can span the whole tile
    uint64_t dims[4] = {(uint64_t)W2 * 2, 2, (uint64_t)H2, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)W2 * 16, (uint64_t)W2 * 32, (uint64_t)W2 * 32 * H2};
    uint32_t box[4] = {(uint32_t)box_px * 2, 2, (uint32_t)th, 1};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map(&tx, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, CU_TENSOR_MAP_SWIZZLE_NONE, planes, 4, dims, box, es,
                        strides);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)(th * tw * 16), (uint64_t)K};
    uint32_t box[2] = {64, 64};
    uint32_t es[2] = {1, 1};
    int rc = encode_bf16(&tmw, w_packed, 2, dims, box, es);
    if (rc) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)Q, (uint64_t)P, (uint64_t)N};
    uint32_t box[4] = {64, (uint32_t)g.TW, 1, 1};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_bf16(&ty, y, 4, dims, box, es);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (stats_partial) return launch_stem<true>(tx, tmw, ty, g, stats_partial, nparts_host, st);
  return launch_stem<false>(tx, tmw, ty, g, nullptr, nullptr, st);
}

// Weight gradient of the first layer from the packed image (see stem_wgrad_kernel).  `dwp_scratch` holds
// tdl_conv_stem_weight_elems() floats; dw is the fp32 [K, R, S, C] gradient (overwritten, or added to if accumulate).
TDL_API int tdl_conv_stem_wgrad(const void* planes, const void* dy, float* dwp_scratch, float* dw, int N, int H, int W,
                                int C, int K, int R, int S, int pad_h, int pad_w, int P, int Q, int accumulate,
                                int variant, void* stream) {
  TDL_REQUIRE(K % 8 == 0 && K >= 8 && K <= 64, "stem kernel handles 8..64 output channels (multiple of 8)");
  TDL_REQUIRE(C >= 1 && C <= 4, "stem handles 1..4 input channels");
  TDL_REQUIRE(variant == 0 || variant == 1, "variant is 0 or 1");
  int th, tw, p2h, p2w;
  stem_taps(R, pad_h, &th, &p2h);
  stem_taps(S, pad_w, &tw, &p2w);
  TDL_REQUIRE(th >= 1 && tw >= 1 && th * tw <= 16, "stem kernel handles at most 16 packed taps (filters up to 7x7)");
  const int H2 = (H + 1) / 2, W2 = (W + 1) / 2;
  StemWgradGeo g;
  memset(&g, 0, sizeof(g));
  g.N = N; g.P = P; g.Q = Q; g.K = K; g.th = th; g.tw = tw; g.pad2_h = p2h; g.pad2_w = p2w; g.variant = variant;
  g.ncols = (8 * tw + 15) / 16 * 16;
  g.nblocks = 2 * th;
  TDL_REQUIRE(g.nblocks * g.ncols <= 256, "packed gradient exceeds 256 TMEM columns");
  const int halo = g.ncols / 8 - 1;  // pixels read past the last output pixel (including rounded-up column groups)
  const int max_twb = (128 - halo) / 16 * 16;
  const int w_tiles = (int)cdiv(Q, max_twb);
  g.TWB = (int)cdiv(cdiv(Q, w_tiles), 16) * 16;
  g.tiles_w = (int)cdiv(Q, g.TWB);
  const int box_px = g.TWB + halo;
  g.sp_plane = box_px * 16;
  g.sp_row = 2 * g.sp_plane;
  g.a_block = g.TWB * 128;
  g.b_off = 2 * g.a_block;
  g.tx_bytes = g.a_block + th * g.sp_row;
  g.stage_bytes = (int)cdiv(g.b_off + th * g.sp_row + 256, 1024) * 1024;
  g.stages = (200 * 1024) / g.stage_bytes;
  if (g.stages > 8) g.stages = 8;
  TDL_REQUIRE(g.stages >= 2, "stem wgrad stage does not fit twice in shared memory");
  const int smem_bytes = g.stages * g.stage_bytes + 1024 + 256;
  CUtensorMap tdy, tx;
  {
    uint64_t dims[4] = {(uint64_t)K, (uint64_t)Q, (uint64_t)P, (uint64_t)N};
    uint32_t box[4] = {64, (uint32_t)g.TWB, 1, 1};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_bf16(&tdy, dy, 4, dims, box, es);
    if (rc) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)W2 * 2, 2, (uint64_t)H2, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)W2 * 16, (uint64_t)W2 * 32, (uint64_t)W2 * 32 * H2};
    uint32_t box[4] = {(uint32_t)box_px * 2, 2, (uint32_t)th, 1};
    uint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map(&tx, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, CU_TENSOR_MAP_SWIZZLE_NONE, planes, 4, dims, box, es,
                        strides);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  void (*kern)(CUtensorMap, CUtensorMap, StemWgradGeo, float*) =
      th == 1 ? stem_wgrad_kernel<2> : th == 2 ? stem_wgrad_kernel<4> : th == 3 ? stem_wgrad_kernel<6>
                                                                                 : stem_wgrad_kernel<8>;
  TDL_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  TDL_CUDA_OK(cudaMemsetAsync(dwp_scratch, 0, sizeof(float) * (size_t)K * th * tw * 16, st));
  const int total = N * P * g.tiles_w;
  const int grid = total < tdl::num_sms() ? total : tdl::num_sms();
  {
    cudaError_t _le = tdl::launch_pdl(kern, dim3(grid), dim3(kThreads), smem_bytes, st, tdy, tx, g, dwp_scratch);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "kern");
  }
  TDL_LAUNCH_OK("stem_wgrad_kernel");
  const int n_out = K * R * S * C;
  {
    cudaError_t _le = tdl::launch_pdl(stem_wgrad_unpack_kernel, dim3((int)cdiv(n_out, 256)), dim3(256), 0, st,
        dwp_scratch, dw, K, R, S, C, th, tw, pad_h, pad_w, p2h, p2w, accumulate);
    if (_le != cudaSuccess) return tdl::cuda_fail(_le, "stem_wgrad_unpack_kernel");
  }
  TDL_LAUNCH_OK("stem_wgrad_unpack_kernel");
  return 0;
}

// 0 = single-CTA MMAs, 1 = CTA pairs (cta_group::2, M = 256) for convolutions with Cout tiles of 128 / 256
TDL_API int tdl_conv2d_set_pair_mode(int mode) {
  TDL_REQUIRE(mode == 0 || mode == 1, "pair mode must be 0 or 1");
  g_pair_mode = mode;
  return 0;
}
TDL_API int tdl_conv2d_get_pair_mode() { return g_pair_mode < 0 ? 0 : g_pair_mode; }

// Experimental: 0 = tap-shifted loads (default), 1 = halo reuse for stride-1 RxS filters on maps >= 12x12 (one patch load
// per 64-channel block, taps read it at a row shift), 2 = the same with "matrix base offset" set (known wrong on B200).
TDL_API int tdl_conv2d_set_halo_mode(int mode) {
  TDL_REQUIRE(mode >= 0 && mode <= 2, "halo mode must be 0, 1 or 2");
  g_halo_mode = mode;
  return 0;
}

// Halo form with the weights resident in shared memory, picked on its own for stride-1 RxS layers with one Cout tile
// whose weights fit beside two patch buffers (3x3 64 -> 64): 1 = on (default), 0 = off.
// CTA-pair weight gradient for Cout >= 256, Cin > 64: 1 = on, 0 = single-CTA kernel (default, see tdl_conv2d_wgrad_bf16)
TDL_API int tdl_conv2d_set_wgrad_pair_mode(int mode) {
  TDL_REQUIRE(mode == 0 || mode == 1, "wgrad pair mode must be 0 or 1");
  tdl::conv::g_wgrad_pair = mode;
  return 0;
}

// wgrad grid tuning (measurement hooks): whole-wave snapping on / off, constant of the split-count model (< 0 = default)
TDL_API int tdl_conv2d_set_wgrad_tuning(int wave_snap, float model_k) {
  tdl::conv::g_wgrad_snap = wave_snap ? 1 : 0;
  tdl::conv::g_wgrad_model_k = model_k < 0.f ? 1.1f : model_k;
  return 0;
}

// wgrad epilogue: 0 = per-thread red.global.add, 1 = bulk reductions through shared memory (default)
TDL_API int tdl_conv2d_set_wgrad_bulk_mode(int mode) {
  TDL_REQUIRE(mode == 0 || mode == 1, "wgrad epilogue mode must be 0 or 1");
  tdl::conv::g_wgrad_bulk = mode;
  return 0;
}

// wgrad row mode (Geometry::wrow): 0 = off, 1 = one N = 192 MMA over row-shifted blocks (default), 2 = three N = 64 MMAs
TDL_API int tdl_conv2d_set_wgrad_row_mode(int mode) {
  TDL_REQUIRE(mode >= 0 && mode <= 2, "wgrad row mode must be 0, 1 or 2");
  g_wrow_mode = mode;
  return 0;
}

TDL_API int tdl_conv2d_set_halo_res_mode(int mode) {
  TDL_REQUIRE(mode == 0 || mode == 1, "resident-weight halo mode must be 0 or 1");
  g_halo_res_mode = mode;
  return 0;
}
