// metrics.cu — per-step segmentation metrics on the device and the argmax-to-uint8 class map
// a mosaic writer needs (row f4).
//
// Reference: src/train_segmentation.py:145 `self.update_metrics(mode, predictions=logits.argmax(dim=1), labels=y)`
// feeding torchmetrics MulticlassConfusionMatrix / IoU / Accuracy / F1 (:53-67), and experiments/inference_demo.py
// which keeps whole logit tensors only to argmax them later.  Every one of those metrics is a function of the
// C x C confusion matrix, so ONE pass over the logits produces everything: counts[target, prediction] += 1 for
// every pixel (all targets, including the ignored class — `ignore_index` is applied on the host by dropping that
// row, which is exactly what torchmetrics does when it filters the ignored samples), plus, optionally, the uint8
// prediction map.  HBM-bound: logits (C x 2 B) + label (8 B) read once, 1 B written per pixel.
#include "common.cuh"

namespace s2lc {

template <typename T, int CLS>
__global__ void __launch_bounds__(256) argmax_confusion_kernel(const T* __restrict__ logits, int64_t ld,
                                                               const int64_t* __restrict__ y,
                                                               unsigned long long* __restrict__ counts,
                                                               uint8_t* __restrict__ pred, int64_t rows) {
  __shared__ unsigned int hist[CLS * CLS];
  for (int i = threadIdx.x; i < CLS * CLS; i += blockDim.x) hist[i] = 0u;
  __syncthreads();
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    float best = ldf<T>(logits + r * ld);
    int arg = 0;
#pragma unroll
    for (int c = 1; c < CLS; ++c) {  // first maximum wins, as torch.argmax; NaN compares false and never wins
      float v = ldf<T>(logits + r * ld + c);
      if (v > best) { best = v; arg = c; }
    }
    if (pred) pred[r] = (uint8_t)arg;
    if (y && counts) {
      int64_t t = y[r];
      if (t >= 0 && t < CLS) atomicAdd(&hist[(int)t * CLS + arg], 1u);
    }
  }
  __syncthreads();
  if (counts)
    for (int i = threadIdx.x; i < CLS * CLS; i += blockDim.x)
      if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

template <typename T>
static int32_t launch_argmax(const void* logits, int64_t ld, const int64_t* y, int64_t* counts, uint8_t* pred,
                             int64_t rows, int n_classes, cudaStream_t st) {
  // a block's shared histogram is 32-bit: bound the rows one block can see (grid-stride) well below 2^32
  int64_t blocks = (rows + 255) / 256;
  const int64_t want = (int64_t)kNumSMs * 8;
  if (blocks > want) blocks = want;
  if (blocks < 1) blocks = 1;
#define AM(C_)                                                                                                   \
  case C_:                                                                                                       \
    argmax_confusion_kernel<T, C_><<<(unsigned)blocks, 256, 0, st>>>((const T*)logits, ld, y,                    \
                                                                     (unsigned long long*)counts, pred, rows);  \
    break;
  switch (n_classes) {
    AM(2) AM(3) AM(4) AM(5) AM(6) AM(7) AM(8) AM(9) AM(10) AM(11) AM(12) AM(13) AM(14) AM(15) AM(16)
    default:
      set_error("argmax_confusion: n_classes=%d unsupported (2..16)", n_classes);
      return S2LC_ERR_SHAPE;
  }
#undef AM
  return check_launch("argmax_confusion");
}

// Mosaic writer (row f4): argmax of B tiles of t x t logits written straight into the uint8 scene mosaic.  Only the
// centre of each tile (margin px trimmed on every side) is kept, and only where it lands inside the mosaic, so tiles
// predicted with overlapping context tile the scene exactly once.  One thread per kept pixel; 1 B written per pixel.
template <typename T, int CLS>
__global__ void __launch_bounds__(256) argmax_mosaic_kernel(const T* __restrict__ logits, int64_t ld,
                                                            const int32_t* __restrict__ origin_yx,
                                                            uint8_t* __restrict__ mosaic, int64_t ld_mosaic, int B, int t,
                                                            int margin, int Hm, int Wm) {
  const int keep = t - 2 * margin;
  const int64_t total = (int64_t)B * keep * keep;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kx = (int)(i % keep) + margin;
    const int ky = (int)((i / keep) % keep) + margin;
    const int b = (int)(i / ((int64_t)keep * keep));
    const int my = origin_yx[2 * b] + ky, mx = origin_yx[2 * b + 1] + kx;
    if (my < 0 || my >= Hm || mx < 0 || mx >= Wm) continue;
    const T* row = logits + (((int64_t)b * t + ky) * t + kx) * ld;
    float best = ldf<T>(row);
    int arg = 0;
#pragma unroll
    for (int c = 1; c < CLS; ++c) {
      float v = ldf<T>(row + c);
      if (v > best) { best = v; arg = c; }
    }
    mosaic[(int64_t)my * ld_mosaic + mx] = (uint8_t)arg;
  }
}

template <typename T>
static int32_t launch_mosaic(const void* logits, int64_t ld, const int32_t* origin_yx, uint8_t* mosaic, int64_t ldm,
                             int B, int t, int margin, int Hm, int Wm, int n_classes, cudaStream_t st) {
  const int keep = t - 2 * margin;
  int64_t blocks = ((int64_t)B * keep * keep + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 8) blocks = (int64_t)kNumSMs * 8;
#define AMM(C_)                                                                                                   \
  case C_:                                                                                                        \
    argmax_mosaic_kernel<T, C_><<<(unsigned)blocks, 256, 0, st>>>((const T*)logits, ld, origin_yx, mosaic, ldm, B, t, \
                                                                  margin, Hm, Wm);                                \
    break;
  switch (n_classes) {
    AMM(2) AMM(3) AMM(4) AMM(5) AMM(6) AMM(7) AMM(8) AMM(9) AMM(10) AMM(11) AMM(12) AMM(13) AMM(14) AMM(15) AMM(16)
    default:
      set_error("argmax_mosaic: n_classes=%d unsupported (2..16)", n_classes);
      return S2LC_ERR_SHAPE;
  }
#undef AMM
  return check_launch("argmax_mosaic");
}

}  // namespace s2lc

using namespace s2lc;

extern "C" int32_t s2lc_argmax_confusion(const void* logits, int64_t ld, const int64_t* y, int64_t* counts,
                                         uint8_t* pred, int64_t rows, int32_t n_classes, int32_t dtype,
                                         void* stream) {
  S2LC_REQUIRE(logits != nullptr && rows >= 0 && ld >= n_classes, S2LC_ERR_SHAPE, "argmax_confusion: bad arguments");
  S2LC_REQUIRE((counts == nullptr) == (y == nullptr) || counts == nullptr, S2LC_ERR_SHAPE,
               "argmax_confusion: counts need labels");
  S2LC_REQUIRE(counts != nullptr || pred != nullptr, S2LC_ERR_SHAPE, "argmax_confusion: nothing to write");
  if (rows == 0) return S2LC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == S2LC_F32) return launch_argmax<float>(logits, ld, y, counts, pred, rows, n_classes, st);
  if (dtype == S2LC_BF16) return launch_argmax<s2lc::bf16>(logits, ld, y, counts, pred, rows, n_classes, st);
  set_error("argmax_confusion: bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}

extern "C" int32_t s2lc_argmax_mosaic(const void* logits, int64_t ld, const int32_t* origin_yx, uint8_t* mosaic,
                                      int64_t ld_mosaic, int32_t B, int32_t tile, int32_t margin, int32_t Hm,
                                      int32_t Wm, int32_t n_classes, int32_t dtype, void* stream) {
  S2LC_REQUIRE(logits && origin_yx && mosaic, S2LC_ERR_SHAPE, "argmax_mosaic: null pointer");
  S2LC_REQUIRE(B >= 0 && tile > 0 && margin >= 0 && 2 * margin < tile && Hm > 0 && Wm > 0 && ld >= n_classes &&
                   ld_mosaic >= Wm,
               S2LC_ERR_SHAPE, "argmax_mosaic: bad geometry (B=%d tile=%d margin=%d mosaic %dx%d)", B, tile, margin, Hm,
               Wm);
  if (B == 0) return S2LC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == S2LC_F32)
    return launch_mosaic<float>(logits, ld, origin_yx, mosaic, ld_mosaic, B, tile, margin, Hm, Wm, n_classes, st);
  if (dtype == S2LC_BF16)
    return launch_mosaic<s2lc::bf16>(logits, ld, origin_yx, mosaic, ld_mosaic, B, tile, margin, Hm, Wm, n_classes, st);
  set_error("argmax_mosaic: bad dtype %d", dtype);
  return S2LC_ERR_SHAPE;
}
