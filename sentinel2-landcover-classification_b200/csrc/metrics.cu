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
