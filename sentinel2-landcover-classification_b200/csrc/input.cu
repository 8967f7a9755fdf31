// input.cu — the input pipeline on the device: raw int16 digital-number tiles + uint8 label
// rasters -> cropped / flipped / normalised fp32 NCHW batch + remapped int64 labels, one pass.
//
// Reference (all on the CPU, per sample, inside the DataLoader workers):
//   src/data/s2osm_dataset.py:51-71       read tile + first label band, remap labels, albumentations, .float()/.long()
//   src/data/s2osm_datamodule.py:75-87    RandomCrop | CenterCrop -> HorizontalFlip -> VerticalFlip -> Normalize(mean, std)
//                                         (A.Normalize default max_pixel_value = 255: (x - 255*mean) * (1 / (255*std)))
//   src/configs/cnes_labell_mappings.py:85-95   np.vectorize(map_func) label remap == a 256-entry lookup table
// HBM-bound: C x 2 B + 1 B read, C x 4 B + 8 B written per output pixel.  The random draws (crop offsets, flips) stay
// on the host so any RNG policy of the caller is reproduced exactly; they arrive as small per-sample arrays.
#include "common.cuh"

namespace s2lc {

__global__ void __launch_bounds__(256) prepare_batch_kernel(const int16_t* __restrict__ raw, const uint8_t* __restrict__ lab,
                                                            const int32_t* __restrict__ crop_yx,
                                                            const uint8_t* __restrict__ flip_hv,
                                                            const float* __restrict__ mean, const float* __restrict__ stdv,
                                                            const int64_t* __restrict__ lut, float* __restrict__ x,
                                                            int64_t* __restrict__ y, int N, int C, int Hs, int Ws,
                                                            int crop, float max_pixel_value) {
  const int64_t total = (int64_t)N * crop * crop;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % crop);
    const int oy = (int)((i / crop) % crop);
    const int n = (int)(i / ((int64_t)crop * crop));
    const int y0 = crop_yx[2 * n], x0 = crop_yx[2 * n + 1];
    // flips act on the CROPPED image: output (oy, ox) comes from cropped (crop-1-oy | oy, crop-1-ox | ox)
    const int cx = flip_hv && flip_hv[2 * n] ? crop - 1 - ox : ox;
    const int cy = flip_hv && flip_hv[2 * n + 1] ? crop - 1 - oy : oy;
    const int64_t sp = (int64_t)(y0 + cy) * Ws + (x0 + cx);
    if (x) {
      for (int c = 0; c < C; ++c) {
        const float v = (float)raw[((int64_t)n * C + c) * Hs * Ws + sp];
        // albumentations: img -= mean * max_pixel_value; img *= reciprocal(std * max_pixel_value)  (fp32)
        const float m = mean[c] * max_pixel_value;
        const float r = 1.0f / (stdv[c] * max_pixel_value);
        x[(((int64_t)n * C + c) * crop + oy) * crop + ox] = (v - m) * r;
      }
    }
    if (y) {
      const uint8_t l = lab[(int64_t)n * Hs * Ws + sp];
      y[i] = lut ? lut[l] : (int64_t)l;
    }
  }
}

// Mosaic inference (row f4): cut B tiles of t x t pixels out of ONE resident int16 scene [C, Hs, Ws] at arbitrary
// (possibly negative / overhanging) origins and normalise them.  Pixels outside the scene replicate the nearest edge
// pixel, so border tiles see plausible context instead of a zero cliff; the mosaic writer never keeps them.
__global__ void __launch_bounds__(256) gather_tiles_kernel(const int16_t* __restrict__ scene,
                                                           const int32_t* __restrict__ origin_yx,
                                                           const float* __restrict__ mean, const float* __restrict__ stdv,
                                                           float* __restrict__ x, int B, int C, int Hs, int Ws, int t,
                                                           float max_pixel_value) {
  const int64_t total = (int64_t)B * t * t;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % t);
    const int oy = (int)((i / t) % t);
    const int b = (int)(i / ((int64_t)t * t));
    int sy = origin_yx[2 * b] + oy, sx = origin_yx[2 * b + 1] + ox;
    sy = sy < 0 ? 0 : (sy >= Hs ? Hs - 1 : sy);
    sx = sx < 0 ? 0 : (sx >= Ws ? Ws - 1 : sx);
    const int64_t sp = (int64_t)sy * Ws + sx;
    for (int c = 0; c < C; ++c) {
      const float v = (float)scene[(int64_t)c * Hs * Ws + sp];
      const float m = mean[c] * max_pixel_value;
      const float r = 1.0f / (stdv[c] * max_pixel_value);
      x[(((int64_t)b * C + c) * t + oy) * t + ox] = (v - m) * r;
    }
  }
}

}  // namespace s2lc

using namespace s2lc;

extern "C" int32_t s2lc_prepare_batch(const int16_t* raw, const uint8_t* labels, const int32_t* crop_yx,
                                      const uint8_t* flip_hv, const float* mean, const float* stdv, const int64_t* lut,
                                      float* x, int64_t* y, int32_t N, int32_t C, int32_t Hs, int32_t Ws, int32_t crop,
                                      float max_pixel_value, void* stream) {
  S2LC_REQUIRE(crop_yx != nullptr && N >= 0 && C > 0 && crop > 0 && crop <= Hs && crop <= Ws, S2LC_ERR_SHAPE,
               "prepare_batch: bad geometry (N=%d C=%d %dx%d crop=%d)", N, C, Hs, Ws, crop);
  S2LC_REQUIRE((x == nullptr) || (raw && mean && stdv), S2LC_ERR_SHAPE, "prepare_batch: x needs raw, mean and std");
  S2LC_REQUIRE((y == nullptr) || labels, S2LC_ERR_SHAPE, "prepare_batch: y needs labels");
  if (N == 0) return S2LC_OK;
  const int64_t total = (int64_t)N * crop * crop;
  int64_t blocks = (total + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  prepare_batch_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(raw, labels, crop_yx, flip_hv, mean, stdv, lut,
                                                                           x, y, N, C, Hs, Ws, crop, max_pixel_value);
  return check_launch("prepare_batch");
}

extern "C" int32_t s2lc_gather_tiles(const int16_t* scene, const int32_t* origin_yx, const float* mean,
                                     const float* stdv, float* x, int32_t B, int32_t C, int32_t Hs, int32_t Ws,
                                     int32_t tile, float max_pixel_value, void* stream) {
  S2LC_REQUIRE(scene && origin_yx && mean && stdv && x, S2LC_ERR_SHAPE, "gather_tiles: null pointer");
  S2LC_REQUIRE(B >= 0 && C > 0 && Hs > 0 && Ws > 0 && tile > 0, S2LC_ERR_SHAPE,
               "gather_tiles: bad geometry (B=%d C=%d %dx%d tile=%d)", B, C, Hs, Ws, tile);
  if (B == 0) return S2LC_OK;
  const int64_t total = (int64_t)B * tile * tile;
  int64_t blocks = (total + 255) / 256;
  if (blocks > (int64_t)kNumSMs * 16) blocks = (int64_t)kNumSMs * 16;
  gather_tiles_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(scene, origin_yx, mean, stdv, x, B, C, Hs, Ws,
                                                                          tile, max_pixel_value);
  return check_launch("gather_tiles");
}
