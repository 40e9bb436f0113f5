// K15 synthetic crack images + masks generated on the device, and K17 crack measurement reductions.
//
// The reference's dataset is not shipped (client_fit_model.py:58-59 point at the author's disk), so benchmarks and
// smoke tests need crack-like data without a host round trip: concrete-like noise texture with 1-2 random-walk
// polyline "cracks" rasterised by distance-to-segment - plain arithmetic, no graphics API.  Same recipe as
// crackfl/data/generator.py::synth_crack_batch (different RNG, statistically equivalent).
//
// K17: the device half of Predict/contour (client_fit_model.py:209-221, test/Segmentation2.py:114-141):
// probability -> *255 -> uint8 -> threshold 127 -> per-image foreground area, boundary-pixel count (4-neighbour
// perimeter estimate) and number of connected-component "seeds" (pixels with no foreground west/north neighbour
// in a 2x2 scan: an upper bound used for triage).  Exact contour tracing / Douglas-Peucker / drawing stay on the
// CPU in crackfl/postproc/measure.py.
#include "common.cuh"
#include "../include/crackfl.h"

namespace {

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ float u01(uint32_t h) { return (h >> 8) * (1.0f / 16777216.0f); }

constexpr int MAX_SEG = 20;

// one block per image: thread 0 draws the polylines into shared memory, all threads rasterise
__global__ void __launch_bounds__(256) synth_cracks_k(uint8_t* __restrict__ img, uint8_t* __restrict__ mask, int H, int W,
                                                      uint32_t seed) {
  __shared__ float seg[MAX_SEG][5];  // x0,y0,x1,y1,width
  __shared__ int nseg;
  __shared__ float base;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    uint32_t s = hash32(seed * 2654435761u + b * 40503u + 17u);
    auto rnd = [&]() { s = hash32(s + 0x9e3779b9u); return u01(s); };
    base = 110.f + 80.f * rnd();
    int n = 0;
    if (rnd() > 0.15f) {
      const int ncr = 1 + (rnd() > 0.5f ? 1 : 0);
      const float scale = fmaxf((float)H, (float)W);
      for (int c = 0; c < ncr; ++c) {
        float px = rnd() * W, py = rnd() * H, ang = rnd() * 6.2831853f;
        const float wd = (0.8f + 1.7f * rnd()) * scale / 128.f;
        const int ns = 4 + (int)(rnd() * 6.f);
        for (int k = 0; k < ns && n < MAX_SEG; ++k) {
          ang += (rnd() + rnd() + rnd() - 1.5f) * 1.0f;
          const float ln = (0.08f + 0.14f * rnd()) * scale;
          const float qx = px + ln * cosf(ang), qy = py + ln * sinf(ang);
          seg[n][0] = px; seg[n][1] = py; seg[n][2] = qx; seg[n][3] = qy; seg[n][4] = wd;
          ++n; px = qx; py = qy;
        }
      }
    }
    nseg = n;
  }
  __syncthreads();
  const int n = nseg;
  for (int p = threadIdx.x; p < H * W; p += blockDim.x) {
    const float x = (float)(p % W), y = (float)(p / W);
    bool in = false;
    for (int k = 0; k < n; ++k) {
      const float dx = seg[k][2] - seg[k][0], dy = seg[k][3] - seg[k][1];
      float t = ((x - seg[k][0]) * dx + (y - seg[k][1]) * dy) / (dx * dx + dy * dy + 1e-6f);
      t = fminf(fmaxf(t, 0.f), 1.f);
      const float ex = x - (seg[k][0] + t * dx), ey = y - (seg[k][1] + t * dy);
      in = in || (ex * ex + ey * ey <= seg[k][4] * seg[k][4]);
    }
    const uint32_t h = hash32(seed ^ (uint32_t)(b * 7919 + 13) ^ ((uint32_t)p * 2246822519u));
    const float g = 12.f * (u01(h) + u01(hash32(h)) + u01(hash32(h + 1)) - 1.5f) * 2.f;
    const long o = ((long)b * H * W + p);
    for (int c = 0; c < 3; ++c) {
      float v = base + g + 4.f * (u01(hash32(h + 11u * (c + 1))) - 0.5f) * 2.f;
      if (in) v *= 0.35f;
      img[o * 3 + c] = (uint8_t)fminf(fmaxf(v, 0.f), 255.f);
    }
    mask[o] = in ? 1 : 0;
  }
}

// out[b] = {foreground area, boundary pixels, component seeds, sum of probabilities}
__global__ void __launch_bounds__(256) measure_masks_k(const float* __restrict__ prob, float* __restrict__ out, int H, int W,
                                                       int threshold) {
  // grid (B, slices): each block reduces a band of rows and adds its four partial sums to the (pre-zeroed) output
  const int b = blockIdx.x;
  const float* p = prob + (long)b * H * W;
  const int rows = (H + gridDim.y - 1) / gridDim.y;
  const int y0 = blockIdx.y * rows, y1 = min(H, y0 + rows);
  auto fg = [&](int y, int x) -> bool {
    if (y < 0 || y >= H || x < 0 || x >= W) return false;
    return (int)(p[y * W + x] * 255.f) > threshold;   // (pred*255).astype(uint8) then threshold(127, 255, BINARY)
  };
  float a[4] = {0, 0, 0, 0};
  for (int i = y0 * W + threadIdx.x; i < y1 * W; i += blockDim.x) {
    const int y = i / W, x = i % W;
    a[3] += p[i];
    if (!fg(y, x)) continue;
    a[0] += 1.f;
    if (!fg(y - 1, x) || !fg(y + 1, x) || !fg(y, x - 1) || !fg(y, x + 1)) a[1] += 1.f;
    if (!fg(y - 1, x) && !fg(y, x - 1) && !fg(y - 1, x - 1) && !fg(y - 1, x + 1)) a[2] += 1.f;
  }
  __shared__ float red[4][8];
  for (int k = 0; k < 4; ++k) {
    const float v = ck::warp_sum(a[k]);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    float v = 0.f;
    for (int i = 0; i < 8; ++i) v += red[threadIdx.x][i];
    atomicAdd(out + b * 4 + threadIdx.x, v);
  }
}

}  // namespace

extern "C" {

int crackfl_synth_cracks(void* img_u8, void* mask_u8, int B, int H, int W, unsigned seed, cudaStream_t st) {
  synth_cracks_k<<<B, 256, 0, st>>>((uint8_t*)img_u8, (uint8_t*)mask_u8, H, W, seed);
  CK_LAUNCH_CHECK();
  return 0;
}

int crackfl_measure_masks(const float* prob, float* out4, int B, int H, int W, int threshold, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(out4, 0, (size_t)B * 4 * sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  int slices = (148 * 4 + B - 1) / B;  // about four blocks per SM over the whole batch (one block per image used 16 SMs)
  if (slices > H) slices = H;
  if (slices < 1) slices = 1;
  measure_masks_k<<<dim3(B, slices), 256, 0, st>>>(prob, out4, H, W, threshold);
  CK_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
