// K9 fused resize + normalise (+ layout / cast) and K14 synthetic data generation.
//
// The reference resizes and normalises on the CPU per sample: train = ToTensor -> Resize (bilinear on
// the tensor) -> Normalize (main.py:62-65); eval pipeline = PIL resize (bicubic) on rank 1, then
// ToTensor + Normalize on rank 2 (evaluation_pipeline.py:89,116-122).  Here one kernel reads the
// uint8 HWC frame and writes the normalised CHW / NHWC fp32 / bf16 tensor; the two pipeline stages
// are also available separately (resize u8->u8, normalise u8->float) to keep the stage structure.
// Filtering is the antialiased separable kernel PIL / torch(antialias=True) use:
//   scale = in/out, support = r*max(scale,1), centre = (o+0.5)*scale, taps j in [xmin, xmin+n),
//   w_j = f((j - centre + 0.5)/max(scale,1)), normalised;  r=1 triangle (bilinear), r=2 Keys a=-0.5.
#include "common.cuh"

using namespace mpb;

namespace {

__device__ __forceinline__ float filt(float x, int bicubic) {
  x = fabsf(x);
  if (!bicubic) return x < 1.f ? 1.f - x : 0.f;
  const float a = -0.5f;
  if (x < 1.f) return ((a + 2.f) * x - (a + 3.f)) * x * x + 1.f;
  if (x < 2.f) return (((x - 5.f) * x + 8.f) * x - 4.f) * a;
  return 0.f;
}

struct Taps {
  int lo, n;
  float centre, inv, norm;
};
__device__ __forceinline__ Taps make_taps(int o, int in_size, int out_size, int bicubic) {
  const float scale = (float)in_size / (float)out_size;
  const float fs = fmaxf(scale, 1.f);
  const float support = (bicubic ? 2.f : 1.f) * fs;
  Taps t;
  t.centre = scale * ((float)o + 0.5f);
  t.inv = 1.f / fs;
  t.lo = max(0, (int)(t.centre - support + 0.5f));
  t.n = min(in_size, (int)(t.centre + support + 0.5f)) - t.lo;
  float s = 0.f;
  for (int j = 0; j < t.n; ++j) s += filt(((float)(j + t.lo) - t.centre + 0.5f) * t.inv, bicubic);
  t.norm = s != 0.f ? 1.f / s : 0.f;
  return t;
}

__constant__ float kMean[3] = {0.485f, 0.456f, 0.406f};
__constant__ float kStd[3] = {0.229f, 0.224f, 0.225f};

// mode: 0 = u8 -> u8 NHWC (round, clamp)            [pipeline stage 1]
//       1 = u8 -> normalised float, quantise through u8 first (eval semantics: PIL resize then ToTensor)
//       2 = u8 -> normalised float directly (train semantics: resize acts on the float tensor)
template <typename OutT>
__global__ void __launch_bounds__(128) resize_kernel(const unsigned char* __restrict__ in, OutT* __restrict__ out, int N,
                                                     int Hi, int Wi, int Ho, int Wo, int bicubic, int mode,
                                                     int out_nhwc) {
  const long long total = (long long)N * Ho * Wo;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int wo = (int)(i % Wo);
    const int ho = (int)((i / Wo) % Ho);
    const int n = (int)(i / ((long long)Wo * Ho));
    const Taps ty = make_taps(ho, Hi, Ho, bicubic), tx = make_taps(wo, Wi, Wo, bicubic);
    float acc[3] = {0.f, 0.f, 0.f};
    const unsigned char* base = in + (long long)n * Hi * Wi * 3;
    for (int a = 0; a < ty.n; ++a) {
      const float wy = filt(((float)(a + ty.lo) - ty.centre + 0.5f) * ty.inv, bicubic) * ty.norm;
      const unsigned char* row = base + ((long long)(a + ty.lo) * Wi + tx.lo) * 3;
      float r0 = 0.f, r1 = 0.f, r2 = 0.f;
      for (int b = 0; b < tx.n; ++b) {
        const float wx = filt(((float)(b + tx.lo) - tx.centre + 0.5f) * tx.inv, bicubic);
        r0 += wx * (float)row[b * 3 + 0];
        r1 += wx * (float)row[b * 3 + 1];
        r2 += wx * (float)row[b * 3 + 2];
      }
      acc[0] += wy * r0 * tx.norm; acc[1] += wy * r1 * tx.norm; acc[2] += wy * r2 * tx.norm;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float v = acc[c];
      if (mode != 2) v = fminf(fmaxf(rintf(v), 0.f), 255.f);
      if (mode == 0) {
        reinterpret_cast<unsigned char*>(out)[i * 3 + c] = (unsigned char)v;
      } else {
        const float f = (v * (1.f / 255.f) - kMean[c]) / kStd[c];
        const long long o = out_nhwc ? (i * 3 + c) : (((long long)n * 3 + c) * Ho + ho) * Wo + wo;
        out[o] = from_f<OutT>(f);
      }
    }
  }
}

template <typename OutT>
__global__ void __launch_bounds__(256) normalize_kernel(const unsigned char* __restrict__ in, OutT* __restrict__ out, int N,
                                                        int H, int W, int out_nhwc) {
  const long long total = (long long)N * H * W;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int w = (int)(i % W);
    const int h = (int)((i / W) % H);
    const int n = (int)(i / ((long long)W * H));
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float f = ((float)in[i * 3 + c] * (1.f / 255.f) - kMean[c]) / kStd[c];
      const long long o = out_nhwc ? (i * 3 + c) : (((long long)n * 3 + c) * H + h) * W + w;
      out[o] = from_f<OutT>(f);
    }
  }
}

// Synthetic images: value(sample, e) with e the NCHW element index; must match data/synthetic.py.
template <typename OutT>
__global__ void __launch_bounds__(256) synth_kernel(OutT* __restrict__ out, const long long* __restrict__ index, int n,
                                                    int C, int H, int W, uint32_t seed_mix, int out_nhwc) {
  const long long chw = (long long)C * H * W, total = (long long)n * chw;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int s = (int)(i / chw);
    const long long e_out = i % chw;
    long long e;  // logical NCHW element index of the value stored at physical position i
    if (out_nhwc) {
      const int c = (int)(e_out % C);
      const long long hw = e_out / C;
      e = (long long)c * H * W + hw;
    } else {
      e = e_out;
    }
    const unsigned long long ctr = (unsigned long long)index[s] * (unsigned long long)chw + (unsigned long long)e;
    const uint32_t h = hash32(((uint32_t)ctr) ^ seed_mix);
    const double u = (double)h * (1.0 / 4294967296.0);
    out[i] = from_f<OutT>((float)((u * 2.0 - 1.0) * 1.7));
  }
}

}  // namespace

// out_dtype: MPB_F32 / MPB_BF16 (ignored for mode 0 which writes uint8)
MPB_API int mpb_resize(const unsigned char* in, void* out, int N, int Hi, int Wi, int Ho, int Wo, int bicubic, int mode,
                       int out_nhwc, int out_dtype, void* stream) {
  const long long total = (long long)N * Ho * Wo;
  if (total <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for(total, 128, 16);
  if (mode == 0 || out_dtype == MPB_F32)
    resize_kernel<float><<<grid, 128, 0, st>>>(in, (float*)out, N, Hi, Wi, Ho, Wo, bicubic, mode, out_nhwc);
  else
    resize_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(in, (__nv_bfloat16*)out, N, Hi, Wi, Ho, Wo, bicubic, mode, out_nhwc);
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_normalize_u8(const unsigned char* in, void* out, int N, int H, int W, int out_nhwc, int out_dtype,
                             void* stream) {
  const long long total = (long long)N * H * W;
  if (total <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == MPB_F32)
    normalize_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(in, (float*)out, N, H, W, out_nhwc);
  else
    normalize_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, st>>>(in, (__nv_bfloat16*)out, N, H, W, out_nhwc);
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_synth_images(void* out, const long long* index, int n, int C, int H, int W, long long seed, int out_nhwc,
                             int out_dtype, void* stream) {
  const long long total = (long long)n * C * H * W;
  if (total <= 0) return 0;
  const uint32_t mix = (uint32_t)(((unsigned long long)seed * 0x9E3779B9ull) & 0xFFFFFFFFull);
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == MPB_F32)
    synth_kernel<float><<<grid_for(total, 256), 256, 0, st>>>((float*)out, index, n, C, H, W, mix, out_nhwc);
  else
    synth_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, st>>>((__nv_bfloat16*)out, index, n, C, H, W, mix, out_nhwc);
  MPB_CHECK_LAUNCH();
  return 0;
}
