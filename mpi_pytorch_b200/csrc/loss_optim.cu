// K10 softmax-cross-entropy (fused forward+backward), flat Adam/SGD (K1's local half),
// K11 argmax+compare+count.  Replaces the ATen CPU log_softmax+nll_loss and
// _single_tensor_adam the reference reaches from main.py:134,150,152,155.
#include "common.cuh"

using namespace mpb;

// ---------------------------------------------------------------------------------------------
// Softmax-CE: one CTA per row.  Pass 1 = online (max, sum) over the row, pass 2 = write
// dlogits = (softmax - onehot) * inv_batch.  Rows up to 64,500 classes stay L2-resident
// between the passes (33 MB for 256 rows in bf16 << 126 MB L2).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) softmax_ce_kernel(const T* __restrict__ logits, const long long* __restrict__ target,
                                                         T* __restrict__ dlogits, float* __restrict__ loss_sum,
                                                         float* __restrict__ row_loss, int C, float inv_batch) {
  __shared__ float red[32];
  const long long row = blockIdx.x;
  const T* x = logits + row * (long long)C;
  T* dx = dlogits + row * (long long)C;
  const int tid = threadIdx.x;
  // Rows need not be vector-aligned (64,500 classes: a bf16 row is 129,000 B, so odd rows start 8 B off): scalar head up
  // to the first 8-element boundary that is 32-byte (fp32) / 16-byte (bf16) aligned, vector body, scalar tail.
  constexpr int kVecBytes = 8 * (int)sizeof(T);
  int head = (int)(((kVecBytes - (reinterpret_cast<uintptr_t>(x) % kVecBytes)) % kVecBytes) / sizeof(T));
  if ((reinterpret_cast<uintptr_t>(dx) % kVecBytes) != (reinterpret_cast<uintptr_t>(x) % kVecBytes) ||
      (reinterpret_cast<uintptr_t>(x) % sizeof(T)) != 0)
    head = C;                                                    // differently aligned output: all scalar
  if (head > C) head = C;
  const int body_end = head + ((C - head) / 8) * 8;
  float m = -INFINITY, s = 0.f;
  auto scalar_stat = [&](int i) {
    const float v = to_f(x[i]);
    const float nm = fmaxf(m, v);
    s = s * __expf(m - nm) + __expf(v - nm);
    m = nm;
  };
  for (int i = tid; i < head; i += blockDim.x) scalar_stat(i);
  for (int i = head + tid * 8; i < body_end; i += blockDim.x * 8) {
    float v[8];
    load8(x + i, v);
    float lm = v[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) lm = fmaxf(lm, v[k]);
    const float nm = fmaxf(m, lm);
    float ls = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) ls += __expf(v[k] - nm);
    s = s * __expf(m - nm) + ls;
    m = nm;
  }
  for (int i = body_end + tid; i < C; i += blockDim.x) scalar_stat(i);
  const float gm = block_max(m, red);
  const float contrib = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_sum(contrib, red);
  const long long t = target[row];
  const float inv_s = 1.f / gs;
  if (tid == 0) {
    const float l = (__logf(gs) + gm) - to_f(x[t]);
    if (row_loss) row_loss[row] = l;
    atomicAdd(loss_sum, l * inv_batch);
  }
  auto scalar_grad = [&](int i) {
    float p = __expf(to_f(x[i]) - gm) * inv_s;
    if (i == t) p -= 1.f;
    dx[i] = from_f<T>(p * inv_batch);
  };
  for (int i = tid; i < head; i += blockDim.x) scalar_grad(i);
  for (int i = head + tid * 8; i < body_end; i += blockDim.x * 8) {
    float v[8];
    load8(x + i, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float p = __expf(v[k] - gm) * inv_s;
      if (i + k == t) p -= 1.f;
      v[k] = p * inv_batch;
    }
    store8(dx + i, v);
  }
  for (int i = body_end + tid; i < C; i += blockDim.x) scalar_grad(i);
}

// loss_sum: 1 float (zeroed here); dlogits same dtype/shape as logits; row_loss optional [B] fp32.
MPB_API int mpb_softmax_ce(const void* logits, const long long* target, void* dlogits, float* loss_sum,
                           float* row_loss, int B, int C, int dtype, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  MPB_CUDA(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
  if (B <= 0) return 0;
  if (dtype == MPB_BF16)
    softmax_ce_kernel<__nv_bfloat16><<<B, 256, 0, st>>>((const __nv_bfloat16*)logits, target, (__nv_bfloat16*)dlogits,
                                                        loss_sum, row_loss, C, 1.f / B);
  else
    softmax_ce_kernel<float><<<B, 256, 0, st>>>((const float*)logits, target, (float*)dlogits, loss_sum, row_loss, C,
                                                1.f / B);
  MPB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// argmax + compare + count (one warp per row).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void count_correct_kernel(const T* __restrict__ logits, const long long* __restrict__ target,
                                     unsigned long long* __restrict__ count, int B, int C) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const T* x = logits + (long long)warp * C;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = lane; i < C; i += 32) {
    const float v = to_f(x[i]);
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  if (lane == 0 && (long long)bi == target[warp]) atomicAdd(count, 1ull);
}

MPB_API int mpb_count_correct(const void* logits, const long long* target, long long* count, int B, int C, int dtype,
                              void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  MPB_CUDA(cudaMemsetAsync(count, 0, sizeof(long long), st));
  if (B <= 0) return 0;
  const int threads = 128, blocks = (B * 32 + threads - 1) / threads;
  if (dtype == MPB_BF16)
    count_correct_kernel<__nv_bfloat16><<<blocks, threads, 0, st>>>((const __nv_bfloat16*)logits, target,
                                                                    (unsigned long long*)count, B, C);
  else
    count_correct_kernel<float><<<blocks, threads, 0, st>>>((const float*)logits, target, (unsigned long long*)count, B, C);
  MPB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Flat Adam / SGD-momentum over an arena slice (fp32 master params, fp32 grads, optional bf16
// shadow copy of the updated params for the compute path).  Same math as torch.optim.Adam
// (non-amsgrad, L2 weight decay) / torch.optim.SGD.
// ---------------------------------------------------------------------------------------------
struct AdamArgs { float lr, b1, b2, eps, wd, bc1, bc2_sqrt, gscale; };

__device__ __forceinline__ float adam_one(float p, float g, float& m, float& v, const AdamArgs& a) {
  g *= a.gscale;
  if (a.wd != 0.f) g += a.wd * p;
  m = a.b1 * m + (1.f - a.b1) * g;
  v = a.b2 * v + (1.f - a.b2) * g * g;
  const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
  return p - (a.lr / a.bc1) * (m / denom);
}

__global__ void __launch_bounds__(256) adam_flat_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        __nv_bfloat16* __restrict__ shadow, long long n, AdamArgs a) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    pp.x = adam_one(pp.x, gg.x, mm.x, vv.x, a);
    pp.y = adam_one(pp.y, gg.y, mm.y, vv.y, a);
    pp.z = adam_one(pp.z, gg.z, mm.z, vv.z, a);
    pp.w = adam_one(pp.w, gg.w, mm.w, vv.w, a);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (shadow) {
      __nv_bfloat162* s2 = reinterpret_cast<__nv_bfloat162*>(shadow) + 2 * i;
      s2[0] = __floats2bfloat162_rn(pp.x, pp.y);
      s2[1] = __floats2bfloat162_rn(pp.z, pp.w);
    }
  }
  // tail (n not multiple of 4): single thread
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (long long i = n4 << 2; i < n; ++i) {
      float mm = m[i], vv = v[i];
      const float np = adam_one(p[i], g[i], mm, vv, a);
      p[i] = np; m[i] = mm; v[i] = vv;
      if (shadow) shadow[i] = __float2bfloat16_rn(np);
    }
  }
}

MPB_API int mpb_adam_flat(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n, int step,
                          float lr, float b1, float b2, float eps, float wd, float grad_scale, void* stream) {
  if (n <= 0) return 0;
  AdamArgs a;
  a.lr = lr; a.b1 = b1; a.b2 = b2; a.eps = eps; a.wd = wd; a.gscale = grad_scale;
  a.bc1 = 1.f - powf(b1, (float)step);
  a.bc2_sqrt = sqrtf(1.f - powf(b2, (float)step));
  adam_flat_kernel<<<grid_for(n / 4 + 1, 256, 4), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (__nv_bfloat16*)shadow_bf16, n, a);
  MPB_CHECK_LAUNCH();
  return 0;
}

__global__ void __launch_bounds__(256) sgd_flat_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ buf, __nv_bfloat16* __restrict__ shadow,
                                                       long long n, float lr, float mom, float wd, float gscale, int first) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gg = g[i] * gscale;
    const float pp = p[i];
    if (wd != 0.f) gg += wd * pp;
    if (mom != 0.f) {
      const float b = first ? gg : mom * buf[i] + gg;
      buf[i] = b;
      gg = b;
    }
    const float np = pp - lr * gg;
    p[i] = np;
    if (shadow) shadow[i] = __float2bfloat16_rn(np);
  }
}

MPB_API int mpb_sgd_flat(float* p, const float* g, float* buf, void* shadow_bf16, long long n, int step, float lr,
                         float momentum, float wd, float grad_scale, void* stream) {
  if (n <= 0) return 0;
  sgd_flat_kernel<<<grid_for(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(p, g, buf, (__nv_bfloat16*)shadow_bf16, n, lr,
                                                                         momentum, wd, grad_scale, step == 1);
  MPB_CHECK_LAUNCH();
  return 0;
}

// fp32 -> bf16 cast of a flat range (shadow refresh after a load / broadcast).
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = __float2bfloat16_rn(src[i]);
}
MPB_API int mpb_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  if (n <= 0) return 0;
  cast_f32_bf16_kernel<<<grid_for(n, 256, 4), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n);
  MPB_CHECK_LAUNCH();
  return 0;
}
