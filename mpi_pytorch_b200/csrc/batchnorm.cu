// K7: BatchNorm (+ residual add) (+ ReLU), NHWC, training forward/backward and eval affine.
// Replaces the ATen CPU batch_norm + relu + add the reference runs for every ResNet / VGG-bn /
// DenseNet / Inception block.  x is [M = N*H*W, C] with C contiguous.
//
//  forward (train):  stats  : sums[0:C] = sum_m x, sums[C:2C] = sum_m x^2   (skipped when the conv
//                             epilogue already produced them)
//                    apply  : out = relu( (x-mean)*invstd*gamma + beta + residual ), saves mean/invstd,
//                             updates running_mean / running_var (unbiased) like torch
//  backward        : reduce : dsum[0:C] = sum dzm,  dsum[C:2C] = sum dzm * xhat,  dzm = dz * (out > 0)
//                    apply  : dx = gamma*invstd*(dzm - dsum0/M - xhat*dsum1/M);  dres = dzm
// Thread mapping: 8 channels per thread (16 B bf16 vectors), blockDim = cvecs * rows_per_block so a
// thread keeps the same channel group for its whole row loop and can accumulate in registers.
#include "common.cuh"
#include "bn_stream.cuh"

using namespace mpb;

namespace {

struct Map {
  int cvecs, rows_per_block, threads;
};
inline Map make_map(int C) {
  Map m;
  m.cvecs = C / 8;
  m.rows_per_block = 256 / m.cvecs;
  if (m.rows_per_block < 1) m.rows_per_block = 1;
  m.threads = m.cvecs * m.rows_per_block;
  return m;
}
inline int rows_grid(long long M, int rows_per_block, int waves) {
  long long b = (M + rows_per_block - 1) / rows_per_block;
  const long long cap = (long long)kSMs * waves;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

// Reduce per-thread acc[8] x2 over the rows_per_block threads sharing a channel group, then atomics.
__device__ __forceinline__ void reduce_pair_to_global(float (&a)[8], float (&b)[8], float* g0, float* g1, int cvecs,
                                                      int rows_per_block, float* smem /* threads*16 floats */) {
  const int t = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 8; ++k) { smem[t * 16 + k] = a[k]; smem[t * 16 + 8 + k] = b[k]; }
  __syncthreads();
  // cvecs*16 output values; thread u handles value u (and strides) summing over rows_per_block partials
  for (int u = t; u < cvecs * 16; u += blockDim.x) {
    const int cv = u >> 4, k = u & 15;
    float s = 0.f;
    for (int r = 0; r < rows_per_block; ++r) s += smem[(r * cvecs + cv) * 16 + k];
    if (k < 8) atomicAdd(g0 + cv * 8 + k, s);
    else atomicAdd(g1 + cv * 8 + (k - 8), s);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) bn_stats_kernel(const T* __restrict__ x, float* __restrict__ sums, long long M,
                                                       int C, int cvecs, int rows_per_block) {
  extern __shared__ float smem[];
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float s[8] = {0}, q[8] = {0};
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float v[8];
    load8(x + m * C + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] += v[k] * v[k]; }
  }
  reduce_pair_to_global(s, q, sums, sums + C, cvecs, rows_per_block, smem);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_fwd_apply_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                           T* __restrict__ out, const float* __restrict__ sums,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           float* __restrict__ run_mean, float* __restrict__ run_var,
                                                           float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                           long long M, int C, int cvecs, int rows_per_block, float eps,
                                                           float momentum, int relu, int training,
                                                           long long* __restrict__ num_batches) {
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  if (num_batches && training && blockIdx.x == 0 && threadIdx.x == 0) *num_batches += 1;
  float sc[8], sh[8];
  const float invM = 1.f / (float)M;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    float mean, var;
    if (training) {
      mean = sums[c] * invM;
      var = fmaxf(sums[C + c] * invM - mean * mean, 0.f);
    } else {
      mean = run_mean[c];
      var = run_var[c];
    }
    const float invstd = rsqrtf(var + eps);
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    sc[k] = g * invstd;
    sh[k] = b - mean * sc[k];
    if (training && blockIdx.x == 0 && r == 0) {
      save_mean[c] = mean;
      save_invstd[c] = invstd;
      if (run_mean) {
        const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
      }
    }
  }
  // one row per trip: these streaming kernels are occupancy-bound (a 2-row unroll measured slower on B200)
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float v[8];
    load8(x + m * C + cv * 8, v);
    if (res) {
      float rr[8];
      load8(res + m * C + cv * 8, rr);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], sc[k], sh[k]) + rr[k];
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], sc[k], sh[k]);
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
    }
    store8(out + m * C + cv * 8, v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const T* __restrict__ dz, const T* __restrict__ z,
                                                            const T* __restrict__ x, const float* __restrict__ save_mean,
                                                            const float* __restrict__ save_invstd,
                                                            float* __restrict__ dsum, long long M, int C, int cvecs,
                                                            int rows_per_block, int relu) {
  extern __shared__ float smem[];
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float mean[8], istd[8], s0[8] = {0}, s1[8] = {0};
#pragma unroll
  for (int k = 0; k < 8; ++k) { mean[k] = save_mean[cv * 8 + k]; istd[k] = save_invstd[cv * 8 + k]; }
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float g[8], xv[8];
    load8(dz + m * C + cv * 8, g);
    load8(x + m * C + cv * 8, xv);
    if (relu) {
      float zv[8];
      load8(z + m * C + cv * 8, zv);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = zv[k] > 0.f ? g[k] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) { s0[k] += g[k]; s1[k] += g[k] * (xv[k] - mean[k]) * istd[k]; }
  }
  reduce_pair_to_global(s0, s1, dsum, dsum + C, cvecs, rows_per_block, smem);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const T* __restrict__ dz, const T* __restrict__ z,
                                                           const T* __restrict__ x, const float* __restrict__ gamma,
                                                           const float* __restrict__ save_mean,
                                                           const float* __restrict__ save_invstd,
                                                           const float* __restrict__ dsum, T* __restrict__ dx,
                                                           T* __restrict__ dres, float* __restrict__ dgamma,
                                                           float* __restrict__ dbeta, long long M, int C, int cvecs,
                                                           int rows_per_block, int relu, int accumulate) {
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float mean[8], istd[8], a[8], b0[8], b1[8];
  const float invM = 1.f / (float)M;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    mean[k] = save_mean[c];
    istd[k] = save_invstd[c];
    const float g = gamma ? gamma[c] : 1.f;
    a[k] = g * istd[k];
    b0[k] = dsum[c] * invM;
    b1[k] = dsum[C + c] * invM;
    if (blockIdx.x == 0 && r == 0) {
      if (dgamma) dgamma[c] = dsum[C + c] + (accumulate ? dgamma[c] : 0.f);
      if (dbeta) dbeta[c] = dsum[c] + (accumulate ? dbeta[c] : 0.f);
    }
  }
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float g[8], xv[8];
    load8(dz + m * C + cv * 8, g);
    load8(x + m * C + cv * 8, xv);
    if (relu) {
      float zv[8];
      load8(z + m * C + cv * 8, zv);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = zv[k] > 0.f ? g[k] : 0.f;
    }
    if (dres) store8(dres + m * C + cv * 8, g);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float xhat = (xv[k] - mean[k]) * istd[k];
      xv[k] = a[k] * (g[k] - b0[k] - xhat * b1[k]);
    }
    store8(dx + m * C + cv * 8, xv);
  }
}

}  // namespace

#define DISPATCH_T(dtype, ...)                               \
  if (dtype == MPB_BF16) { using T = __nv_bfloat16; __VA_ARGS__ } \
  else { using T = float; __VA_ARGS__ }

// sums: 2*C floats, zeroed here.
MPB_API int mpb_bn_stats(const void* x, float* sums, long long M, int C, int dtype, void* stream) {
  if (C % 8 != 0 || C > 2048) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  MPB_CUDA(cudaMemsetAsync(sums, 0, 2 * C * sizeof(float), st));
  if (mpbbn::stream_ok(M, C, dtype)) return mpbbn::stream_stats(x, sums, M, C, st);
  const Map mp = make_map(C);
  const int grid = rows_grid(M, mp.rows_per_block, 4);
  const size_t smem = (size_t)mp.threads * 16 * sizeof(float);
  DISPATCH_T(dtype, bn_stats_kernel<T><<<grid, mp.threads, smem, st>>>((const T*)x, sums, M, C, mp.cvecs, mp.rows_per_block);)
  MPB_CHECK_LAUNCH();
  return 0;
}

// `num_batches` (may be null): the module's int64 num_batches_tracked counter, bumped by the kernel in training mode
// so the step issues no separate elementwise launch for it.
MPB_API int mpb_bn_fwd_apply_ex(const void* x, const void* res, void* out, const float* sums, const float* gamma,
                                const float* beta, float* run_mean, float* run_var, float* save_mean,
                                float* save_invstd, long long* num_batches, long long M, int C, float eps,
                                float momentum, int relu, int training, int dtype, void* stream) {
  if (C % 8 != 0 || C > 2048) return -1;
  if (mpbbn::stream_ok(M, C, dtype))      // large bf16 tensors: bulk-copy streaming form (bn_stream.cu)
    return mpbbn::stream_fwd(x, res, out, sums, gamma, beta, run_mean, run_var, save_mean, save_invstd, num_batches, M, C,
                             eps, momentum, relu, training, (cudaStream_t)stream);
  const Map mp = make_map(C);
  const int grid = rows_grid(M, mp.rows_per_block, 8);
  DISPATCH_T(dtype, bn_fwd_apply_kernel<T><<<grid, mp.threads, 0, (cudaStream_t)stream>>>(
                        (const T*)x, (const T*)res, (T*)out, sums, gamma, beta, run_mean, run_var, save_mean,
                        save_invstd, M, C, mp.cvecs, mp.rows_per_block, eps, momentum, relu, training, num_batches);)
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_bn_fwd_apply(const void* x, const void* res, void* out, const float* sums, const float* gamma,
                             const float* beta, float* run_mean, float* run_var, float* save_mean, float* save_invstd,
                             long long M, int C, float eps, float momentum, int relu, int training, int dtype,
                             void* stream) {
  return mpb_bn_fwd_apply_ex(x, res, out, sums, gamma, beta, run_mean, run_var, save_mean, save_invstd, nullptr, M, C,
                             eps, momentum, relu, training, dtype, stream);
}

// dsum: 2*C floats scratch, zeroed here by the reduce call.
// flags bit 0: dgamma/dbeta are added to what the buffers hold (they point into the gradient arena);
// flags bit 1: `dsum` is already zero (a slice of the engine's per-step pre-zeroed pool) -- skip the memset.
MPB_API int mpb_bn_bwd_ex(const void* dz, const void* z, const void* x, const float* gamma, const float* save_mean,
                          const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta,
                          long long M, int C, int relu, int flags, int dtype, void* stream) {
  if (C % 8 != 0 || C > 2048) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int accumulate = flags & 1;
  if (!(flags & 2)) MPB_CUDA(cudaMemsetAsync(dsum, 0, 2 * C * sizeof(float), st));
  if (mpbbn::stream_ok(M, C, dtype))
    return mpbbn::stream_bwd(dz, z, x, gamma, nullptr, save_mean, save_invstd, dsum, dx, dres, dgamma, dbeta, M, C,
                             relu ? 1 : 0, accumulate, st);
  const Map mp = make_map(C);
  const size_t smem = (size_t)mp.threads * 16 * sizeof(float);
  const int g1 = rows_grid(M, mp.rows_per_block, 4), g2 = rows_grid(M, mp.rows_per_block, 8);
  DISPATCH_T(dtype,
             bn_bwd_reduce_kernel<T><<<g1, mp.threads, smem, st>>>((const T*)dz, (const T*)z, (const T*)x, save_mean,
                                                                   save_invstd, dsum, M, C, mp.cvecs, mp.rows_per_block, relu);
             bn_bwd_apply_kernel<T><<<g2, mp.threads, 0, st>>>((const T*)dz, (const T*)z, (const T*)x, gamma, save_mean,
                                                               save_invstd, dsum, (T*)dx, (T*)dres, dgamma, dbeta, M, C,
                                                               mp.cvecs, mp.rows_per_block, relu, accumulate);)
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_bn_bwd(const void* dz, const void* z, const void* x, const float* gamma, const float* save_mean,
                       const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta,
                       long long M, int C, int relu, int dtype, void* stream) {
  return mpb_bn_bwd_ex(dz, z, x, gamma, save_mean, save_invstd, dsum, dx, dres, dgamma, dbeta, M, C, relu, 0, dtype,
                       stream);
}

// ---------------------------------------------------------------------------------------------------------
// BatchNorm+ReLU backward WITHOUT reading the forward output: for BN -> ReLU (no residual) the ReLU mask is
// recomputed from x with the same fp32 expression the forward used (fmaf(x, gamma*invstd, beta - mean*gamma*invstd)
// > 0), so each pass reads two tensors (dz, x) instead of three.
// ---------------------------------------------------------------------------------------------------------
namespace {
template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_reduce_rx_kernel(const T* __restrict__ dz, const T* __restrict__ x,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               const float* __restrict__ save_mean,
                                                               const float* __restrict__ save_invstd, float* __restrict__ dsum,
                                                               long long M, int C, int cvecs, int rows_per_block) {
  extern __shared__ float smem[];
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float mean[8], istd[8], sc[8], sh[8], s0[8] = {0}, s1[8] = {0};
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    mean[k] = save_mean[c]; istd[k] = save_invstd[c];
    sc[k] = (gamma ? gamma[c] : 1.f) * istd[k];
    sh[k] = (beta ? beta[c] : 0.f) - mean[k] * sc[k];
  }
  // reduction pass: few registers are live, so two rows per trip (4 independent 16 B loads) pay off here (measured),
  // unlike in the apply passes
  const long long stride = (long long)gridDim.x * rows_per_block;
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += 2 * stride) {
    float g[2][8], xv[2][8];
    const bool two = m + stride < M;
    load8(dz + m * C + cv * 8, g[0]);
    load8(x + m * C + cv * 8, xv[0]);
    if (two) { load8(dz + (m + stride) * C + cv * 8, g[1]); load8(x + (m + stride) * C + cv * 8, xv[1]); }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float gg = fmaf(xv[u][k], sc[k], sh[k]) > 0.f ? g[u][k] : 0.f;
        s0[k] += gg; s1[k] += gg * (xv[u][k] - mean[k]) * istd[k];
      }
    }
  }
  reduce_pair_to_global(s0, s1, dsum, dsum + C, cvecs, rows_per_block, smem);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_apply_rx_kernel(const T* __restrict__ dz, const T* __restrict__ x,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              const float* __restrict__ save_mean,
                                                              const float* __restrict__ save_invstd,
                                                              const float* __restrict__ dsum, T* __restrict__ dx,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta, long long M,
                                                              int C, int cvecs, int rows_per_block, int accumulate) {
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float mean[8], istd[8], sc[8], sh[8], b0[8], b1[8];
  const float invM = 1.f / (float)M;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    mean[k] = save_mean[c]; istd[k] = save_invstd[c];
    sc[k] = (gamma ? gamma[c] : 1.f) * istd[k];
    sh[k] = (beta ? beta[c] : 0.f) - mean[k] * sc[k];
    b0[k] = dsum[c] * invM; b1[k] = dsum[C + c] * invM;
    if (blockIdx.x == 0 && r == 0) {
      if (dgamma) dgamma[c] = dsum[C + c] + (accumulate ? dgamma[c] : 0.f);
      if (dbeta) dbeta[c] = dsum[c] + (accumulate ? dbeta[c] : 0.f);
    }
  }
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float g[8], xv[8];
    load8(dz + m * C + cv * 8, g);
    load8(x + m * C + cv * 8, xv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float gg = fmaf(xv[k], sc[k], sh[k]) > 0.f ? g[k] : 0.f;
      const float xhat = (xv[k] - mean[k]) * istd[k];
      xv[k] = sc[k] * (gg - b0[k] - xhat * b1[k]);
    }
    store8(dx + m * C + cv * 8, xv);
  }
}
}  // namespace

// BN -> ReLU backward (no residual branch); the mask is recomputed from x, the forward output is not needed.
MPB_API int mpb_bn_relu_bwd_recompute_ex(const void* dz, const void* x, const float* gamma, const float* beta,
                                         const float* save_mean, const float* save_invstd, float* dsum, void* dx,
                                         float* dgamma, float* dbeta, long long M, int C, int flags, int dtype,
                                         void* stream) {
  if (C % 8 != 0 || C > 2048) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  const int accumulate = flags & 1;
  if (!(flags & 2)) MPB_CUDA(cudaMemsetAsync(dsum, 0, 2 * C * sizeof(float), st));
  if (mpbbn::stream_ok(M, C, dtype))
    return mpbbn::stream_bwd(dz, nullptr, x, gamma, beta, save_mean, save_invstd, dsum, dx, nullptr, dgamma, dbeta, M, C, 2,
                             accumulate, st);
  const Map mp = make_map(C);
  const size_t smem = (size_t)mp.threads * 16 * sizeof(float);
  const int g1 = rows_grid(M, mp.rows_per_block, 8), g2 = rows_grid(M, mp.rows_per_block, 8);
  DISPATCH_T(dtype,
             bn_bwd_reduce_rx_kernel<T><<<g1, mp.threads, smem, st>>>((const T*)dz, (const T*)x, gamma, beta, save_mean,
                                                                      save_invstd, dsum, M, C, mp.cvecs, mp.rows_per_block);
             bn_bwd_apply_rx_kernel<T><<<g2, mp.threads, 0, st>>>((const T*)dz, (const T*)x, gamma, beta, save_mean,
                                                                  save_invstd, dsum, (T*)dx, dgamma, dbeta, M, C, mp.cvecs,
                                                                  mp.rows_per_block, accumulate);)
  MPB_CHECK_LAUNCH();
  return 0;
}

// ReLU backward fused with the per-channel column sum of its result: dx = dy * (y > 0), sums[c] += sum_m dx[m][c].
// For a Conv2d(+bias) -> ReLU pair (VGG / AlexNet / SqueezeNet, models.py:47-72) dx is the convolution's output
// gradient and its column sums are the bias gradient, so one pass replaces relu_bwd + a separate reduction over dx.
// sums: 2*C floats, zeroed by the caller (the second half is scratch of the shared reduction helper).
namespace {
template <typename T>
__global__ void __launch_bounds__(256) relu_bwd_colsum_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                              T* __restrict__ dx, float* __restrict__ sums, long long M,
                                                              int C, int cvecs, int rows_per_block) {
  extern __shared__ float smem[];
  const int cv = threadIdx.x % cvecs, r = threadIdx.x / cvecs;
  float s[8] = {0}, z[8] = {0};
  for (long long m = (long long)blockIdx.x * rows_per_block + r; m < M; m += (long long)gridDim.x * rows_per_block) {
    float g[8], v[8];
    load8(dy + m * C + cv * 8, g);
    load8(y + m * C + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) { g[k] = v[k] > 0.f ? g[k] : 0.f; s[k] += g[k]; }
    store8(dx + m * C + cv * 8, g);
  }
  reduce_pair_to_global(s, z, sums, sums + C, cvecs, rows_per_block, smem);
}
}  // namespace

MPB_API int mpb_relu_bwd_colsum(const void* dy, const void* y, void* dx, float* sums, long long M, int C, int dtype,
                                void* stream) {
  if (C <= 0 || C % 8 != 0 || C > 2048) return -1;
  if (M <= 0) return 0;
  const Map mp = make_map(C);
  const size_t smem = (size_t)mp.threads * 16 * sizeof(float);
  const int g = rows_grid(M, mp.rows_per_block, 8);
  cudaStream_t st = (cudaStream_t)stream;
  DISPATCH_T(dtype, relu_bwd_colsum_kernel<T><<<g, mp.threads, smem, st>>>((const T*)dy, (const T*)y, (T*)dx, sums, M, C,
                                                                             mp.cvecs, mp.rows_per_block);)
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_bn_relu_bwd_recompute(const void* dz, const void* x, const float* gamma, const float* beta,
                                      const float* save_mean, const float* save_invstd, float* dsum, void* dx,
                                      float* dgamma, float* dbeta, long long M, int C, int dtype, void* stream) {
  return mpb_bn_relu_bwd_recompute_ex(dz, x, gamma, beta, save_mean, save_invstd, dsum, dx, dgamma, dbeta, M, C, 0,
                                      dtype, stream);
}

// ---------------------------------------------------------------------------------------------------------
// BatchNorm backward for callers that cannot allocate device memory.
//
// mpb_bn_bwd() needs 2*C floats of scratch (`dsum`) that the caller must allocate.  Some callers cannot do
// that: a host application that has handed the whole GPU to its own memory pool (so cudaMalloc fails or is
// forbidden by policy), or a latency-critical loop in which the implicit device synchronisation of
// cudaMalloc / cudaFree is not acceptable.  For them this variant keeps the scratch inside the library, in a
// fixed file-scope device array.  Consequences the caller has to know about:
//   * capacity: at most MPB_NOALLOC_MAX_CHANNELS (2048) channels -- 16 KB of device memory that every GPU which
//     loads this library holds until the process exits.  Larger C is rejected with -4 and a message on stderr.
//   * set-up: the first call on a given device resolves the array's device address (cudaGetSymbolAddress, with
//     that device current) and caches it per device.  Make that first call before any stream capture begins.
//   * sharing: every call on a device uses the same array.  It is cleared at the start of each call (the sums
//     must start at zero), and it still holds the previous call's sums afterwards.  Two calls whose kernels can
//     overlap on the GPU (different streams, no dependency) would race on it -- serialise them.
// No cudaMalloc / cudaFree / allocating library routine is called here.
// ---------------------------------------------------------------------------------------------------------
#define MPB_NOALLOC_MAX_CHANNELS 2048
#define MPB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) float g_bn_bwd_scratch[2 * MPB_NOALLOC_MAX_CHANNELS];

// In host code `g_bn_bwd_scratch` is only a handle for the symbol, not a device address; kernels must be given
// the address obtained below (one per device, because each device has its own copy of the array).
static float* g_bn_bwd_scratch_addr[MPB_NOALLOC_MAX_DEVICES] = {nullptr};

MPB_API int mpb_bn_bwd_noalloc(const void* dz, const void* z, const void* x, const float* gamma,
                               const float* save_mean, const float* save_invstd, void* dx, void* dres, float* dgamma,
                               float* dbeta, long long M, int C, int relu, int dtype, void* stream) {
  if (C > MPB_NOALLOC_MAX_CHANNELS) {
    fprintf(stderr, "mpb_bn_bwd_noalloc: C=%d exceeds the built-in scratch capacity of %d channels; "
                    "use mpb_bn_bwd with caller-provided scratch instead\n", C, MPB_NOALLOC_MAX_CHANNELS);
    return -4;
  }
  if (C <= 0 || C % 8 != 0) return -1;
  if (M <= 0) {                                         // an empty batch would launch zero-sized grids
    fprintf(stderr, "mpb_bn_bwd_noalloc: M=%lld rows; need at least one\n", M);
    return -1;
  }
  int dev = -1;
  MPB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MPB_NOALLOC_MAX_DEVICES) return -5;
  if (g_bn_bwd_scratch_addr[dev] == nullptr) {           // one-time set-up for this device (device `dev` is current)
    float* ptr = nullptr;
    MPB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_bn_bwd_scratch));
    g_bn_bwd_scratch_addr[dev] = ptr;
  }
  float* dsum = g_bn_bwd_scratch_addr[dev];
  cudaStream_t st = (cudaStream_t)stream;
  MPB_CUDA(cudaMemsetAsync(dsum, 0, 2 * C * sizeof(float), st));     // the sums accumulate with atomics: start from zero
  const Map mp = make_map(C);
  const size_t smem = (size_t)mp.threads * 16 * sizeof(float);
  const int g1 = rows_grid(M, mp.rows_per_block, 4), g2 = rows_grid(M, mp.rows_per_block, 8);
  DISPATCH_T(dtype,
             bn_bwd_reduce_kernel<T><<<g1, mp.threads, smem, st>>>((const T*)dz, (const T*)z, (const T*)x, save_mean,
                                                                   save_invstd, dsum, M, C, mp.cvecs, mp.rows_per_block, relu);
             bn_bwd_apply_kernel<T><<<g2, mp.threads, 0, st>>>((const T*)dz, (const T*)z, (const T*)x, gamma, save_mean,
                                                               save_invstd, dsum, (T*)dx, (T*)dres, dgamma, dbeta, M, C,
                                                               mp.cvecs, mp.rows_per_block, relu, 0);)
  MPB_CHECK_LAUNCH();
  return 0;
}
