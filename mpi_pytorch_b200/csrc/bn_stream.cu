// K7, streaming form: BatchNorm apply / backward-reduce / backward-apply passes fed by 1-D bulk copies (bn_stream.cuh).
// Same maths, thread mapping and outputs as the register-resident kernels in batchnorm.cu (which remain the fp32 and
// small-tensor path); mpb_bn_*_ex dispatch here for bf16 tensors of at least kMinBytes.  Reference counterpart: the
// ATen CPU batch_norm / relu / add calls behind every BatchNorm2d built in models.py:33..87.
#include "bn_stream.cuh"

using namespace mpb;
using namespace mpbbn;

namespace {

typedef __nv_bfloat16 bf16;

#define BNS_PROLOGUE(NIN_)                                                                                     \
  extern __shared__ __align__(128) uint8_t smem[];                                                            \
  __shared__ Ring ring;                                                                                        \
  const int t = threadIdx.x, cv = t % g.cvecs, r0 = t / g.cvecs;                                               \
  const int mine = ((int)blockIdx.x < g.nchunks) ? (g.nchunks - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0; \
  const int chunk_bytes = g.chunk_elems * 2;                                                                   \
  if (t == 0) {                                                                                                \
    for (int s = 0; s < kStages; ++s) bar_init(&ring.full[s], 1);                                              \
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");                                         \
  }                                                                                                            \
  __syncthreads();                                                                                             \
  if (t == 0)                                                                                                  \
    for (int i = 0; i < kStages && i < mine; ++i) issue_chunk<NIN_>(g, in, nin, &ring, smem, i);

// rows of chunk i that exist, and the stage's base address
#define BNS_CHUNK(NIN_)                                                                         \
  const int s = i % kStages;                                                                    \
  bar_wait(&ring.full[s], (uint32_t)((i / kStages) & 1));                                       \
  const long long row0 = ((long long)blockIdx.x + (long long)i * gridDim.x) * g.chunk_rows;     \
  const int rows_here = (int)((g.M - row0 < g.chunk_rows) ? (g.M - row0) : g.chunk_rows);       \
  const uint8_t* base = smem + (size_t)s * NIN_ * chunk_bytes;

#define BNS_RELEASE(NIN_)                                                                       \
  __syncthreads();                                                                              \
  if (t == 0 && i + kStages < mine) issue_chunk<NIN_>(g, in, nin, &ring, smem, i + kStages);

__global__ void __launch_bounds__(256) bns_fwd_kernel(Inputs<2> in, int nin, bf16* __restrict__ out,
                                                      const float* __restrict__ sums, const float* __restrict__ gamma,
                                                      const float* __restrict__ beta, float* __restrict__ run_mean,
                                                      float* __restrict__ run_var, float* __restrict__ save_mean,
                                                      float* __restrict__ save_invstd, long long* __restrict__ num_batches,
                                                      const StreamGeom g, float eps, float momentum, int relu, int training) {
  BNS_PROLOGUE(2)
  if (num_batches && training && blockIdx.x == 0 && t == 0) *num_batches += 1;
  float sc[8], sh[8];
  const float invM = 1.f / (float)g.M;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    float mean, var;
    if (training) {
      mean = sums[c] * invM;
      var = fmaxf(sums[g.C + c] * invM - mean * mean, 0.f);
    } else {
      mean = run_mean[c];
      var = run_var[c];
    }
    const float invstd = rsqrtf(var + eps);
    const float ga = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
    sc[k] = ga * invstd;
    sh[k] = be - mean * sc[k];
    if (training && blockIdx.x == 0 && r0 == 0) {
      save_mean[c] = mean;
      save_invstd[c] = invstd;
      if (run_mean) {
        const float unbiased = g.M > 1 ? var * ((float)g.M / (float)(g.M - 1)) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
      }
    }
  }
  for (int i = 0; i < mine; ++i) {
    BNS_CHUNK(2)
    for (int r = r0; r < rows_here; r += g.rows_per_block) {
      const size_t e = (size_t)r * g.C + cv * 8;
      float v[8];
      unpack8(*reinterpret_cast<const uint4*>(base + e * 2), v);
      if (nin > 1) {
        float rr[8];
        unpack8(*reinterpret_cast<const uint4*>(base + chunk_bytes + e * 2), rr);
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
      store8(out + (size_t)row0 * g.C + e, v);
    }
    BNS_RELEASE(2)
  }
}

// Per-channel sum and sum of squares (BatchNorm statistics of a tensor that did not come out of one of our conv
// epilogues: DenseNet's concatenated features, a conv-bias gradient).  sums must be zero on entry.
__global__ void __launch_bounds__(256) bns_stats_kernel(Inputs<1> in, int nin, float* __restrict__ sums, const StreamGeom g) {
  BNS_PROLOGUE(1)
  float s0[8] = {0}, s1[8] = {0};
  for (int i = 0; i < mine; ++i) {
    BNS_CHUNK(1)
    for (int r = r0; r < rows_here; r += g.rows_per_block) {
      float v[8];
      unpack8(*reinterpret_cast<const uint4*>(base + ((size_t)r * g.C + cv * 8) * 2), v);
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += v[k]; s1[k] = fmaf(v[k], v[k], s1[k]); }
    }
    BNS_RELEASE(1)
  }
  float* red = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int k = 0; k < 8; ++k) { red[t * 16 + k] = s0[k]; red[t * 16 + 8 + k] = s1[k]; }
  __syncthreads();
  for (int u = t; u < g.cvecs * 16; u += blockDim.x) {
    const int c8 = u >> 4, k = u & 15;
    float s = 0.f;
    for (int r = 0; r < g.rows_per_block; ++r) s += red[(r * g.cvecs + c8) * 16 + k];
    if (k < 8) atomicAdd(sums + c8 * 8 + k, s);
    else atomicAdd(sums + g.C + c8 * 8 + (k - 8), s);
  }
}

// mask_mode: 0 = no ReLU, 1 = ReLU mask from the saved output z (input 2), 2 = mask recomputed from x
__global__ void __launch_bounds__(256) bns_reduce_kernel(Inputs<3> in, int nin, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, const float* __restrict__ save_mean,
                                                         const float* __restrict__ save_invstd, float* __restrict__ dsum,
                                                         const StreamGeom g, int mask_mode) {
  BNS_PROLOGUE(3)
  float mean[8], istd[8], sc[8], sh[8], s0[8] = {0}, s1[8] = {0};
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    mean[k] = save_mean[c]; istd[k] = save_invstd[c];
    sc[k] = (gamma ? gamma[c] : 1.f) * istd[k];
    sh[k] = (beta ? beta[c] : 0.f) - mean[k] * sc[k];
  }
  for (int i = 0; i < mine; ++i) {
    BNS_CHUNK(3)
    for (int r = r0; r < rows_here; r += g.rows_per_block) {
      const size_t e = (size_t)r * g.C + cv * 8;
      float gg[8], xv[8];
      unpack8(*reinterpret_cast<const uint4*>(base + e * 2), gg);
      unpack8(*reinterpret_cast<const uint4*>(base + chunk_bytes + e * 2), xv);
      if (mask_mode == 1) {
        float zv[8];
        unpack8(*reinterpret_cast<const uint4*>(base + 2 * chunk_bytes + e * 2), zv);
#pragma unroll
        for (int k = 0; k < 8; ++k) gg[k] = zv[k] > 0.f ? gg[k] : 0.f;
      } else if (mask_mode == 2) {
#pragma unroll
        for (int k = 0; k < 8; ++k) gg[k] = fmaf(xv[k], sc[k], sh[k]) > 0.f ? gg[k] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0[k] += gg[k]; s1[k] += gg[k] * (xv[k] - mean[k]) * istd[k]; }
    }
    BNS_RELEASE(3)
  }
  // block reduction over the rows_per_block threads that share a channel group (the ring is idle now: reuse it)
  float* red = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int k = 0; k < 8; ++k) { red[t * 16 + k] = s0[k]; red[t * 16 + 8 + k] = s1[k]; }
  __syncthreads();
  for (int u = t; u < g.cvecs * 16; u += blockDim.x) {
    const int c8 = u >> 4, k = u & 15;
    float s = 0.f;
    for (int r = 0; r < g.rows_per_block; ++r) s += red[(r * g.cvecs + c8) * 16 + k];
    if (k < 8) atomicAdd(dsum + c8 * 8 + k, s);
    else atomicAdd(dsum + g.C + c8 * 8 + (k - 8), s);
  }
}

__global__ void __launch_bounds__(256) bns_apply_kernel(Inputs<3> in, int nin, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, const float* __restrict__ save_mean,
                                                        const float* __restrict__ save_invstd, const float* __restrict__ dsum,
                                                        bf16* __restrict__ dx, bf16* __restrict__ dres,
                                                        float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                        const StreamGeom g, int mask_mode, int accumulate) {
  BNS_PROLOGUE(3)
  float mean[8], istd[8], sc[8], sh[8], b0[8], b1[8];
  const float invM = 1.f / (float)g.M;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cv * 8 + k;
    mean[k] = save_mean[c]; istd[k] = save_invstd[c];
    sc[k] = (gamma ? gamma[c] : 1.f) * istd[k];
    sh[k] = (beta ? beta[c] : 0.f) - mean[k] * sc[k];
    b0[k] = dsum[c] * invM; b1[k] = dsum[g.C + c] * invM;
    if (blockIdx.x == 0 && r0 == 0) {
      if (dgamma) dgamma[c] = dsum[g.C + c] + (accumulate ? dgamma[c] : 0.f);
      if (dbeta) dbeta[c] = dsum[c] + (accumulate ? dbeta[c] : 0.f);
    }
  }
  for (int i = 0; i < mine; ++i) {
    BNS_CHUNK(3)
    for (int r = r0; r < rows_here; r += g.rows_per_block) {
      const size_t e = (size_t)r * g.C + cv * 8;
      float gg[8], xv[8];
      unpack8(*reinterpret_cast<const uint4*>(base + e * 2), gg);
      unpack8(*reinterpret_cast<const uint4*>(base + chunk_bytes + e * 2), xv);
      if (mask_mode == 1) {
        float zv[8];
        unpack8(*reinterpret_cast<const uint4*>(base + 2 * chunk_bytes + e * 2), zv);
#pragma unroll
        for (int k = 0; k < 8; ++k) gg[k] = zv[k] > 0.f ? gg[k] : 0.f;
      } else if (mask_mode == 2) {
#pragma unroll
        for (int k = 0; k < 8; ++k) gg[k] = fmaf(xv[k], sc[k], sh[k]) > 0.f ? gg[k] : 0.f;
      }
      if (dres) store8(dres + (size_t)row0 * g.C + e, gg);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float xhat = (xv[k] - mean[k]) * istd[k];
        xv[k] = sc[k] * (gg[k] - b0[k] - xhat * b1[k]);
      }
      store8(dx + (size_t)row0 * g.C + e, xv);
    }
    BNS_RELEASE(3)
  }
}

template <typename K>
int set_smem(K kernel, int bytes) {
  return (int)cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

int grid_for_chunks(int nchunks, int ctas_per_sm) {
  const int cap = kSMs * ctas_per_sm;
  return nchunks < cap ? (nchunks < 1 ? 1 : nchunks) : cap;
}

}  // namespace

namespace mpbbn {

bool stream_ok(long long M, int C, int dtype) {
  static const int enabled = [] { const char* v = getenv("MPB_BN_STREAM"); return (v && *v) ? atoi(v) : 1; }();
  return enabled && dtype == MPB_BF16 && C % 8 == 0 && C <= 2048 && M * C * 2 >= kMinBytes;
}

int stream_fwd(const void* x, const void* res, void* out, const float* sums, const float* gamma, const float* beta,
               float* run_mean, float* run_var, float* save_mean, float* save_invstd, long long* num_batches, long long M,
               int C, float eps, float momentum, int relu, int training, cudaStream_t st) {
  const StreamGeom g = make_geom(M, C);
  Inputs<2> in;
  in.p[0] = (const bf16*)x; in.p[1] = (const bf16*)res;
  const int nin = res ? 2 : 1;
  const int smem = kStages * 2 * g.chunk_elems * 2;
  static bool configured = false;
  if (!configured) { int rc = set_smem(bns_fwd_kernel, 200 * 1024); if (rc) return rc; configured = true; }
  bns_fwd_kernel<<<grid_for_chunks(g.nchunks, 3), g.threads, smem, st>>>(in, nin, (bf16*)out, sums, gamma, beta, run_mean,
                                                                         run_var, save_mean, save_invstd, num_batches, g,
                                                                         eps, momentum, relu, training);
  return (int)cudaGetLastError();
}

int stream_stats(const void* x, float* sums, long long M, int C, cudaStream_t st) {
  const StreamGeom g = make_geom(M, C);
  Inputs<1> in;
  in.p[0] = (const bf16*)x;
  int smem = kStages * g.chunk_elems * 2;
  if (smem < g.threads * 16 * 4) smem = g.threads * 16 * 4;
  static bool configured = false;
  if (!configured) { int rc = set_smem(bns_stats_kernel, 200 * 1024); if (rc) return rc; configured = true; }
  bns_stats_kernel<<<grid_for_chunks(g.nchunks, 4), g.threads, smem, st>>>(in, 1, sums, g);
  return (int)cudaGetLastError();
}

int stream_bwd(const void* dz, const void* z, const void* x, const float* gamma, const float* beta, const float* save_mean,
               const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta, long long M, int C,
               int mask_mode, int accumulate, cudaStream_t st) {
  const StreamGeom g = make_geom(M, C);
  Inputs<3> in;
  in.p[0] = (const bf16*)dz; in.p[1] = (const bf16*)x; in.p[2] = (const bf16*)z;
  const int nin = (mask_mode == 1) ? 3 : 2;
  int smem = kStages * 3 * g.chunk_elems * 2;
  if (smem < g.threads * 16 * 4) smem = g.threads * 16 * 4;
  static bool configured = false;
  if (!configured) {
    int rc = set_smem(bns_reduce_kernel, 200 * 1024);
    if (rc) return rc;
    rc = set_smem(bns_apply_kernel, 200 * 1024);
    if (rc) return rc;
    configured = true;
  }
  const int grid = grid_for_chunks(g.nchunks, 2);
  bns_reduce_kernel<<<grid, g.threads, smem, st>>>(in, nin, gamma, beta, save_mean, save_invstd, dsum, g, mask_mode);
  bns_apply_kernel<<<grid, g.threads, smem, st>>>(in, nin, gamma, beta, save_mean, save_invstd, dsum, (bf16*)dx, (bf16*)dres,
                                                  dgamma, dbeta, g, mask_mode, accumulate);
  return (int)cudaGetLastError();
}

}  // namespace mpbbn
