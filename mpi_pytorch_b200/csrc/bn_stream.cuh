// Bulk-copy streaming pipeline for the per-channel elementwise passes of BatchNorm (bf16, NHWC).
//
// The register-resident versions in batchnorm.cu keep one or two 16-byte loads per thread in flight; with 60-100
// registers per thread that is 16-36 KB per SM, and the passes sit at 50-75 % of the HBM copy rate.  An NHWC
// activation is one flat array, so here a single thread per CTA streams it through shared memory with 1-D
// cp.async.bulk copies (TMA without a tensor map) into a 4-stage mbarrier ring -- ~100 KB in flight per SM no matter
// how many registers the math needs -- and all threads consume 16-byte vectors from shared memory.
#pragma once
#include "common.cuh"

namespace mpbbn {

constexpr int kStages = 4;
constexpr int kChunkBytes = 8192;         // target bytes per input per stage
constexpr long long kMinBytes = 4ll << 20;  // smaller tensors stay on the register-resident kernels

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count));
}
__device__ __forceinline__ void bar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred P1;\n\t"
      "BNS_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra BNS_DONE;\n\t"
      "bra BNS_WAIT;\n\t"
      "BNS_DONE:\n\t"
      "}" ::"r"(s_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(s_u32(dst_smem)), "l"(src), "r"(bytes), "r"(s_u32(bar)) : "memory");
}

// Geometry shared by host and device: how a [M][C] bf16 array is cut into chunks of whole rows.
struct StreamGeom {
  int C, cvecs, rows_per_block, threads;  // thread t -> channel group t % cvecs, first row t / cvecs
  int chunk_rows, chunk_elems;            // rows (a multiple of rows_per_block) and elements per chunk
  long long M, total;                     // rows and elements of the tensor
  int nchunks;
};

inline StreamGeom make_geom(long long M, int C) {
  StreamGeom g;
  g.C = C; g.cvecs = C / 8;
  g.rows_per_block = 256 / g.cvecs;
  if (g.rows_per_block < 1) g.rows_per_block = 1;
  g.threads = g.cvecs * g.rows_per_block;
  int rows = kChunkBytes / (C * 2) / g.rows_per_block * g.rows_per_block;
  if (rows < g.rows_per_block) rows = g.rows_per_block;
  g.chunk_rows = rows; g.chunk_elems = rows * C;
  g.M = M; g.total = M * C;
  g.nchunks = (int)((M + rows - 1) / rows);
  return g;
}

// Ring of kStages stages, each holding `nin` input chunks back to back.  One instance per CTA (in shared memory).
struct Ring {
  uint64_t full[kStages];
};

template <int NIN>
struct Inputs { const __nv_bfloat16* p[NIN]; };

// Issue the loads of this CTA's i-th chunk into stage i % kStages (called by one thread).
template <int NIN>
__device__ __forceinline__ void issue_chunk(const StreamGeom& g, const Inputs<NIN>& in, int nin, Ring* ring, uint8_t* smem,
                                            int i) {
  const int s = i % kStages;
  const long long j = (long long)blockIdx.x + (long long)i * gridDim.x;
  const long long off = j * g.chunk_elems;
  long long n = g.total - off;
  if (n > g.chunk_elems) n = g.chunk_elems;
  const uint32_t bytes = (uint32_t)(n * 2);
  bar_expect_tx(&ring->full[s], bytes * (uint32_t)nin);
  uint8_t* dst = smem + (size_t)s * NIN * (g.chunk_elems * 2);
#pragma unroll
  for (int k = 0; k < NIN; ++k)
    if (k < nin) bulk_load(dst + (size_t)k * (g.chunk_elems * 2), in.p[k] + off, bytes, &ring->full[s]);
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&v)[8]) {
  const unsigned w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u); }
}

// host entry points (bn_stream.cu); mask_mode: 0 none, 1 from z, 2 recomputed from x
bool stream_ok(long long M, int C, int dtype);
int stream_fwd(const void* x, const void* res, void* out, const float* sums, const float* gamma, const float* beta,
               float* run_mean, float* run_var, float* save_mean, float* save_invstd, long long* num_batches, long long M,
               int C, float eps, float momentum, int relu, int training, cudaStream_t st);
int stream_stats(const void* x, float* sums, long long M, int C, cudaStream_t st);
int stream_bwd(const void* dz, const void* z, const void* x, const float* gamma, const float* beta, const float* save_mean,
               const float* save_invstd, float* dsum, void* dx, void* dres, float* dgamma, float* dbeta, long long M, int C,
               int mask_mode, int accumulate, cudaStream_t st);

}  // namespace mpbbn
