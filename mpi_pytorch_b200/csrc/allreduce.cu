// K1 / K2 / K13: gradient all-reduce fused with averaging + optimizer update, broadcast and scalar
// reduction, written as sm_100a kernels that read and write *peer GPUs' memory directly* over
// NVLink 5 / NVSwitch.  Replaces mpi_tools.py:30-37 (per-tensor blocking MPI.Allreduce + host divide)
// + main.py:155 (optimizer.step) + mpi_tools.py:47-53 (per-tensor Bcast) of the reference.
//
// Memory model.  Every rank owns symmetric buffers (cudaMalloc + cudaIpc handles exchanged by the
// Python bootstrap, opened by every peer), so `bufs[p] + off` is rank p's copy of the same logical
// element.  Cross-GPU ordering uses a per-block flag barrier in a symmetric signal pad:
//   signal:  st.release.sys  peer_pad[block][my_rank] = epoch
//   wait  :  ld.acquire.sys  my_pad[block][peer]     >= epoch
// Epochs increase monotonically (device-resident counter per block, so CUDA-graph replays stay
// correct) -> no flag reset, no ABA.  All ranks must launch the same sequence of collectives with the
// same grid (they do: the bucket schedule is deterministic).  Grids are capped well below 148 CTAs so
// every CTA of every rank is co-resident and the rest of the GPU keeps running backward kernels.
//
// Algorithms (chosen per bucket by the host from the measured crossover):
//   one-shot : barrier; every rank loads all W copies of every element, sums in rank order 0..W-1
//              (bit-identical on all ranks), keeps the result in registers across the end barrier,
//              then applies the epilogue locally.  One element vector per thread -> n <= grid*512*4.
//   two-shot : barrier; rank r reduces only slice r (reduce-scatter by peer loads), applies the
//              epilogue, and stores the result into slice r of *every* peer (all-gather by peer
//              stores); barrier.  With the optimizer epilogue the owner keeps Adam moments for its
//              slice only (ZeRO-1 data flow) and all-gathers updated fp32 params + the bf16 shadow.
#include "common.cuh"

using namespace mpb;

#define MPB_MAX_WORLD 8
#define MPB_MAX_COMM_BLOCKS 128
#define MPB_COMM_THREADS 512

struct PeerBufs { void* p[MPB_MAX_WORLD]; };

struct OptArgs {
  int kind;           // 0 = none (write averaged grads), 1 = adam, 2 = sgd-momentum
  float lr, b1, b2, eps, wd, bc1, bc2_sqrt, momentum;
  int first_step;
  const int* step_ptr;  // optional device-resident step counter (CUDA-graph replays: bias corrections follow it)
};

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer data loads: system-scope relaxed, no L1 allocation (peer lines must never be served stale from L1).
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// NVLS: one load returns the sum over every GPU bound to the multicast object (reduced inside the NVSwitch), one
// store is replicated to all of them.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x4(void* mc, const uint2& v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(mc), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void st_peer_u2(void* p, const uint2& v) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// Per-block barrier across ranks.  pads[p] = rank p's signal pad: uint32 [MPB_MAX_COMM_BLOCKS][MPB_MAX_WORLD].
__device__ __forceinline__ void rank_barrier(const PeerBufs& pads, int rank, int W, uint32_t epoch) {
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < W) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(pads.p[threadIdx.x]) + blockIdx.x * MPB_MAX_WORLD + rank;
    st_release_sys(dst, epoch);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(pads.p[rank]) + blockIdx.x * MPB_MAX_WORLD + threadIdx.x;
    while ((int32_t)(ld_acquire_sys(src) - epoch) < 0) { __nanosleep(20); }
  }
  __syncthreads();
}

__device__ __forceinline__ float opt_one(float p, float g, float& m, float& v, const OptArgs& a) {
  if (a.wd != 0.f) g += a.wd * p;
  if (a.kind == 1) {
    m = a.b1 * m + (1.f - a.b1) * g;
    v = a.b2 * v + (1.f - a.b2) * g * g;
    return p - (a.lr / a.bc1) * (m / (sqrtf(v) / a.bc2_sqrt + a.eps));
  }
  if (a.momentum != 0.f) { m = a.first_step ? g : a.momentum * m + g; g = m; }
  return p - a.lr * g;
}

struct ArParams {
  PeerBufs grads;     // float* per rank, already offset to the bucket start
  PeerBufs params;    // float* per rank (two-shot optimizer epilogue writes all of them)
  PeerBufs shadows;   // bf16* per rank or null
  PeerBufs pads;
  uint32_t* epochs;   // local, [MPB_MAX_COMM_BLOCKS]
  float* m; float* v; // local optimizer state (bucket-offset)
  long long n;        // elements in the bucket (multiple of 4*W)
  int rank, W, algo;  // algo 0 = one-shot, 1 = two-shot, 2 = two-shot through NVLS multicast addresses
  float scale;        // 1/W
  float* mc_grads; float* mc_params; __nv_bfloat16* mc_shadow;   // multicast views of the same ranges (algo 2)
  OptArgs opt;
};

template <int W>
__device__ __forceinline__ float4 sum_peers(const ArParams& P, long long i4) {
  float4 v[W];
#pragma unroll
  for (int p = 0; p < W; ++p)                         // all W loads are in flight together
    v[p] = ld_peer_f4(reinterpret_cast<const float*>(P.grads.p[p]) + i4 * 4);
  float4 s = v[0];                                    // summation order fixed 0..W-1: same bits on every rank
#pragma unroll
  for (int p = 1; p < W; ++p) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
  s.x *= P.scale; s.y *= P.scale; s.z *= P.scale; s.w *= P.scale;
  return s;
}

template <int W>
__global__ void __launch_bounds__(MPB_COMM_THREADS, 1) allreduce_kernel(ArParams P) {
  if (P.opt.kind != 0 && P.opt.step_ptr != nullptr) {
    const int step = *P.opt.step_ptr;
    P.opt.first_step = (step == 1);
    P.opt.bc1 = 1.f - powf(P.opt.b1, (float)step);
    P.opt.bc2_sqrt = sqrtf(1.f - powf(P.opt.b2, (float)step));
  }
  const uint32_t e0 = P.epochs[blockIdx.x];
  rank_barrier(P.pads, P.rank, W, e0 + 1);            // every peer's gradients are complete and visible
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * blockDim.x;
  if (P.algo == 0) {
    // ---- one-shot: one float4 per thread, result lives in registers across the end barrier
    const long long n4 = P.n >> 2;
    const bool active = tid < n4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) s = sum_peers<W>(P, tid);
    rank_barrier(P.pads, P.rank, W, e0 + 2);          // everybody finished reading my buffer
    if (active) {
      float* g = reinterpret_cast<float*>(P.grads.p[P.rank]) + tid * 4;
      if (P.opt.kind == 0) {
        *reinterpret_cast<float4*>(g) = s;
      } else {
        float* pp = reinterpret_cast<float*>(P.params.p[P.rank]) + tid * 4;
        float4 p4 = *reinterpret_cast<float4*>(pp);
        float4 m4 = *reinterpret_cast<float4*>(P.m + tid * 4);
        float4 v4 = P.opt.kind == 1 ? *reinterpret_cast<float4*>(P.v + tid * 4) : make_float4(0, 0, 0, 0);
        p4.x = opt_one(p4.x, s.x, m4.x, v4.x, P.opt); p4.y = opt_one(p4.y, s.y, m4.y, v4.y, P.opt);
        p4.z = opt_one(p4.z, s.z, m4.z, v4.z, P.opt); p4.w = opt_one(p4.w, s.w, m4.w, v4.w, P.opt);
        *reinterpret_cast<float4*>(pp) = p4;
        *reinterpret_cast<float4*>(P.m + tid * 4) = m4;
        if (P.opt.kind == 1) *reinterpret_cast<float4*>(P.v + tid * 4) = v4;
        if (P.shadows.p[P.rank]) {
          uint2 u;
          *reinterpret_cast<__nv_bfloat162*>(&u.x) = __floats2bfloat162_rn(p4.x, p4.y);
          *reinterpret_cast<__nv_bfloat162*>(&u.y) = __floats2bfloat162_rn(p4.z, p4.w);
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(P.shadows.p[P.rank]) + tid * 4) = u;
        }
      }
    }
  } else if (P.algo == 2) {
    // ---- NVLS two-shot: the switch does the reduce-scatter (ld_reduce) and the all-gather (multicast store);
    // each GPU moves only 1/W of the bucket over its links in each direction.
    constexpr int U = 8;
    const long long slice4 = (P.n / W) >> 2;
    const long long base4 = slice4 * P.rank;
    for (long long j0 = 0; j0 < slice4; j0 += nthreads * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = j0 + (long long)u * nthreads + tid;
        if (j < slice4) v[u] = multimem_ld_reduce_f4(P.mc_grads + (base4 + j) * 4);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = j0 + (long long)u * nthreads + tid;
        if (j >= slice4) continue;
        const long long i4 = base4 + j;
        float4 s = v[u];
        s.x *= P.scale; s.y *= P.scale; s.z *= P.scale; s.w *= P.scale;
        if (P.opt.kind == 0) {
          multimem_st_f4(P.mc_grads + i4 * 4, s);
        } else {
          float* pp = reinterpret_cast<float*>(P.params.p[P.rank]) + i4 * 4;
          float4 p4 = *reinterpret_cast<float4*>(pp);
          float4 m4 = *reinterpret_cast<float4*>(P.m + i4 * 4);
          float4 v4 = P.opt.kind == 1 ? *reinterpret_cast<float4*>(P.v + i4 * 4) : make_float4(0, 0, 0, 0);
          p4.x = opt_one(p4.x, s.x, m4.x, v4.x, P.opt); p4.y = opt_one(p4.y, s.y, m4.y, v4.y, P.opt);
          p4.z = opt_one(p4.z, s.z, m4.z, v4.z, P.opt); p4.w = opt_one(p4.w, s.w, m4.w, v4.w, P.opt);
          *reinterpret_cast<float4*>(P.m + i4 * 4) = m4;
          if (P.opt.kind == 1) *reinterpret_cast<float4*>(P.v + i4 * 4) = v4;
          multimem_st_f4(P.mc_params + i4 * 4, p4);
          if (P.mc_shadow) {
            uint2 uu;
            *reinterpret_cast<__nv_bfloat162*>(&uu.x) = __floats2bfloat162_rn(p4.x, p4.y);
            *reinterpret_cast<__nv_bfloat162*>(&uu.y) = __floats2bfloat162_rn(p4.z, p4.w);
            multimem_st_bf16x4(P.mc_shadow + i4 * 4, uu);
          }
        }
      }
    }
    rank_barrier(P.pads, P.rank, W, e0 + 2);
  } else {
    // ---- two-shot: my slice only; result pushed to every peer.  U vectors per thread per trip keep
    // 16 x 16 B peer loads in flight per thread (NVLink latency ~2-3 us needs ~2 MB in flight per GPU).
    constexpr int U = (W <= 2) ? 8 : (W <= 4 ? 4 : 2);
    const long long slice4 = (P.n / W) >> 2;
    const long long base4 = slice4 * P.rank;
    for (long long j0 = 0; j0 < slice4; j0 += nthreads * U) {
      float4 acc[U];
      float4 v[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = j0 + (long long)u * nthreads + tid;
        if (j < slice4) {
#pragma unroll
          for (int p = 0; p < W; ++p) v[u][p] = ld_peer_f4(reinterpret_cast<const float*>(P.grads.p[p]) + (base4 + j) * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long j = j0 + (long long)u * nthreads + tid;
        if (j >= slice4) continue;
        float4 s = v[u][0];
#pragma unroll
        for (int p = 1; p < W; ++p) { s.x += v[u][p].x; s.y += v[u][p].y; s.z += v[u][p].z; s.w += v[u][p].w; }
        s.x *= P.scale; s.y *= P.scale; s.z *= P.scale; s.w *= P.scale;
        acc[u] = s;
        const long long i4 = base4 + j;
        if (P.opt.kind == 0) {
#pragma unroll
          for (int p = 0; p < W; ++p) st_peer_f4(reinterpret_cast<float*>(P.grads.p[p]) + i4 * 4, s);
        } else {
          float* pp = reinterpret_cast<float*>(P.params.p[P.rank]) + i4 * 4;
          float4 p4 = *reinterpret_cast<float4*>(pp);
          float4 m4 = *reinterpret_cast<float4*>(P.m + i4 * 4);
          float4 v4 = P.opt.kind == 1 ? *reinterpret_cast<float4*>(P.v + i4 * 4) : make_float4(0, 0, 0, 0);
          p4.x = opt_one(p4.x, s.x, m4.x, v4.x, P.opt); p4.y = opt_one(p4.y, s.y, m4.y, v4.y, P.opt);
          p4.z = opt_one(p4.z, s.z, m4.z, v4.z, P.opt); p4.w = opt_one(p4.w, s.w, m4.w, v4.w, P.opt);
          *reinterpret_cast<float4*>(P.m + i4 * 4) = m4;
          if (P.opt.kind == 1) *reinterpret_cast<float4*>(P.v + i4 * 4) = v4;
          uint2 uu;
          *reinterpret_cast<__nv_bfloat162*>(&uu.x) = __floats2bfloat162_rn(p4.x, p4.y);
          *reinterpret_cast<__nv_bfloat162*>(&uu.y) = __floats2bfloat162_rn(p4.z, p4.w);
#pragma unroll
          for (int p = 0; p < W; ++p) {
            st_peer_f4(reinterpret_cast<float*>(P.params.p[p]) + i4 * 4, p4);
            if (P.shadows.p[p]) st_peer_u2(reinterpret_cast<__nv_bfloat16*>(P.shadows.p[p]) + i4 * 4, uu);
          }
        }
      }
      (void)acc;
    }
    rank_barrier(P.pads, P.rank, W, e0 + 2);          // every slice has landed everywhere
  }
  if (threadIdx.x == 0) P.epochs[blockIdx.x] = e0 + 2;
}

// Broadcast: every rank pulls root's buffer (root == me: no copy).  Barriers fence root's writes before
// and root's reuse after.
struct BcParams { PeerBufs bufs; PeerBufs pads; uint32_t* epochs; long long n4; int rank, W, root; };

__global__ void __launch_bounds__(MPB_COMM_THREADS, 1) broadcast_kernel(BcParams P) {
  const uint32_t e0 = P.epochs[blockIdx.x];
  rank_barrier(P.pads, P.rank, P.W, e0 + 1);
  if (P.rank != P.root) {
    const float* src = reinterpret_cast<const float*>(P.bufs.p[P.root]);
    float* dst = reinterpret_cast<float*>(P.bufs.p[P.rank]);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < P.n4; i += (long long)gridDim.x * blockDim.x)
      *reinterpret_cast<float4*>(dst + i * 4) = ld_peer_f4(src + i * 4);
  }
  rank_barrier(P.pads, P.rank, P.W, e0 + 2);
  if (threadIdx.x == 0) P.epochs[blockIdx.x] = e0 + 2;
}

// All-gather of equal slices in place (slice r of every peer <- rank r's slice); used to rebuild
// full optimizer state from ZeRO-1 slices before a checkpoint.
__global__ void __launch_bounds__(MPB_COMM_THREADS, 1) allgather_kernel(BcParams P) {
  const uint32_t e0 = P.epochs[blockIdx.x];
  rank_barrier(P.pads, P.rank, P.W, e0 + 1);
  const long long slice4 = P.n4 / P.W;
  float* dst = reinterpret_cast<float*>(P.bufs.p[P.rank]);
  for (int k = 1; k < P.W; ++k) {
    const int p = (P.rank + k) % P.W;
    const float* src = reinterpret_cast<const float*>(P.bufs.p[p]) + slice4 * p * 4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < slice4; i += (long long)gridDim.x * blockDim.x)
      *reinterpret_cast<float4*>(dst + (slice4 * p + i) * 4) = ld_peer_f4(src + i * 4);
  }
  rank_barrier(P.pads, P.rank, P.W, e0 + 2);
  if (threadIdx.x == 0) P.epochs[blockIdx.x] = e0 + 2;
}

// int64 sum of one value per rank (correct counts of the eval pipeline); one block, one-shot.
struct CountParams { PeerBufs vals; PeerBufs pads; uint32_t* epochs; long long* out; int rank, W; };
__global__ void count_allreduce_kernel(CountParams P) {
  const uint32_t e0 = P.epochs[0];
  rank_barrier(P.pads, P.rank, P.W, e0 + 1);
  long long s = 0;
  if (threadIdx.x == 0)
    for (int p = 0; p < P.W; ++p) {
      long long v;
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(P.vals.p[p]) : "memory");
      s += v;
    }
  rank_barrier(P.pads, P.rank, P.W, e0 + 2);
  if (threadIdx.x == 0) { *P.out = s; P.epochs[0] = e0 + 2; }
}

// bf16 buffers (gradients kept in bf16 by the caller: half the NVLink bytes of the fp32 path): two-shot average with
// fp32 accumulation in rank order, result rounded once and pushed to every peer.
struct Bf16Params { PeerBufs bufs; PeerBufs pads; uint32_t* epochs; long long n8; int rank, W; float scale; };

__device__ __forceinline__ uint4 ld_peer_u4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_u4(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int W>
__global__ void __launch_bounds__(MPB_COMM_THREADS, 1) allreduce_bf16_kernel(Bf16Params P) {
  const uint32_t e0 = P.epochs[blockIdx.x];
  rank_barrier(P.pads, P.rank, W, e0 + 1);
  constexpr int U = (W <= 2) ? 8 : (W <= 4 ? 4 : 2);          // U * W 16-byte peer loads in flight per thread
  const long long slice8 = P.n8 / W, base = slice8 * P.rank;
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x, nthreads = (long long)gridDim.x * blockDim.x;
  for (long long j0 = 0; j0 < slice8; j0 += nthreads * U) {
    uint4 u[U][W];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const long long j = j0 + (long long)k * nthreads + tid;
      if (j < slice8) {
#pragma unroll
        for (int p = 0; p < W; ++p) u[k][p] = ld_peer_u4(reinterpret_cast<const uint4*>(P.bufs.p[p]) + base + j);
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const long long j = j0 + (long long)k * nthreads + tid;
      if (j >= slice8) continue;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int p = 0; p < W; ++p) {                            // rank order: identical bits on every rank
        const unsigned w[4] = {u[k][p].x, u[k][p].y, u[k][p].z, u[k][p].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[2 * i] += __uint_as_float(w[i] << 16); acc[2 * i + 1] += __uint_as_float(w[i] & 0xFFFF0000u); }
      }
      uint4 r;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
      for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(acc[2 * i] * P.scale, acc[2 * i + 1] * P.scale);
#pragma unroll
      for (int p = 0; p < W; ++p) st_peer_u4(reinterpret_cast<uint4*>(P.bufs.p[p]) + base + j, r);
    }
  }
  rank_barrier(P.pads, P.rank, W, e0 + 2);
  if (threadIdx.x == 0) P.epochs[blockIdx.x] = e0 + 2;
}

// ---------------------------------------------------------------------------------------------------------
// bf16 WIRE format inside the fused engine (GRAD_DTYPE=bf16).  wgrad keeps accumulating split-K partial sums in
// the fp32 gradient arena; this kernel (1) packs its own fp32 bucket to a symmetric bf16 wire buffer, (2) after the
// barrier reduces slice `rank` from every peer's wire buffer with fp32 accumulation in rank order -- 2 bytes per
// element cross NVLink in the reduce-scatter instead of 4 -- (3) applies 1/W and Adam / SGD to the fp32 masters of
// its slice and (4) all-gathers updated fp32 params + the bf16 compute copy by peer stores (or, algo 2, does (2)
// with multimem.ld_reduce.acc::f32 bf16x2 through the switch and (4) with multimem.st).
// The pack uses the same element -> block mapping as the reduce, so the per-block flag barrier is sufficient.
struct WireParams {
  const float* grads;           // local fp32 gradient bucket
  PeerBufs wire;                // bf16* per rank (symmetric)
  PeerBufs params, shadows, pads;
  uint32_t* epochs;
  float* m; float* v;
  long long n;                  // multiple of 8*W
  int rank, W, algo;            // 1 = peer loads/stores, 2 = NVLS
  float scale;
  __nv_bfloat16* mc_wire; float* mc_params; __nv_bfloat16* mc_shadow;
  OptArgs opt;
};

__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}

template <int W>
__global__ void __launch_bounds__(MPB_COMM_THREADS, 1) allreduce_wire16_kernel(WireParams P) {
  if (P.opt.kind != 0 && P.opt.step_ptr != nullptr) {
    const int step = *P.opt.step_ptr;
    P.opt.first_step = (step == 1);
    P.opt.bc1 = 1.f - powf(P.opt.b1, (float)step);
    P.opt.bc2_sqrt = sqrtf(1.f - powf(P.opt.b2, (float)step));
  }
  const uint32_t e0 = P.epochs[blockIdx.x];
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * blockDim.x;
  const long long slice8 = (P.n / W) >> 3;
  // (1) pack: fp32 -> bf16, my copy of every slice, element -> thread mapping identical to the reduce below
  {
    uint4* dst = reinterpret_cast<uint4*>(P.wire.p[P.rank]);
    for (int q = 0; q < W; ++q)
      for (long long j = tid; j < slice8; j += nthreads) {
        const long long i8 = (long long)q * slice8 + j;
        const float4 a = *reinterpret_cast<const float4*>(P.grads + i8 * 8);
        const float4 b = *reinterpret_cast<const float4*>(P.grads + i8 * 8 + 4);
        uint4 r;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
        h[0] = __floats2bfloat162_rn(a.x, a.y); h[1] = __floats2bfloat162_rn(a.z, a.w);
        h[2] = __floats2bfloat162_rn(b.x, b.y); h[3] = __floats2bfloat162_rn(b.z, b.w);
        dst[i8] = r;
      }
  }
  rank_barrier(P.pads, P.rank, W, e0 + 1);
  constexpr int U = (W <= 2) ? 4 : 2;
  const long long base8 = slice8 * P.rank;
  for (long long j0 = 0; j0 < slice8; j0 += nthreads * U) {
    uint4 u[U][W];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const long long j = j0 + (long long)k * nthreads + tid;
      if (j < slice8) {
        if (P.algo == 2) {
          u[k][0] = multimem_ld_reduce_bf16x8(reinterpret_cast<const uint4*>(P.mc_wire) + base8 + j);
        } else {
#pragma unroll
          for (int p = 0; p < W; ++p) u[k][p] = ld_peer_u4(reinterpret_cast<const uint4*>(P.wire.p[p]) + base8 + j);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const long long j = j0 + (long long)k * nthreads + tid;
      if (j >= slice8) continue;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const int np = (P.algo == 2) ? 1 : W;
#pragma unroll
      for (int p = 0; p < W; ++p) {
        if (p >= np) break;
        const unsigned w[4] = {u[k][p].x, u[k][p].y, u[k][p].z, u[k][p].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { acc[2 * i] += __uint_as_float(w[i] << 16); acc[2 * i + 1] += __uint_as_float(w[i] & 0xFFFF0000u); }
      }
      const long long i8 = base8 + j;
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const long long i4 = i8 * 2 + hh;
        float4 s = make_float4(acc[4 * hh] * P.scale, acc[4 * hh + 1] * P.scale, acc[4 * hh + 2] * P.scale, acc[4 * hh + 3] * P.scale);
        float* pp = reinterpret_cast<float*>(P.params.p[P.rank]) + i4 * 4;
        float4 p4 = *reinterpret_cast<float4*>(pp);
        float4 m4 = *reinterpret_cast<float4*>(P.m + i4 * 4);
        float4 v4 = P.opt.kind == 1 ? *reinterpret_cast<float4*>(P.v + i4 * 4) : make_float4(0, 0, 0, 0);
        p4.x = opt_one(p4.x, s.x, m4.x, v4.x, P.opt); p4.y = opt_one(p4.y, s.y, m4.y, v4.y, P.opt);
        p4.z = opt_one(p4.z, s.z, m4.z, v4.z, P.opt); p4.w = opt_one(p4.w, s.w, m4.w, v4.w, P.opt);
        *reinterpret_cast<float4*>(P.m + i4 * 4) = m4;
        if (P.opt.kind == 1) *reinterpret_cast<float4*>(P.v + i4 * 4) = v4;
        uint2 uu;
        *reinterpret_cast<__nv_bfloat162*>(&uu.x) = __floats2bfloat162_rn(p4.x, p4.y);
        *reinterpret_cast<__nv_bfloat162*>(&uu.y) = __floats2bfloat162_rn(p4.z, p4.w);
        if (P.algo == 2) {
          multimem_st_f4(P.mc_params + i4 * 4, p4);
          if (P.mc_shadow) multimem_st_bf16x4(P.mc_shadow + i4 * 4, uu);
        } else {
#pragma unroll
          for (int p = 0; p < W; ++p) {
            st_peer_f4(reinterpret_cast<float*>(P.params.p[p]) + i4 * 4, p4);
            if (P.shadows.p[p]) st_peer_u2(reinterpret_cast<__nv_bfloat16*>(P.shadows.p[p]) + i4 * 4, uu);
          }
        }
      }
    }
  }
  rank_barrier(P.pads, P.rank, W, e0 + 2);
  if (threadIdx.x == 0) P.epochs[blockIdx.x] = e0 + 2;
}

template <int W>
int launch_ar(const ArParams& P, int grid, cudaStream_t st) {
  allreduce_kernel<W><<<grid, MPB_COMM_THREADS, 0, st>>>(P);
  return (int)cudaGetLastError();
}

}  // namespace

// ------------------------------------------------------------------------------------------ C ABI
// Symmetric memory helpers (plain cudaMalloc so cudaIpc handles cover exactly this allocation).
MPB_API int mpb_sym_alloc(void** ptr, long long bytes) {
  MPB_CUDA(cudaMalloc(ptr, (size_t)bytes));
  MPB_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
  return 0;
}
MPB_API int mpb_sym_free(void* ptr) { return (int)cudaFree(ptr); }
MPB_API int mpb_ipc_get_handle(void* ptr, void* handle_out64) {
  cudaIpcMemHandle_t h;
  MPB_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out64, &h, sizeof(h));
  return 0;
}
MPB_API int mpb_ipc_open_handle(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  MPB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
MPB_API int mpb_ipc_close_handle(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }
MPB_API int mpb_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
MPB_API int mpb_signal_pad_bytes(void) { return MPB_MAX_COMM_BLOCKS * MPB_MAX_WORLD * (int)sizeof(uint32_t); }
MPB_API int mpb_epoch_bytes(void) { return MPB_MAX_COMM_BLOCKS * (int)sizeof(uint32_t); }
MPB_API long long mpb_oneshot_max_elems(int blocks) { return (long long)blocks * MPB_COMM_THREADS * 4; }

// grads/params/shadows/pads: arrays of W device pointers (index = rank), already offset to the bucket.
// opt_kind 0: grads are averaged in place.  1/2: params (fp32) + optional bf16 shadows are updated, m/v
// are this rank's local moments (bucket-offset; v may be null for sgd).  n must be a multiple of 4*W.
MPB_API int mpb_allreduce_fused(void* const* grads, void* const* params, void* const* shadows, void* const* pads,
                                void* epochs, float* m, float* v, long long n, int rank, int W, int algo, int blocks,
                                int opt_kind, int step, float lr, float b1, float b2, float eps, float wd,
                                float momentum, const int* step_dev, void* mc_grads, void* mc_params, void* mc_shadow,
                                void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD || n <= 0 || (n % (4 * W)) != 0) return -1;
  if (blocks < 1) blocks = 1;
  if (blocks > MPB_MAX_COMM_BLOCKS) blocks = MPB_MAX_COMM_BLOCKS;
  if (algo == 0 && n > mpb_oneshot_max_elems(blocks)) return -2;
  if (algo == 2 && (mc_grads == nullptr || (opt_kind != 0 && mc_params == nullptr))) return -3;
  ArParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) {
    P.grads.p[i] = grads[i];
    P.params.p[i] = params ? params[i] : nullptr;
    P.shadows.p[i] = shadows ? shadows[i] : nullptr;
    P.pads.p[i] = pads[i];
  }
  P.epochs = (uint32_t*)epochs; P.m = m; P.v = v; P.n = n; P.rank = rank; P.W = W; P.algo = algo;
  P.mc_grads = (float*)mc_grads; P.mc_params = (float*)mc_params; P.mc_shadow = (__nv_bfloat16*)mc_shadow;
  P.scale = 1.f / (float)W;
  P.opt.kind = opt_kind; P.opt.lr = lr; P.opt.b1 = b1; P.opt.b2 = b2; P.opt.eps = eps; P.opt.wd = wd;
  P.opt.momentum = momentum; P.opt.first_step = (step == 1); P.opt.step_ptr = step_dev;
  P.opt.bc1 = 1.f - powf(b1, (float)step);
  P.opt.bc2_sqrt = sqrtf(1.f - powf(b2, (float)step));
  cudaStream_t st = (cudaStream_t)stream;
  switch (W) {
    case 1: return launch_ar<1>(P, blocks, st);
    case 2: return launch_ar<2>(P, blocks, st);
    case 3: return launch_ar<3>(P, blocks, st);
    case 4: return launch_ar<4>(P, blocks, st);
    case 5: return launch_ar<5>(P, blocks, st);
    case 6: return launch_ar<6>(P, blocks, st);
    case 7: return launch_ar<7>(P, blocks, st);
    default: return launch_ar<8>(P, blocks, st);
  }
}

// Fused reduce + optimizer with a bf16 wire format: `grads` is this rank's LOCAL fp32 bucket, `wire` the W symmetric
// bf16 staging buffers (index = rank) the peers read; everything else as mpb_allreduce_fused.  n % (8*W) == 0,
// opt_kind must be 1 or 2 (the averaged gradient itself is never materialised), algo 1 (peer) or 2 (NVLS).
MPB_API int mpb_allreduce_fused_bf16wire(const float* grads, void* const* wire, void* const* params,
                                         void* const* shadows, void* const* pads, void* epochs, float* m, float* v,
                                         long long n, int rank, int W, int algo, int blocks, int opt_kind, int step,
                                         float lr, float b1, float b2, float eps, float wd, float momentum,
                                         const int* step_dev, void* mc_wire, void* mc_params, void* mc_shadow,
                                         void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD || n <= 0 || (n % (8 * W)) != 0) return -1;
  if (opt_kind != 1 && opt_kind != 2) return -1;
  if (algo == 2 && (mc_wire == nullptr || mc_params == nullptr)) return -3;
  if (algo != 2) algo = 1;
  if (blocks < 1) blocks = 1;
  if (blocks > MPB_MAX_COMM_BLOCKS) blocks = MPB_MAX_COMM_BLOCKS;
  WireParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) {
    P.wire.p[i] = wire[i];
    P.params.p[i] = params[i];
    P.shadows.p[i] = shadows ? shadows[i] : nullptr;
    P.pads.p[i] = pads[i];
  }
  P.grads = grads; P.epochs = (uint32_t*)epochs; P.m = m; P.v = v; P.n = n; P.rank = rank; P.W = W; P.algo = algo;
  P.mc_wire = (__nv_bfloat16*)mc_wire; P.mc_params = (float*)mc_params; P.mc_shadow = (__nv_bfloat16*)mc_shadow;
  P.scale = 1.f / (float)W;
  P.opt.kind = opt_kind; P.opt.lr = lr; P.opt.b1 = b1; P.opt.b2 = b2; P.opt.eps = eps; P.opt.wd = wd;
  P.opt.momentum = momentum; P.opt.first_step = (step == 1); P.opt.step_ptr = step_dev;
  P.opt.bc1 = 1.f - powf(b1, (float)step);
  P.opt.bc2_sqrt = sqrtf(1.f - powf(b2, (float)step));
  cudaStream_t st = (cudaStream_t)stream;
  switch (W) {
    case 1: allreduce_wire16_kernel<1><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 2: allreduce_wire16_kernel<2><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 3: allreduce_wire16_kernel<3><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 4: allreduce_wire16_kernel<4><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 5: allreduce_wire16_kernel<5><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 6: allreduce_wire16_kernel<6><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 7: allreduce_wire16_kernel<7><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    default: allreduce_wire16_kernel<8><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
  }
  return (int)cudaGetLastError();
}

// In-place average of W symmetric bf16 buffers of n elements (n a multiple of 8*W).
MPB_API int mpb_allreduce_bf16(void* const* bufs, void* const* pads, void* epochs, long long n, int rank, int W, int blocks,
                               void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD || n <= 0 || (n % (8 * W)) != 0) return -1;
  if (blocks < 1) blocks = 1;
  if (blocks > MPB_MAX_COMM_BLOCKS) blocks = MPB_MAX_COMM_BLOCKS;
  Bf16Params P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) { P.bufs.p[i] = bufs[i]; P.pads.p[i] = pads[i]; }
  P.epochs = (uint32_t*)epochs; P.n8 = n / 8; P.rank = rank; P.W = W; P.scale = 1.f / (float)W;
  cudaStream_t st = (cudaStream_t)stream;
  switch (W) {
    case 1: allreduce_bf16_kernel<1><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 2: allreduce_bf16_kernel<2><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 3: allreduce_bf16_kernel<3><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 4: allreduce_bf16_kernel<4><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 5: allreduce_bf16_kernel<5><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 6: allreduce_bf16_kernel<6><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    case 7: allreduce_bf16_kernel<7><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
    default: allreduce_bf16_kernel<8><<<blocks, MPB_COMM_THREADS, 0, st>>>(P); break;
  }
  return (int)cudaGetLastError();
}

MPB_API int mpb_broadcast(void* const* bufs, void* const* pads, void* epochs, long long n_floats, int rank, int W,
                          int root, int blocks, void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD || (n_floats % 4) != 0) return -1;
  if (blocks < 1) blocks = 1;
  if (blocks > MPB_MAX_COMM_BLOCKS) blocks = MPB_MAX_COMM_BLOCKS;
  BcParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) { P.bufs.p[i] = bufs[i]; P.pads.p[i] = pads[i]; }
  P.epochs = (uint32_t*)epochs; P.n4 = n_floats / 4; P.rank = rank; P.W = W; P.root = root;
  broadcast_kernel<<<blocks, MPB_COMM_THREADS, 0, (cudaStream_t)stream>>>(P);
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_allgather_slices(void* const* bufs, void* const* pads, void* epochs, long long n_floats, int rank, int W,
                                 int blocks, void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD || (n_floats % (4 * W)) != 0) return -1;
  if (blocks < 1) blocks = 1;
  if (blocks > MPB_MAX_COMM_BLOCKS) blocks = MPB_MAX_COMM_BLOCKS;
  BcParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) { P.bufs.p[i] = bufs[i]; P.pads.p[i] = pads[i]; }
  P.epochs = (uint32_t*)epochs; P.n4 = n_floats / 4; P.rank = rank; P.W = W; P.root = 0;
  allgather_kernel<<<blocks, MPB_COMM_THREADS, 0, (cudaStream_t)stream>>>(P);
  MPB_CHECK_LAUNCH();
  return 0;
}

MPB_API int mpb_count_allreduce(void* const* vals, void* const* pads, void* epochs, long long* out, int rank, int W,
                                void* stream) {
  if (W < 1 || W > MPB_MAX_WORLD) return -1;
  CountParams P;
  memset(&P, 0, sizeof(P));
  for (int i = 0; i < W; ++i) { P.vals.p[i] = vals[i]; P.pads.p[i] = pads[i]; }
  P.epochs = (uint32_t*)epochs; P.out = out; P.rank = rank; P.W = W;
  count_allreduce_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P);
  MPB_CHECK_LAUNCH();
  return 0;
}
