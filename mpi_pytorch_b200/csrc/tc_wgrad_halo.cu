// Halo-tile weight gradient for stride-1 filters larger than 1x1, sm_100a (tcgen05 + TMEM + TMA).
//
//      dw[k][r][s][c] = sum over output pixels (n, p, q) of  dy[n][p][q][k] * x[n][p + r - ph][q + s - pw][c]
//
// The im2col formulation in tc_gemm.cu launches one CTA per (filter tap, split of the pixel range): every tap re-reads
// the activations through TMA im2col mode (R*S reads of x from L2), a 64-channel layer fills half of the 128-row MMA,
// and the epilogue issues strided fp32 atomics.  Here the reduction (GEMM-K) index runs over the *padded* pixel grid
// of a tile of TR full image rows, kk = tr * pitch + j with pitch = Wo + S - 1:
//   * dy tile: one tiled TMA box {64 k, pitch, TR, 1} at (w = 0, h = h0).  Columns j >= Wo and rows >= Ho are out of
//     range, so TMA zero-fills them -- those are exactly the padded-grid positions that must not contribute.
//   * x tile:  one tiled TMA box {64 c, pitch, rows, 1} at (w = -pw, h = h0 - ph + r0); out-of-range = conv padding.
//   * tap (r, s): dy[kk] pairs with x[kk + (r - r0) * pitch + s], i.e. the SAME x tile viewed through a UMMA
//     descriptor whose start address is ((r - r0) * pitch + s) * 128 bytes further in (128B swizzle is a function of
//     the absolute shared-memory address, so shifted views de-swizzle correctly -- measured in tc_gemm.cu/tc_halo.cu).
// Both operands are MN-major (the reduction index is the strided one in NHWC), M = input channels, N = output
// channels, one fp32 accumulator per tap in TMEM, kept there across ALL pixel tiles the CTA walks (persistent CTA);
// one coalesced fp32 RED epilogue per CTA at the very end.
//
// Two decompositions of the work over blockIdx.y:
//   row mode  (C % 128 == 0): CTA = (128-channel block, k block, filter row r); S accumulators of 128 x BNk.
//   pair mode (C == 64):      CTA = all R*S taps; two taps share one M = 128 MMA -- the second 64-row panel of the
//                             MN-major A operand is the same x tile at the NEXT tap's shift, expressed through the
//                             descriptor's leading-dimension byte offset (LBO = shift difference * 128 B).
// blockIdx.x strides over the (image, row-tile) list.  Reference call this replaces: loss.backward() (main.py:152)
// reaching ATen's CPU convolution_backward for every 3x3 nn.Conv2d built in models.py:33..87.
#include "tc_ptx.cuh"

using namespace mpb;
using namespace mpbtc;

namespace {

constexpr int kThreads = 192;
constexpr int kMaxJobs = 8;
constexpr int kSmemBudget = 206 * 1024;      // dynamic shared memory the stages may use (of 227 KB per CTA)

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

struct WgParams {
  int N, Ho, Wo, C, K, R, S, ph, pw;
  int TR, pitch, tiles_h, kpix;            // rows per tile, padded row length, tiles per image, GEMM-K per tile (x16)
  int pair;                                // 1 = pair mode
  int BNk;                                 // output channels per CTA (64 or 128)
  int x_panels, dy_panels, x_panel_bytes, dy_panel_bytes, x_box_bytes, dy_box_bytes, stage_bytes, stages;
  int kgroups, rgroups;                    // blockIdx.y = (cg * kgroups + kg) * rgroups + rg
  int njobs;
  int job_shift[kMaxJobs];                 // A-view shift in pixel rows (128 B each)
  int job_lbo[kMaxJobs];                   // A descriptor leading-dimension byte offset
  uint32_t idesc;
  int dbg;                                 // MPB_WG_DBG: bit 0 skip the MMAs, bit 1 skip the loads (timing experiments)
  float* dw;
};

__global__ void __launch_bounds__(kThreads, 1)
tc_wgrad_halo_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, const WgParams p,
                     int num_tiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full_bar[8], empty_bar[8], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int y = blockIdx.y;
  const int rg = y % p.rgroups; y /= p.rgroups;
  const int kg = y % p.kgroups;
  const int cg = y / p.kgroups;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmX);
    prefetch_tmap(&tmDY);
    for (int i = 0; i < p.stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  // The shifted views and the round-up of the pixel count to 16 read a few rows past each TMA box: those bytes are
  // never written by TMA, so zero them once (the dy side must be exactly zero, the x side merely finite).
  {
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = p.stages * p.stage_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int my_tiles = ((int)blockIdx.x < num_tiles) ? (num_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (elect_one()) {
      const uint32_t tx = (uint32_t)(p.x_panels * p.x_box_bytes + p.dy_panels * p.dy_box_bytes);
      const int c_base = p.pair ? 0 : cg * 128, k_base = kg * p.BNk;
      const int r0 = p.pair ? 0 : rg;
      for (int li = 0; li < my_tiles; ++li) {
        const int t = (int)blockIdx.x + li * (int)gridDim.x;
        const int img = t / p.tiles_h, h0 = (t % p.tiles_h) * p.TR;
        const int st = li % p.stages;
        mbar_wait(&empty_bar[st], ((li / p.stages) & 1) ^ 1);
        if (p.dbg & 2) { mbar_expect_tx(&full_bar[st], 0); continue; }
        mbar_expect_tx(&full_bar[st], tx);
        uint8_t* xs = smem + (size_t)st * p.stage_bytes;
        uint8_t* ds = xs + (size_t)p.x_panels * p.x_panel_bytes;
        for (int j = 0; j < p.x_panels; ++j)
          tma_load_4d(&tmX, &full_bar[st], xs + (size_t)j * p.x_panel_bytes, c_base + j * 64, -p.pw, h0 - p.ph + r0, img);
        for (int j = 0; j < p.dy_panels; ++j)
          tma_load_4d(&tmDY, &full_bar[st], ds + (size_t)j * p.dy_panel_bytes, k_base + j * 64, 0, h0, img);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const int ksteps = p.kpix / 16;
    for (int li = 0; li < my_tiles; ++li) {
      const int st = li % p.stages;
      mbar_wait(&full_bar[st], (li / p.stages) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t xbase = smem_u32(smem + (size_t)st * p.stage_bytes);
        const uint32_t dbase = xbase + (uint32_t)(p.x_panels * p.x_panel_bytes);
        const uint64_t bdesc = make_desc(dbase, (uint32_t)p.dy_panel_bytes, 1024);
        uint64_t adesc[kMaxJobs];
#pragma unroll
        for (int j = 0; j < kMaxJobs; ++j)
          adesc[j] = (j < p.njobs) ? make_desc(xbase + (uint32_t)p.job_shift[j] * 128u, (uint32_t)p.job_lbo[j], 1024) : 0;
        const int nks = (p.dbg & 1) ? (li == 0 ? 1 : 0) : ksteps;
        for (int ks = 0; ks < nks; ++ks) {
          const uint64_t koff = (uint64_t)(ks * 128);            // 16 pixel rows * 128 B, in 16-byte units
          const uint32_t acc = (li | ks) != 0;
#pragma unroll
          for (int j = 0; j < kMaxJobs; ++j)
            if (j < p.njobs) umma_bf16(tmem_base + (uint32_t)(j * p.BNk), adesc[j] + koff, bdesc + koff, p.idesc, acc);
        }
        umma_commit(&empty_bar[st]);                             // the stage is free once these MMAs have read it
        if (li == my_tiles - 1) umma_commit(&tmem_full_bar);
      }
      __syncwarp();
    }
  } else if (my_tiles > 0) {
    // ================================ epilogue (once per CTA) ================================
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    const int q = warp & 3;                                      // TMEM lane quarter this warp may read
    const int m = q * 32 + lane;
    const int RS = p.R * p.S;
    // Every CTA of a column group adds into the same dw slots: rotate the (job, chunk) order by blockIdx.x so CTAs
    // that finish together start on different addresses instead of queueing on one L2 atomic unit.
    const int nchunks = p.BNk / 32, nphase = p.njobs * nchunks;
    for (int ph = 0; ph < nphase; ++ph) {
      const int rot = (ph + (int)blockIdx.x) % nphase;
      const int j = rot / nchunks, cb = (rot % nchunks) * 32;
      int tap, c;
      if (p.pair) { tap = 2 * j + (m >> 6); c = m & 63; }
      else        { tap = rg * p.S + j;     c = cg * 128 + m; }
      const bool live = tap < RS && c < p.C;
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * p.BNk + cb), v);
      if (live) {
        const int k0 = kg * p.BNk + cb;
        float* dst = p.dw + ((size_t)k0 * RS + tap) * p.C + c;
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (k0 + i < p.K) atomicAdd(dst + (size_t)i * RS * p.C, __uint_as_float(v[i]));     // 32 lanes = 32 consecutive c
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}


}  // namespace

namespace {

inline int round_up(int v, int a) { return (v + a - 1) / a * a; }

// Fill the geometry for one layer; false when the layer does not fit this kernel (the caller then uses tc_gemm.cu).
bool plan_wgrad(int C, int K, int R, int S, int Ho, int Wo, WgParams* out) {
  WgParams p;
  memset(&p, 0, sizeof(p));
  if (R * S <= 1 || C % 64 != 0 || K % 64 != 0) return false;
  p.pair = (C == 64) ? 1 : 0;
  if (!p.pair && C % 128 != 0) return false;
  p.BNk = (K % 128 == 0) ? 128 : 64;
  p.pitch = Wo + S - 1;
  if (p.pitch > 256) return false;
  p.njobs = p.pair ? (R * S + 1) / 2 : S;
  if (p.njobs > kMaxJobs || p.njobs * p.BNk > 512) {
    if (p.BNk == 128 && p.njobs <= kMaxJobs && p.njobs * 64 <= 512) p.BNk = 64;
    else return false;
  }
  p.x_panels = p.pair ? 1 : 2;
  p.dy_panels = p.BNk / 64;
  const int halo_rows = p.pair ? R - 1 : 0;
  const int max_shift = p.pair ? (R - 1) * p.pitch + S : S - 1;     // pair mode: +1 for the dummy partner of an odd tap
  // rows per tile: the most pixel-efficient choice that still leaves a >= 2-deep ring, preferring >= 3 stages
  const int forced = env_int("MPB_WG_TR", 0);
  double best = -1.0;
  for (int TR = 1; TR <= Ho && TR + halo_rows <= 256; ++TR) {
    if (forced && TR != forced) continue;
    const int kpix = round_up(TR * p.pitch, 16);
    const int xrows = TR + halo_rows;
    const int need = (xrows * p.pitch > kpix + max_shift) ? xrows * p.pitch : kpix + max_shift;
    const int xpb = round_up(need * 128, 1024), dpb = round_up(kpix * 128, 1024);
    const int stage = p.x_panels * xpb + p.dy_panels * dpb;
    const int stages = kSmemBudget / stage;
    if (stages < 2) break;
    const int tiles_h = (Ho + TR - 1) / TR;
    // cost model fitted on B200: a tile costs its MMAs (kpix/16 per job) plus ~4.5 MMA-times of hand-over, a
    // 2-stage ring exposes some load latency, halo rows are extra x traffic; score = 1 / cost
    double cost = (double)tiles_h * ((double)(kpix / 16) * p.njobs + 4.5);
    if (stages < 3) cost *= 1.1;
    cost *= 1.0 + 0.15 * (double)(xrows - TR) / (double)xrows;
    const double score = 1.0 / cost;
    if (score > best) {
      best = score;
      p.TR = TR; p.kpix = kpix; p.tiles_h = tiles_h; p.x_panel_bytes = xpb; p.dy_panel_bytes = dpb;
      p.x_box_bytes = xrows * p.pitch * 128; p.dy_box_bytes = TR * p.pitch * 128;
      p.stage_bytes = stage; p.stages = stages > 6 ? 6 : stages;
    }
  }
  if (best < 0) return false;
  p.kgroups = (K + p.BNk - 1) / p.BNk;
  p.rgroups = p.pair ? 1 : R;
  if ((p.pair ? 1 : C / 128) * p.kgroups * p.rgroups > env_int("MPB_WG_MAX_Y", 40)) return false;   // few pixel groups per column group: tc_gemm.cu's split-K form is faster (l4)
  for (int j = 0; j < p.njobs; ++j) {
    if (p.pair) {
      const int t0 = 2 * j, t1 = t0 + 1;
      const int s0 = (t0 / S) * p.pitch + t0 % S;
      const int s1 = (t1 < R * S) ? (t1 / S) * p.pitch + t1 % S : s0 + 1;
      p.job_shift[j] = s0; p.job_lbo[j] = (s1 - s0) * 128;
    } else {
      p.job_shift[j] = j; p.job_lbo[j] = p.x_panel_bytes;
    }
  }
  p.idesc = make_idesc(p.BNk, 1, 1);
  p.dbg = env_int("MPB_WG_DBG", 0);
  *out = p;
  return true;
}

}  // namespace

MPB_API int mpb_tc_wgrad_halo_ok(int C, int K, int R, int S, int Ho, int Wo) {
  WgParams p;
  return plan_wgrad(C, K, R, S, Ho, Wo, &p) ? 1 : 0;
}

namespace {
void wg_grid(const WgParams& p, int N, int* gx, int* ygroups, int* num_tiles) {
  const int cgroups = p.pair ? 1 : p.C / 128;
  *ygroups = cgroups * p.kgroups * p.rgroups;
  *num_tiles = N * p.tiles_h;
  int g = kSMs / *ygroups;
  if (g < 1) g = 1;
  if (g > *num_tiles) g = *num_tiles;
  *gx = g;
}
}  // namespace

// dw[K,R,S,C] (fp32, zero-initialised or partially accumulated) += wgrad of a stride-1, dilation-1 convolution.
// Every CTA adds its accumulators into dw with coalesced fp32 RED operations (parking per-CTA partial sums in a
// workspace and folding them with a second kernel measured 5-20 us SLOWER per layer on B200).
static int wgrad_halo_launch(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R, int S,
                             int P, int Q, int pad_h, int pad_w, long long xs_w, long long xs_h, long long xs_n,
                             void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  WgParams p;
  if (!plan_wgrad(C, K, R, S, P, Q, &p)) return -2;
  p.N = N; p.Ho = P; p.Wo = Q; p.C = C; p.K = K; p.R = R; p.S = S; p.ph = pad_h; p.pw = pad_w; p.dw = dw;
  CUtensorMap tx, td;
  rc = make_map_nhwc_tiled(&tx, x, N, H, W, C, p.pitch, p.TR + (p.pair ? R - 1 : 0), xs_w, xs_h, xs_n);
  if (rc) return rc;
  rc = make_map_nhwc_tiled(&td, dy, N, P, Q, K, p.pitch, p.TR);
  if (rc) return rc - 1000;
  int gx, ygroups, num_tiles;
  wg_grid(p, N, &gx, &ygroups, &num_tiles);
  const int smem = p.stages * p.stage_bytes + 1024;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(tc_wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)gx, (unsigned)ygroups, 1);
  tc_wgrad_halo_kernel<<<grid, kThreads, smem, st>>>(tx, td, p, num_tiles);
  return (int)cudaGetLastError();
}

MPB_API int mpb_tc_conv_wgrad_halo(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R,
                                   int S, int P, int Q, int pad_h, int pad_w, void* stream) {
  return wgrad_halo_launch(dy, x, dw, N, H, W, C, K, R, S, P, Q, pad_h, pad_w, 0, 0, 0, stream);
}

// x given with explicit element strides between pixels / rows / images (see mpb_tc_conv_fprop_halo_strided).
MPB_API int mpb_tc_conv_wgrad_halo_strided(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K,
                                           int R, int S, int P, int Q, int pad_h, int pad_w, long long xs_w,
                                           long long xs_h, long long xs_n, void* stream) {
  return wgrad_halo_launch(dy, x, dw, N, H, W, C, K, R, S, P, Q, pad_h, pad_w, xs_w, xs_h, xs_n, stream);
}
