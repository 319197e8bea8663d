// Halo-tile implicit-GEMM convolution for stride-1 filters larger than 1x1 (fprop and dgrad), sm_100a.
//
// tc_gemm.cu feeds the tensor core through TMA im2col mode: simple, but an R x S filter re-reads every activation
// tile R*S times from L2, and on B200 those kernels sit at the L2->SM fabric limit (~12 TB/s, profiles/).  Here the
// activation tile is fetched ONCE per 64-channel block, with its halo, as a plain 4-D tiled TMA box
//      {64 ch, pitch = TW + S - 1 pixels, TR + R - 1 rows, 1 image}   starting at (w0 - pad_w, h0 - pad_h)
// (out-of-range coordinates are zero-filled, which is the convolution's padding).  In shared memory that is a dense,
// 128B-swizzled array of pixel rows.  The GEMM-M index runs over the *padded* output grid m = tr * pitch + j, so the
// A operand of filter tap (r, s) is the same array shifted by (r * pitch + s) rows -- a UMMA descriptor whose start
// address is (r * pitch + s) * 128 bytes further in.  (The 128B swizzle is a function of the absolute smem address,
// measured in tc_gemm.cu's a_shift experiment, so the shifted view de-swizzles correctly with base_offset = 0.)
// Columns j >= TW of each row and rows m >= TR * pitch compute garbage that the epilogue drops.
//
// Per 64-channel block: 1 activation load + R*S weight-tile loads + R*S*4 tcgen05.mma; activation traffic falls by
// ~R*S*(TR*TW)/((TR+R-1)*pitch)  (4-5x for 3x3).  Two mbarrier rings: A (halo tiles, 2 deep) and B (weight tiles).
// Warp roles as in tc_gemm.cu: warp 0 TMA producer, warp 1 TMEM + MMA issuer, warps 2-5 epilogue.
#include "tc_ptx.cuh"

using namespace mpb;
using namespace mpbtc;

namespace {

constexpr int kThreads = 192;
constexpr int kThreadsPersist = 320;      // persistent kernel: 8 epilogue warps, two per TMEM lane quarter (each takes half the columns)

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

struct HaloParams {
  int N, Ho, Wo, Co;          // output tensor [N, Ho, Wo, Co]
  int Ci;                     // reduction channels of the input tensor
  int R, S, ph, pw;           // filter taps; input pixel = output pixel + tap - pad
  int TR, TW, pitch;          // output tile TR x TW pixels; pitch = TW + S - 1 buffer pixels per row
  int tiles_h, tiles_w;
  int BN, b_mn, flip;         // tile width; weights as MN-major regions (dgrad); flip taps (dgrad)
  int w_cols;                 // dgrad: C (columns per tap in the [K][R*S*C] weight matrix)
  uint32_t idesc;
  int a_stages, b_stages, a_stage_bytes, b_stage_bytes, a_box_bytes;
  void* D; long long ldd;
  const float* bias; float* stats;
  int relu;                 // epilogue: max(x, 0) after the bias
};


// ------------------------------------------------------------------------------------------------------------
// Persistent variant for layers whose whole filter bank (all taps x channel blocks of this CTA's BN output channels)
// fits in shared memory next to the activation ring -- the 64-channel 3x3 layers (ResNet layer1, VGG block 1).
// One CTA per SM loops over output tiles:
//   * the weights are loaded ONCE per CTA and stay resident; only halo tiles stream (a_stages-deep ring);
//   * the accumulator is double-buffered in TMEM, so the epilogue of tile i overlaps the MMAs of tile i+1
//     (tmem_full / tmem_empty barriers);
//   * BatchNorm statistics are accumulated in shared memory over all tiles of the CTA and flushed once.
// Per tile the tensor core issues R*S*cblocks*4 MMAs while TMA brings one halo tile per channel block: the loop is
// MMA-bound instead of L2-fabric-bound.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreadsPersist, 1)
tc_halo_persist_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const HaloParams p,
                       int num_tiles) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int cblocks = (p.Ci + BK - 1) / BK, taps = p.R * p.S;
  uint8_t* smem_w = smem + (size_t)p.a_stages * p.a_stage_bytes;       // resident weights: [cb][tap] tiles of b_stage_bytes
  __shared__ uint64_t full_a[8], empty_a[8], full_w, tmem_full[4], tmem_empty[4];
  __shared__ uint32_t tmem_base_smem;
  __shared__ float stat_smem[2][256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * p.BN;
  uint32_t acc_cols = 32;
  while ((int)acc_cols < p.BN) acc_cols <<= 1;
  // accumulator ring in TMEM: 4 deep when it fits in the 512 columns (the MMA warp then never waits for the epilogue)
  const int n_acc = (acc_cols * 4 <= 512) ? 4 : 2;
  const uint32_t tmem_cols = acc_cols * (uint32_t)n_acc;

  if (threadIdx.x == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < p.a_stages; ++i) { mbar_init(&full_a[i], 1); mbar_init(&empty_a[i], 1); }
    mbar_init(&full_w, 1);
    for (int i = 0; i < 4; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 8); }   // 8 epilogue warps
    fence_barrier_init();
  }
  if (p.stats) for (int j = threadIdx.x; j < 256; j += blockDim.x) { stat_smem[0][j] = 0.f; stat_smem[1][j] = 0.f; }
  if (warp == 1) tmem_alloc(&tmem_base_smem, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (elect_one()) {
      mbar_expect_tx(&full_w, (uint32_t)(cblocks * taps * p.b_stage_bytes));
      for (int cb = 0; cb < cblocks; ++cb)
        for (int tap = 0; tap < taps; ++tap) {
          uint8_t* dst = smem_w + (size_t)(cb * taps + tap) * p.b_stage_bytes;
          if (!p.b_mn) {
            tma_load_2d(&tmB, &full_w, dst, tap * p.Ci + cb * BK, n0);
          } else {
            const int r = tap / p.S, s = tap % p.S;
            const int wt = p.flip ? ((p.R - 1 - r) * p.S + (p.S - 1 - s)) : tap;
            for (int j = 0; j < p.BN / 64; ++j) tma_load_2d(&tmB, &full_w, dst + j * 8192, wt * p.w_cols + n0 + j * 64, cb * BK);
          }
        }
      int ia = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int t = tile;
        const int tw = t % p.tiles_w; t /= p.tiles_w;
        const int th = t % p.tiles_h;
        const int img = t / p.tiles_h;
        for (int cb = 0; cb < cblocks; ++cb, ++ia) {
          const int sa = ia % p.a_stages;
          mbar_wait(&empty_a[sa], ((ia / p.a_stages) & 1) ^ 1);
          mbar_expect_tx(&full_a[sa], (uint32_t)p.a_box_bytes);
          tma_load_4d(&tmA, &full_a[sa], smem + (size_t)sa * p.a_stage_bytes, cb * BK, tw * p.TW - p.pw, th * p.TR - p.ph, img);
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    const uint32_t b_lbo = p.b_mn ? 8192u : 16u, b_step = p.b_mn ? 128u : 2u;
    mbar_wait(&full_w, 0);
    int ia = 0, it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it % n_acc;
      mbar_wait(&tmem_empty[acc], ((it / n_acc) & 1) ^ 1);   // the epilogue drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)acc * acc_cols;
      for (int cb = 0; cb < cblocks; ++cb, ++ia) {
        const int sa = ia % p.a_stages;
        mbar_wait(&full_a[sa], (ia / p.a_stages) & 1);
        tc_fence_after();
        if (elect_one()) {
          // The issuing thread's own instruction stream is the limit for short (N = 64) MMAs, so descriptors are built
          // once and only their 32-bit low words (start address >> 4) are advanced: +8 per pixel row for A taps,
          // +b_stage_bytes/16 per resident weight tile, +2 (K-major) or +128 (MN-major) per K = 16 step.
          const uint64_t a0 = make_desc(smem_u32(smem + (size_t)sa * p.a_stage_bytes), 16u, 1024);
          const uint64_t b0 = make_desc(smem_u32(smem_w + (size_t)cb * taps * p.b_stage_bytes), b_lbo, 1024);
          const uint32_t b_tile = (uint32_t)p.b_stage_bytes >> 4;
          uint32_t a_off = 0, b_off = 0;
          int s = 0;
          for (int tap = 0; tap < taps; ++tap) {
            const uint64_t adesc = a0 + a_off, bdesc = b0 + b_off;
            umma_bf16(d_tmem, adesc, bdesc, p.idesc, (cb | tap) != 0);
            umma_bf16(d_tmem, adesc + 2u, bdesc + b_step, p.idesc, 1u);
            umma_bf16(d_tmem, adesc + 4u, bdesc + 2u * b_step, p.idesc, 1u);
            umma_bf16(d_tmem, adesc + 6u, bdesc + 3u * b_step, p.idesc, 1u);
            b_off += b_tile;
            if (++s == p.S) { s = 0; a_off += (uint32_t)(p.pitch - p.S + 1) * 8u; } else { a_off += 8u; }
          }
          umma_commit(&empty_a[sa]);
          if (cb == cblocks - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
      }
    }
  } else {
    // ================================ epilogue ================================
    const int quarter = warp & 3;
    // the epilogue is the longer side of the tile pipeline on 64-channel layers (elimination experiments):
    // two warps share each lane quarter and split the columns
    const int c_lo = ((warp - 2) >> 2) * (p.BN / 2), c_hi = c_lo + p.BN / 2;
    const int m = quarter * 32 + lane;
    const int tr = m / p.pitch, j = m % p.pitch;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it % n_acc;
      int t = tile;
      const int tw = t % p.tiles_w; t /= p.tiles_w;
      const int th = t % p.tiles_h;
      const int img = t / p.tiles_h;
      const int p0 = th * p.TR, j0 = tw * p.TW;
      const bool row_ok = tr < p.TR && j < p.TW && (p0 + tr) < p.Ho && (j0 + j) < p.Wo;
      const long long pix = ((long long)img * p.Ho + p0 + tr) * p.Wo + j0 + j;
      mbar_wait(&tmem_full[acc], (it / n_acc) & 1);
      tc_fence_after();
      for (int c = c_lo; c < c_hi; c += 32) {
        uint32_t rr[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)acc * acc_cols + (uint32_t)c, rr);
        float v[32];
#pragma unroll
        for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(rr[q]);
        if (c + 32 >= c_hi) {                      // last chunk read: hand the accumulator back to the MMA warp early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty[acc])) : "memory");
        }
        if (p.bias) {
#pragma unroll
          for (int q = 0; q < 32; ++q) if (n0 + c + q < p.Co) v[q] += p.bias[n0 + c + q];
        }
        if (p.relu) {
#pragma unroll
          for (int q = 0; q < 32; ++q) v[q] = fmaxf(v[q], 0.f);
        }
        if (p.stats) {
          float s1[32], s2[32];
#pragma unroll
          for (int q = 0; q < 32; ++q) { const float x = row_ok ? v[q] : 0.f; s1[q] = x; s2[q] = x * x; }
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int q = 0; q < off; ++q) {
              const float send1 = upper ? s1[q] : s1[q + off], send2 = upper ? s2[q] : s2[q + off];
              const float keep1 = upper ? s1[q + off] : s1[q], keep2 = upper ? s2[q + off] : s2[q];
              s1[q] = keep1 + __shfl_xor_sync(0xffffffffu, send1, off);
              s2[q] = keep2 + __shfl_xor_sync(0xffffffffu, send2, off);
            }
          }
          atomicAdd(&stat_smem[0][c + lane], s1[0]);
          atomicAdd(&stat_smem[1][c + lane], s2[0]);
        }
        if (row_ok) {
          const int ncols = min(32, p.Co - (n0 + c));
          if (ncols > 0) {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + pix * p.ldd + n0 + c;
            if (ncols == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
              for (int q = 0; q < 32; q += 8) {
                uint4 u;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
                h[0] = __floats2bfloat162_rn(v[q], v[q + 1]); h[1] = __floats2bfloat162_rn(v[q + 2], v[q + 3]);
                h[2] = __floats2bfloat162_rn(v[q + 4], v[q + 5]); h[3] = __floats2bfloat162_rn(v[q + 6], v[q + 7]);
                *reinterpret_cast<uint4*>(dst + q) = u;
              }
            } else {
              for (int q = 0; q < ncols; ++q) dst[q] = __float2bfloat16_rn(v[q]);
            }
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (p.stats) {
    for (int q = threadIdx.x; q < p.BN; q += blockDim.x) {
      if (n0 + q < p.Co) {
        atomicAdd(p.stats + n0 + q, stat_smem[0][q]);
        atomicAdd(p.stats + p.Co + n0 + q, stat_smem[1][q]);
      }
    }
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// Tile geometry: TR rows x TW columns of output with TR * (TW + S - 1) <= 128 GEMM rows.  Among the splits of a row
// into 1..8 column parts pick the one that fills the 128 rows best, discounted by how much halo the tile drags in
// ((TR + R - 1) * pitch pixels loaded per TR * TW outputs): narrow-and-tall tiles re-read fewer halo rows.
void pick_tile(int Ho, int Wo, int R, int S, int* TR, int* TW) {
  const int forced = env_int("MPB_HALO_PARTS", 0);
  double best = -1.0;
  for (int parts = 1; parts <= 8; ++parts) {
    if (forced && parts != forced) continue;
    const int tw = (Wo + parts - 1) / parts;
    const int pitch = tw + S - 1;
    if (pitch > 128) continue;
    int tr = 128 / pitch;
    if (tr > Ho) tr = Ho;
    if (tr < 1) continue;
    const int tiles = ((Ho + tr - 1) / tr) * ((Wo + tw - 1) / tw);
    const double eff = (double)Ho * Wo / ((double)tiles * 128.0);
    const double halo = (double)(tr + R - 1) * pitch / ((double)tr * tw);
    const double score = eff / (1.0 + 0.15 * (halo - 1.0));
    if (score > best * 1.0001) { best = score; *TR = tr; *TW = tw; }
    if (tw == 1) break;
  }
  if (best < 0) { *TW = 128 - (S - 1); *TR = 1; }
}

// Can the persistent, weights-resident variant run this layer, and with which tile width?  Prefers BN = 128 and
// falls back to two 64-wide passes when only half of the filter bank fits (activation tiles are then fetched twice,
// which is still far less traffic than streaming weights per tap).
bool plan_persistent(int Ci, int Co, int R, int S, int Ho, int Wo, int* BN_out, int* a_stages_out) {
  int TR, TW;
  pick_tile(Ho, Wo, R, S, &TR, &TW);
  const int pitch = TW + S - 1, box_h = TR + R - 1;
  if (pitch > 256 || box_h > 256) return false;
  const int cblocks = (Ci + BK - 1) / BK;
  if (cblocks > 2) return false;
  const long long a_stage = ((long long)box_h * pitch * 128 + 1023) / 1024 * 1024;
  const int first = (Co > 64) ? 128 : 64;
  for (int BN = first; BN >= 64; BN -= 64) {
    const long long wbytes = (long long)cblocks * R * S * BN * BK * 2;
    const long long room = 208 * 1024 - 2048 - wbytes;
    if (room >= 2 * a_stage) {
      int st = (int)(room / a_stage);
      *BN_out = BN; *a_stages_out = st > 4 ? 4 : st;
      return true;
    }
  }
  return false;
}

int launch_halo(const void* in, const void* w, void* out, const float* bias, float* stats, int N, int Hi, int Wi, int Ci,
                int Ho, int Wo, int Co, int R, int S, int ph, int pw, int b_mn, int flip, int w_rows, long long w_cols_total,
                int w_cols, cudaStream_t st, long long xs_w = 0, long long xs_h = 0, long long xs_n = 0, int relu = 0) {
  int rc = load_driver_fns();
  if (rc) return rc;
  HaloParams p;
  memset(&p, 0, sizeof(p));
  p.N = N; p.Ho = Ho; p.Wo = Wo; p.Co = Co; p.Ci = Ci; p.R = R; p.S = S; p.ph = ph; p.pw = pw;
  pick_tile(Ho, Wo, R, S, &p.TR, &p.TW);
  p.pitch = p.TW + S - 1;
  p.tiles_h = (Ho + p.TR - 1) / p.TR; p.tiles_w = (Wo + p.TW - 1) / p.TW;
  p.BN = (Co > 64) ? 128 : 64; p.b_mn = b_mn; p.flip = flip; p.w_cols = w_cols;
  int bn_p = 0, a_st_p = 0;
  if (!plan_persistent(Ci, Co, R, S, Ho, Wo, &bn_p, &a_st_p)) return -2;    // callers check mpb_tc_halo_persistent_ok
  p.BN = bn_p;
  p.idesc = make_idesc(p.BN, 0, b_mn);
  const int box_h = p.TR + R - 1;
  if (p.pitch > 256 || box_h > 256) return -2;
  p.a_box_bytes = box_h * p.pitch * 128;
  p.a_stage_bytes = (p.a_box_bytes + 1023) / 1024 * 1024;
  p.b_stage_bytes = p.BN * BK * 2;
  p.D = out; p.ldd = Co; p.bias = bias; p.stats = stats; p.relu = relu;
  CUtensorMap ta, tb;
  rc = make_map_nhwc_tiled(&ta, in, N, Hi, Wi, Ci, p.pitch, box_h, xs_w, xs_h, xs_n);
  if (rc) return rc;
  rc = b_mn ? make_map_2d(&tb, w, w_rows, w_cols_total, w_cols_total, 64, 64)
            : make_map_2d(&tb, w, w_rows, w_cols_total, w_cols_total, 64, p.BN);
  if (rc) return rc - 1000;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(tc_halo_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int cblocks = (Ci + BK - 1) / BK;
  const long long wbytes = (long long)cblocks * R * S * p.b_stage_bytes;
  const int n_tiles_y = (Co + p.BN - 1) / p.BN;
  p.a_stages = a_st_p;
  const int smem_p = p.a_stages * p.a_stage_bytes + (int)wbytes + 2048;
  const int num_tiles = N * p.tiles_h * p.tiles_w;
  int ctas = kSMs / n_tiles_y;
  if (ctas < 1) ctas = 1;
  if (ctas > num_tiles) ctas = num_tiles;
  dim3 gridp((unsigned)ctas, (unsigned)n_tiles_y, 1);
  tc_halo_persist_kernel<<<gridp, kThreadsPersist, smem_p, st>>>(ta, tb, p, num_tiles);
  return (int)cudaGetLastError();
}

}  // namespace

// 1 when the persistent, weights-resident variant applies (the CTA's whole filter bank fits in shared memory beside
// a >= 2-deep halo ring); that is the variant that beats the im2col kernel, the streaming one does not (measured).
MPB_API int mpb_tc_halo_persistent_ok(int Ci, int Co, int R, int S, int Ho, int Wo) {
  int bn = 0, st = 0;
  return plan_persistent(Ci, Co, R, S, Ho, Wo, &bn, &st) ? 1 : 0;
}

// Fraction of the 128 GEMM rows that are real output pixels for this geometry (the caller's go / no-go signal).
MPB_API float mpb_tc_halo_efficiency(int Ho, int Wo, int R, int S) {
  int TR, TW;
  pick_tile(Ho, Wo, R, S, &TR, &TW);
  const int th = (Ho + TR - 1) / TR, tw = (Wo + TW - 1) / TW;
  return (float)((double)Ho * Wo / ((double)th * tw * 128.0));
}

// y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R,S,C]), stride 1, dilation 1 (P = H + 2*pad_h - R + 1).  stats as in mpb_tc_conv_fprop.
MPB_API int mpb_tc_conv_fprop_halo_act(const void* x, const void* w, void* y, const float* bias, float* stats, int N,
                                       int H, int W, int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w,
                                       int relu, void* stream) {
  if (C % 8 != 0 || K % 8 != 0) return -2;
  return launch_halo(x, w, y, bias, stats, N, H, W, C, P, Q, K, R, S, pad_h, pad_w, /*b_mn=*/0, /*flip=*/0,
                     /*w_rows=*/K, /*w_cols_total=*/(long long)R * S * C, /*w_cols=*/0, (cudaStream_t)stream, 0, 0, 0, relu);
}
MPB_API int mpb_tc_conv_fprop_halo(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H,
                                   int W, int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w, void* stream) {
  return mpb_tc_conv_fprop_halo_act(x, w, y, bias, stats, N, H, W, C, K, R, S, P, Q, pad_h, pad_w, 0, stream);
}

// Same, with explicit element strides of x between pixels / rows / images.  A pixel stride smaller than C describes
// overlapping super-pixels: the space-to-depth form of the 7x7/2 stem reads 4 neighbouring 16-channel pixels as one
// 64-channel pixel that advances by 16 elements (mpb_s2d_stem below builds that buffer).
MPB_API int mpb_tc_conv_fprop_halo_strided(const void* x, const void* w, void* y, const float* bias, float* stats, int N,
                                           int H, int W, int C, int K, int R, int S, int P, int Q, int pad_h, int pad_w,
                                           long long xs_w, long long xs_h, long long xs_n, void* stream) {
  if (C % 8 != 0 || K % 8 != 0) return -2;
  return launch_halo(x, w, y, bias, stats, N, H, W, C, P, Q, K, R, S, pad_h, pad_w, 0, 0, K, (long long)R * S * C, 0,
                     (cudaStream_t)stream, xs_w, xs_h, xs_n);
}

// dx[N,H,W,C] for a stride-1 convolution: a stride-1 convolution of dy with the flipped filter, pad' = R-1-pad.
MPB_API int mpb_tc_conv_dgrad_halo(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R,
                                   int S, int P, int Q, int pad_h, int pad_w, void* stream) {
  if (C % 8 != 0 || K % 8 != 0) return -2;
  return launch_halo(dy, w, dx, nullptr, nullptr, N, P, Q, K, H, W, C, R, S, (R - 1) - pad_h, (S - 1) - pad_w, /*b_mn=*/1,
                     /*flip=*/1, /*w_rows=*/K, /*w_cols_total=*/(long long)R * S * C, /*w_cols=*/C, (cudaStream_t)stream);
}
