// K3-K6: tensor-core GEMM and implicit-GEMM convolution for sm_100a, hand-written with
// tcgen05.mma (accumulators in TMEM), TMA (cp.async.bulk.tensor, tiled and im2col modes) and mbarrier
// pipelines.  Replaces the ATen CPU conv / addmm calls the reference reaches from every nn.Conv2d and
// nn.Linear: fprop, dgrad, wgrad and the linear fwd/dX/dW GEMMs, bf16 in, fp32
// accumulate.  No cuDNN / cuBLAS / CUTLASS on this path.
//
// One CTA computes one 128 x BN output tile (BN in {32..256}).  Warp roles (192 threads):
//   warp 0   : TMA producer  -- fills a kStages-deep ring of {A tile 128x64, B tile BNx64} bf16 in
//              128B-swizzled shared memory; completion via mbarrier complete_tx
//   warp 1   : TMEM allocator + MMA issuer -- one elected lane issues 4 x tcgen05.mma (K=16) per stage,
//              tcgen05.commit releases the smem stage and finally signals the epilogue
//   warps 2-5: epilogue -- tcgen05.ld the fp32 accumulator (lane = output row), add bias, optionally
//              accumulate per-channel sum / sum-of-squares for BatchNorm, convert, store (or fp32
//              atomic-add for split-K wgrad)
//
// The same kernel serves every problem; only the producer's coordinate generation changes:
//   GEMM      : A[M,K], B[N,K] each either K-major ([rows][K]) or MN-major ([K][rows])
//   conv fprop: A = im2col(x)  via TMA im2col mode (NHWC, C%64==0), B = W[Cout][R*S*C] K-major
//   conv dgrad: A = im2col(dy) (stride-1 convs; flipped taps), B = W read as MN-major (no transpose copy)
//   conv wgrad: A = dy^T MN-major, B = im2col(x) MN-major, reduction over output pixels, split-K
//
// UMMA shared-memory descriptor (64-bit), SWIZZLE_128B, bf16:
//   K-major  tile [rows][64]: rows of 128 B; 8-row groups 1024 B apart (SBO=1024); +32 B per K=16 step
//   MN-major tile [64 k][64 mn] regions: rows (k) of 128 B; 8-k groups 1024 B apart (SBO=1024);
//            64-wide mn regions LBO bytes apart (LBO = 8192); +2048 B per K=16 step
#include "tc_ptx.cuh"

using namespace mpb;
using namespace mpbtc;

namespace {

constexpr int kThreads = 192;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB

enum Mode { MODE_GEMM = 0, MODE_FPROP = 1, MODE_DGRAD = 2, MODE_WGRAD = 3, MODE_DGRAD_STRIDED = 4 };
enum OutMode { OUT_BF16 = 0, OUT_F32 = 1, OUT_ATOMIC_F32 = 2 };

struct TcParams {
  int mode;
  int M, N;                 // output tile space (rows, cols)
  int BN;                   // tile width (multiple of 16, <= 256)
  int a_mn, b_mn;           // operand majorness
  int num_kb;               // total K blocks (of 64)
  int kb_per_split;         // K blocks per blockIdx.z
  uint32_t idesc;
  int out_mode;
  void* D; long long ldd;   // output, row stride in elements
  const float* bias;        // [N] or null
  float* stats;             // [2*N] per-channel sum / sumsq of the fp32 accumulator, or null
  // conv geometry (im2col coordinate generation)
  int C, R, S;              // reduction channels per tap (fprop: Cin, dgrad: Cout), filter size
  int Hout, Wout;           // spatial size of the GEMM-M pixel space (fprop: P,Q; dgrad: H,W; wgrad: P,Q)
  int stride, pad_h, pad_w, dil;
  int tap_cols;             // wgrad: grid.y enumerates (tap, c-block); number of c-blocks per tap
  int Cin_total;            // wgrad: Cin (row stride pieces of dW)
  int stages;               // smem ring depth (fewer stages -> more co-resident CTAs per SM)
  // strided dgrad, one launch per output parity class (ph, pw): valid filter taps r = r0 + st*i, s = s0 + st*j
  int st, ph, pw, r0, s0, nr, ns, lower_h, lower_w, Hfull, Wfull;
  int relu;                 // epilogue: max(x, 0) after the bias (inference: conv + folded BatchNorm + ReLU in one kernel)
  int nclass;               // strided dgrad as ONE launch: blockIdx.z = output parity class, geometry in cls[z]
  struct StridedClass { int M, Hout, Wout, num_kb, ph, pw, r0, s0, nr, ns, lower_h, lower_w; } cls[4];
  int cluster;              // 1, or 2: CTA pair (cta_group::2) computes a 256 x BN tile; each CTA stages its 128 A rows and HALF of B
};

struct ExtraMaps { CUtensorMap m[3]; };   // im2col maps of parity classes 1..3 (class 0 uses tmA)

// kTwo = CTA-pair (cta_group::2) build.  It is a separate instantiation because a kernel that contains
// cta_group::2 instructions can only be launched with an even cluster size (plain launches fail with error 912).
template <bool kTwo>
__global__ void __launch_bounds__(kThreads, 4)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ ExtraMaps tmX, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages][A 16KB][B BN*128B] | barriers
  const int b_stage_bytes = (p.BN / p.cluster) * BK * 2;   // a CTA pair splits the B tile between its two shared memories
  constexpr int a_bytes = A_STAGE_BYTES;
  const int stage_bytes = a_bytes + b_stage_bytes;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full_bar[8], empty_bar[8], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ float stat_smem[2][256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stages = p.stages;
  const int m0 = blockIdx.x * BM;
  int n0 = blockIdx.y * p.BN;
  int tap_fixed = 0, cblk_fixed = 0;
  if (p.mode == MODE_WGRAD) {           // grid.y = (tap, c-block)
    tap_fixed = blockIdx.y / p.tap_cols;
    cblk_fixed = blockIdx.y % p.tap_cols;
    n0 = 0;
  }
  // geometry: the launch's own, or (single-launch strided dgrad) that of the parity class this CTA works on
  const bool multi = p.mode == MODE_DGRAD_STRIDED && p.nclass > 1;
  const int cz = multi ? (int)blockIdx.z : 0;
  const CUtensorMap* mapA = (multi && cz > 0) ? &tmX.m[cz - 1] : &tmA;
  int gM = p.M, gH = p.Hout, gW = p.Wout, g_ph = p.ph, g_pw = p.pw, g_r0 = p.r0, g_s0 = p.s0, g_nr = p.nr, g_ns = p.ns;
  int g_lower_h = p.lower_h, g_lower_w = p.lower_w, g_num_kb = p.num_kb, g_kbps = p.kb_per_split;
  if (multi) {
    const TcParams::StridedClass& c = p.cls[cz];
    gM = c.M; gH = c.Hout; gW = c.Wout; g_ph = c.ph; g_pw = c.pw; g_r0 = c.r0; g_s0 = c.s0; g_nr = c.nr; g_ns = c.ns;
    g_lower_h = c.lower_h; g_lower_w = c.lower_w; g_num_kb = c.num_kb; g_kbps = c.num_kb;
    // a class with fewer rows than the grid covers: the whole CTA (pair) leaves before touching barriers or TMEM
    const int lead_m0 = (kTwo ? (int)(blockIdx.x & ~1u) : (int)blockIdx.x) * BM;
    if (lead_m0 >= gM) return;
  }
  const int kb_begin = multi ? 0 : blockIdx.z * g_kbps;
  const int kb_end = min(g_num_kb, kb_begin + g_kbps);
  const int nkb = kb_end - kb_begin;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < p.BN) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    prefetch_tmap(mapA);
    prefetch_tmap(&tmB);
    for (int i = 0; i < stages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    mbar_init(&tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (p.stats) for (int j = threadIdx.x; j < 256; j += blockDim.x) { stat_smem[0][j] = 0.f; stat_smem[1][j] = 0.f; }
  constexpr bool two = kTwo;
  if (warp == 1) { if constexpr (two) tmem_alloc_2sm(&tmem_base_smem, tmem_cols); else tmem_alloc(&tmem_base_smem, tmem_cols); }
  tc_fence_before();
  __syncthreads();
  if constexpr (two) cluster_sync_all();     // both CTAs' barriers and TMEM exist before the pair starts signalling
  tc_fence_after();
  uint32_t crank = 0u;
  if constexpr (two) crank = cluster_ctarank();
  const uint32_t tmem_base = tmem_base_smem;

  if (nkb > 0) {
    if (warp == 0) {
      // ================================ TMA producer ================================
      if (elect_one()) {
        // pixel decomposition of the tile's first row (fprop / dgrad)
        int q0 = 0, p0 = 0, img0 = 0;
        if (p.mode == MODE_FPROP || p.mode == MODE_DGRAD || p.mode == MODE_DGRAD_STRIDED) {
          q0 = m0 % gW;
          const int t = m0 / gW;
          p0 = t % gH;
          img0 = t / gH;
        }
        const int cblocks = (p.C + BK - 1) / BK;   // channel tail: TMA zero-fills c >= C, so the overhang multiplies by 0
        for (int i = 0; i < nkb; ++i) {
          const int kb = kb_begin + i;
          const int st = i % stages;
          const uint32_t ph = (i / stages) & 1;
          mbar_wait(&empty_bar[st], ph ^ 1);
          uint8_t* sa = smem + (size_t)st * stage_bytes;
          uint8_t* sb = sa + a_bytes;
          // pair mode: every byte of both CTAs completes on the leader's barrier, which the leader arms for both
          if constexpr (!two) mbar_expect_tx(&full_bar[st], (uint32_t)stage_bytes);
          else { if (crank == 0) mbar_expect_tx(&full_bar[st], (uint32_t)(2 * stage_bytes)); }
          uint64_t* fb = &full_bar[st];
          auto LD2 = [&](const CUtensorMap* m, void* dst, int c0, int c1) {
            if constexpr (two) tma_load_2d_2sm(m, fb, dst, c0, c1); else tma_load_2d(m, fb, dst, c0, c1);
          };
          auto LDI = [&](const CUtensorMap* m, void* dst, int c, int w, int h, int n, int ow, int oh) {
            if constexpr (two) tma_load_im2col_4d_2sm(m, fb, dst, c, w, h, n, (uint16_t)ow, (uint16_t)oh);
            else tma_load_im2col_4d(m, fb, dst, c, w, h, n, (uint16_t)ow, (uint16_t)oh);
          };
          // this CTA's share of the B tile: all BN columns, or the crank-th half of them in pair mode
          const int bn_cta = p.BN / p.cluster, n_off = (int)crank * bn_cta, regions = bn_cta / 64;
          if (p.mode == MODE_GEMM) {
            if (!p.a_mn) LD2(mapA, sa, kb * BK, m0);
            else { LD2(mapA, sa, m0, kb * BK); LD2(mapA, sa + 8192, m0 + 64, kb * BK); }
            if (!p.b_mn) LD2(&tmB, sb, kb * BK, n0 + n_off);
            else for (int j = 0; j < regions; ++j) LD2(&tmB, sb + j * 8192, n0 + n_off + j * 64, kb * BK);
          } else if (p.mode == MODE_FPROP) {
            const int tap = kb / cblocks, cb = kb % cblocks;
            const int r = tap / p.S, s = tap % p.S;
            LDI(mapA, sa, cb * BK, q0 * p.stride - p.pad_w, p0 * p.stride - p.pad_h, img0, s * p.dil, r * p.dil);
            LD2(&tmB, sb, tap * p.C + cb * BK, n0 + n_off);
          } else if (p.mode == MODE_DGRAD) {
            // A = im2col(dy) with pad' = (R-1) - pad, taps (r'', s''); B = W[k][r][s][c] with r = R-1-r'', MN-major
            const int tap = kb / cblocks, cb = kb % cblocks;       // cblocks = ceil(Cout/64) here
            const int rr = tap / p.S, ss = tap % p.S;
            LDI(mapA, sa, cb * BK, q0 - p.pad_w, p0 - p.pad_h, img0, ss, rr);
            const int r = p.R - 1 - rr, s = p.S - 1 - ss;
            const int col = (r * p.S + s) * p.Cin_total + n0 + n_off;
            for (int j = 0; j < regions; ++j) LD2(&tmB, sb + j * 8192, col + j * 64, cb * BK);
          } else if (p.mode == MODE_DGRAD_STRIDED) {
            // rows = (n, h', w') of one parity class; dy row = h' + lower_h + (nr-1-i) for filter row r = r0 + st*i
            const int tap = kb / cblocks, cb = kb % cblocks;
            const int ir = tap / g_ns, is = tap % g_ns;
            LDI(mapA, sa, cb * BK, q0 + g_lower_w, p0 + g_lower_h, img0, g_ns - 1 - is, g_nr - 1 - ir);
            const int r = g_r0 + p.st * ir, s = g_s0 + p.st * is;
            const int col = (r * p.S + s) * p.Cin_total + n0 + n_off;
            for (int j = 0; j < regions; ++j) LD2(&tmB, sb + j * 8192, col + j * 64, cb * BK);
          } else {  // MODE_WGRAD: reduction over output pixels; kb = pixel chunk of 64
            const int pix = kb * BK;
            const int q = pix % p.Wout;
            const int t = pix / p.Wout;
            const int pp = t % p.Hout, img = t / p.Hout;
            LD2(mapA, sa, m0, pix);
            LD2(mapA, sa + 8192, m0 + 64, pix);
            const int r = tap_fixed / p.S, s = tap_fixed % p.S;
            for (int j = 0; j < regions; ++j)
              LDI(&tmB, sb + j * 8192, cblk_fixed * p.BN + n_off + j * 64, q * p.stride - p.pad_w, pp * p.stride - p.pad_h,
                  img, s * p.dil, r * p.dil);
          }
        }
      }
    } else if (warp == 1) {
      // ================================ MMA issuer ================================
      const uint32_t a_lbo = p.a_mn ? 8192u : 16u, b_lbo = p.b_mn ? 8192u : 16u;
      const uint32_t a_step = p.a_mn ? 128u : 2u, b_step = p.b_mn ? 128u : 2u;   // descriptor advance per K=16 (16 B units)
      for (int i = 0; i < nkb && crank == 0; ++i) {      // in pair mode only the leader CTA issues (for both)
        const int st = i % stages;
        const uint32_t ph = (i / stages) & 1;
        mbar_wait(&full_bar[st], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + (size_t)st * stage_bytes);
          const uint32_t sb = sa + a_bytes;
          const uint64_t adesc = make_desc(sa, a_lbo, 1024);
          const uint64_t bdesc = make_desc(sb, b_lbo, 1024);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if constexpr (two) umma_bf16_2sm(tmem_base, adesc + (uint64_t)(a_step * k), bdesc + (uint64_t)(b_step * k), p.idesc, (i | k) != 0);
            else umma_bf16(tmem_base, adesc + (uint64_t)(a_step * k), bdesc + (uint64_t)(b_step * k), p.idesc, (i | k) != 0);
          }
          if constexpr (!two) {
            umma_commit(&empty_bar[st]);               // frees the smem stage once these MMAs retire
            if (i == nkb - 1) umma_commit(&tmem_full_bar);
          } else {                                     // ... in both CTAs of the pair
            umma_commit_2sm(&empty_bar[st], 3);
            if (i == nkb - 1) umma_commit_2sm(&tmem_full_bar, 3);
          }
        }
        __syncwarp();
      }
    } else {
      // ================================ epilogue ================================
      mbar_wait(&tmem_full_bar, 0);
      tc_fence_after();
      const int quarter = warp & 3;                      // TMEM lane quarter this warp may access
      const int row = m0 + quarter * 32 + lane;
      const bool row_ok = row < gM;
      int col_base = n0;
      long long out_off = 0;
      long long out_row = row;
      if (p.mode == MODE_DGRAD_STRIDED && row_ok) {      // scatter to (n, st*h'+ph, st*w'+pw) of the full-resolution dx
        const int wq = row % gW, t = row / gW;
        const int hq = t % gH, img = t / gH;
        out_row = ((long long)img * p.Hfull + (long long)p.st * hq + g_ph) * p.Wfull + (long long)p.st * wq + g_pw;
      }
      if (p.mode == MODE_WGRAD) { col_base = cblk_fixed * p.BN; out_off = (long long)tap_fixed * p.Cin_total; }
      for (int c = 0; c < p.BN; c += 32) {
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c, r);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (p.bias) {
#pragma unroll
          for (int j = 0; j < 32; ++j) if (col_base + c + j < p.N) v[j] += p.bias[col_base + c + j];
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (p.stats) {
          // per-column sums over this warp's 32 rows: transpose-reduce (31 shuffles per quantity)
          float s1[32], s2[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) { const float x = row_ok ? v[j] : 0.f; s1[j] = x; s2[j] = x * x; }
#pragma unroll
          for (int off = 16; off >= 1; off >>= 1) {
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int j = 0; j < off; ++j) {
              // lane keeps columns [j + (upper?off:0)], sends the other half
              const float send1 = upper ? s1[j] : s1[j + off], send2 = upper ? s2[j] : s2[j + off];
              const float keep1 = upper ? s1[j + off] : s1[j], keep2 = upper ? s2[j + off] : s2[j];
              s1[j] = keep1 + __shfl_xor_sync(0xffffffffu, send1, off);
              s2[j] = keep2 + __shfl_xor_sync(0xffffffffu, send2, off);
            }
          }
          // lane L now holds the total of column bitrev-free index: column = L (see derivation in DESIGN.md)
          atomicAdd(&stat_smem[0][c + lane], s1[0]);
          atomicAdd(&stat_smem[1][c + lane], s2[0]);
        }
        if (row_ok) {
          const int ncols = min(32, p.N - (col_base + c));
          if (ncols > 0) {
            if (p.out_mode == OUT_BF16) {
              __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + out_row * p.ldd + out_off + col_base + c;
              if (ncols == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                  uint4 u;
                  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
                  h[0] = __floats2bfloat162_rn(v[j], v[j + 1]); h[1] = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
                  h[2] = __floats2bfloat162_rn(v[j + 4], v[j + 5]); h[3] = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
                  *reinterpret_cast<uint4*>(dst + j) = u;
                }
              } else {
                for (int j = 0; j < ncols; ++j) dst[j] = __float2bfloat16_rn(v[j]);
              }
            } else if (p.out_mode == OUT_F32) {
              float* dst = reinterpret_cast<float*>(p.D) + out_row * p.ldd + out_off + col_base + c;
              if (ncols == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
              } else {
                for (int j = 0; j < ncols; ++j) dst[j] = v[j];
              }
            } else {
              float* dst = reinterpret_cast<float*>(p.D) + out_row * p.ldd + out_off + col_base + c;
              if (ncols == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 4)       // 16-byte vector reduction: a quarter of the RED instructions
                  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(v[j]), "f"(v[j + 1]),
                               "f"(v[j + 2]), "f"(v[j + 3]) : "memory");
              } else {
                for (int j = 0; j < ncols; ++j) atomicAdd(dst + j, v[j]);
              }
            }
          }
        }
      }
      tc_fence_before();
    }
  } else if (multi && warp >= 2) {
    // a parity class no filter tap reaches (e.g. three of the four classes of a 1x1 stride-2 convolution): its input
    // gradient is exactly zero, written here instead of by a memset over the whole tensor
    const int row = m0 + (warp & 3) * 32 + lane;
    if (row < gM) {
      const int wq = row % gW, t = row / gW;
      const int hq = t % gH, img = t / gH;
      const long long out_row = ((long long)img * p.Hfull + (long long)p.st * hq + g_ph) * p.Wfull + (long long)p.st * wq + g_pw;
      __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + out_row * p.ldd + n0;
      const int ncols = min(p.BN, p.N - n0);
      if ((ncols & 7) == 0 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
        for (int j = 0; j < ncols; j += 8) *reinterpret_cast<uint4*>(dst + j) = make_uint4(0u, 0u, 0u, 0u);
      } else {
        for (int j = 0; j < ncols; ++j) dst[j] = __float2bfloat16_rn(0.f);
      }
    }
  }
  __syncthreads();
  if (p.stats && nkb > 0) {
    for (int j = threadIdx.x; j < p.BN; j += blockDim.x) {
      if (n0 + j < p.N) {
        atomicAdd(p.stats + n0 + j, stat_smem[0][j]);
        atomicAdd(p.stats + p.N + n0 + j, stat_smem[1][j]);
      }
    }
  }
  if constexpr (two) cluster_sync_all();     // the leader may still read this CTA's smem / signal its barriers until it is done
  if (warp == 1) {
    tc_fence_after();
    if constexpr (two) tmem_dealloc_2sm(tmem_base, tmem_cols); else tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------ host side
int default_stages(int BN, int num_kb) {
  int s = env_int("MPB_TC_STAGES", 0);
  if (s <= 0) s = 2;   // measured: co-resident CTAs (3-4 per SM) beat a deeper ring on every ResNet layer
  if (s > 8) s = 8;
  if (s > num_kb && num_kb > 0) s = num_kb < 2 ? 2 : num_kb;
  while (s > 2 && s * (A_STAGE_BYTES + BN * BK * 2) + 1024 > 200 * 1024) --s;
  return s;
}

// CTA pairs need >= 2 M tiles, and a B share per CTA that is a whole number of 64-wide regions when B is MN-major
int want_cluster(int m_tiles, int BN, int b_mn) {
  if (env_int("MPB_TC_CLUSTER", 1) == 0 || m_tiles < 2) return 1;
  if (b_mn ? (BN % 128 != 0) : (BN % 16 != 0)) return 1;
  return 2;
}

int launch(const CUtensorMap& a, const CUtensorMap& b, TcParams p, dim3 grid, cudaStream_t st,
           const ExtraMaps* extra = nullptr) {
  static const ExtraMaps no_extra = {};
  const ExtraMaps& x = extra ? *extra : no_extra;
  if (p.cluster < 1) p.cluster = 1;
  if (p.cluster == 2) p.idesc = (p.idesc & ~(0x1Fu << 24)) | ((uint32_t)(2 * BM >> 4) << 24);   // UMMA M = 256 across the pair
  p.stages = default_stages(p.BN / p.cluster, p.kb_per_split);
  const int smem = p.stages * (A_STAGE_BYTES + (p.BN / p.cluster) * BK * 2) + 1024;
  static int configured = 0;
  if (configured < smem) {
    cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 1024);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(tc_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 1024);
    if (e != cudaSuccess) return (int)e;
    configured = 200 * 1024 + 1024;
  }
  if (p.cluster == 1) {
    tc_gemm_kernel<false><<<grid, kThreads, smem, st>>>(a, b, x, p);
    return (int)cudaGetLastError();
  }
  grid.x = (grid.x + 1) / 2 * 2;              // CTA pairs along M; a padding CTA only sees out-of-range rows
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = (size_t)smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, tc_gemm_kernel<true>, a, b, x, p);
}

int pick_bn(int n) {
  const int forced = env_int("MPB_TC_BN", 0);
  if (forced > 0 && n % forced == 0) return forced;
  if (n % 128 == 0 || n > 192) return 128;
  if (n > 64) return 128;
  if (n > 32) return 64;
  return 32;
}

}  // namespace

// D[M,N] = A[M,K] * B[N,K]^T   (bf16 operands, fp32 accumulate)
//   a_mn = 0: A stored [M][K] (row stride lda)      a_mn = 1: A stored [K][M] (row stride lda)
//   b_mn = 0: B stored [N][K] (row stride ldb)      b_mn = 1: B stored [K][N] (row stride ldb)
//   out_mode 0: bf16 D, 1: fp32 D, 2: fp32 atomic accumulate with split-K (D must be pre-zeroed)
//   stats (optional, no split-K): 2*N floats zeroed by the caller, receive per-column sum / sum of squares of D
MPB_API int mpb_tc_gemm(const void* A, const void* B, void* D, const float* bias, float* stats, int M, int N, int K,
                        long long lda, long long ldb, long long ldd, int a_mn, int b_mn, int out_mode, int splitk,
                        void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 8) || (ldb % 8) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15)) return -2;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.mode = MODE_GEMM; p.M = M; p.N = N; p.BN = pick_bn(N); p.a_mn = a_mn; p.b_mn = b_mn;
  if (b_mn && p.BN < 64) p.BN = 64;
  p.num_kb = (K + BK - 1) / BK;
  if (splitk < 1) splitk = 1;
  if (out_mode != OUT_ATOMIC_F32) splitk = 1;
  p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
  p.idesc = make_idesc(p.BN, a_mn, b_mn);
  p.out_mode = out_mode; p.D = D; p.ldd = ldd; p.bias = bias; p.stats = (splitk == 1) ? stats : nullptr;
  CUtensorMap ta, tb;
  rc = a_mn ? make_map_2d(&ta, A, K, M, lda, 64, 64) : make_map_2d(&ta, A, M, K, lda, 64, BM);
  if (rc) return rc;
  p.cluster = want_cluster((M + BM - 1) / BM, p.BN, b_mn);
  rc = b_mn ? make_map_2d(&tb, B, K, N, ldb, 64, 64) : make_map_2d(&tb, B, N, K, ldb, 64, p.BN / p.cluster);
  if (rc) return rc - 1000;
  dim3 grid((M + BM - 1) / BM, (N + p.BN - 1) / p.BN, (p.num_kb + p.kb_per_split - 1) / p.kb_per_split);
  return launch(ta, tb, p, grid, (cudaStream_t)stream);
}

// y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R,S,C]) (+bias); NHWC bf16; C % 64 == 0, K % 16 == 0.
// stats (optional): fp32 [2*K], must be zeroed by the caller; receives per-channel sum / sum of squares of y (fp32).
MPB_API int mpb_tc_conv_fprop_act(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H,
                                  int W, int C, int K, int R, int S, int P, int Q, int stride, int pad_h, int pad_w,
                                  int dil, int relu, void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  if (C % 8 != 0 || K % 8 != 0) return -2;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.mode = MODE_FPROP; p.M = N * P * Q; p.N = K; p.BN = pick_bn(K);
  p.num_kb = R * S * ((C + 63) / 64); p.kb_per_split = p.num_kb;
  p.idesc = make_idesc(p.BN, 0, 0);
  p.out_mode = OUT_BF16; p.D = y; p.ldd = K; p.bias = bias; p.stats = stats; p.relu = relu;
  p.C = C; p.R = R; p.S = S; p.Hout = P; p.Wout = Q; p.stride = stride; p.pad_h = pad_h; p.pad_w = pad_w; p.dil = dil;
  CUtensorMap ta, tb;
  rc = make_map_im2col(&ta, x, N, H, W, C, -pad_w, -pad_h, pad_w - (S - 1) * dil, pad_h - (R - 1) * dil, stride, BM);
  if (rc) return rc;
  p.cluster = want_cluster((p.M + BM - 1) / BM, p.BN, 0);
  rc = make_map_2d(&tb, w, K, (long long)R * S * C, (long long)R * S * C, 64, p.BN / p.cluster);
  if (rc) return rc - 1000;
  dim3 grid((p.M + BM - 1) / BM, (K + p.BN - 1) / p.BN, 1);
  return launch(ta, tb, p, grid, (cudaStream_t)stream);
}

MPB_API int mpb_tc_conv_fprop(const void* x, const void* w, void* y, const float* bias, float* stats, int N, int H, int W,
                              int C, int K, int R, int S, int P, int Q, int stride, int pad_h, int pad_w, int dil,
                              void* stream) {
  return mpb_tc_conv_fprop_act(x, w, y, bias, stats, N, H, W, C, K, R, S, P, Q, stride, pad_h, pad_w, dil, 0, stream);
}

// dx[N,H,W,C] = conv_transpose(dy[N,P,Q,K], w[K,R,S,C]) for stride-1, dilation-1 convolutions (P = H + 2*pad - R + 1).
MPB_API int mpb_tc_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R, int S,
                              int P, int Q, int pad_h, int pad_w, void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  if (K % 8 != 0 || C % 8 != 0) return -2;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.mode = MODE_DGRAD; p.M = N * H * W; p.N = C; p.BN = (C > 64) ? 128 : 64; p.a_mn = 0; p.b_mn = 1;
  if (env_int("MPB_TC_BN", 0) == 256 && C % 256 == 0) p.BN = 256;
  p.num_kb = R * S * ((K + 63) / 64); p.kb_per_split = p.num_kb;
  p.idesc = make_idesc(p.BN, 0, 1);
  p.out_mode = OUT_BF16; p.D = dx; p.ldd = C;
  p.C = K; p.R = R; p.S = S; p.Hout = H; p.Wout = W; p.stride = 1; p.dil = 1; p.Cin_total = C;
  p.pad_h = (R - 1) - pad_h; p.pad_w = (S - 1) - pad_w;
  CUtensorMap ta, tb;
  // gemm-M space is the H x W input grid: lower corner = pad - (S-1) = -pad', upper corner = lower + (W - Q)
  // (so that Q + upper - lower == W base pixels per row), cf. the dgrad corner rules of TMA im2col.
  rc = make_map_im2col(&ta, dy, N, P, Q, K, -p.pad_w, -p.pad_h, pad_w - (S - 1) + (W - Q), pad_h - (R - 1) + (H - P), 1, BM);
  if (rc) return rc;
  p.cluster = want_cluster((p.M + BM - 1) / BM, p.BN, 1);
  rc = make_map_2d(&tb, w, K, (long long)R * S * C, (long long)R * S * C, 64, 64);
  if (rc) return rc - 1000;
  dim3 grid((p.M + BM - 1) / BM, (C + p.BN - 1) / p.BN, 1);
  return launch(ta, tb, p, grid, (cudaStream_t)stream);
}

// dx for strided convolutions (stride >= 2), decomposed by output parity class (ph, pw): each class is a dense
// stride-1 problem over the filter taps r = r0 + stride*i, s = s0 + stride*j that reach it.  Stride 2 (every strided
// convolution with a data gradient in the zoo) runs as ONE launch: blockIdx.z selects the class, classes no tap
// reaches write their zeros themselves, so there is no memset and no per-class launch.  Other strides keep the
// per-class loop over a cleared dx.
MPB_API int mpb_tc_conv_dgrad_strided(const void* dy, const void* w, void* dx, int N, int H, int W, int C, int K, int R,
                                      int S, int P, int Q, int stride, int pad_h, int pad_w, void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  if (K % 8 != 0 || C % 8 != 0 || stride < 2) return -2;
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap tb;
  rc = make_map_2d(&tb, w, K, (long long)R * S * C, (long long)R * S * C, 64, 64);
  if (rc) return rc - 1000;
  const bool fused = (stride == 2) && env_int("MPB_TC_DGRAD_FUSED", 1) != 0;
  if (!fused) MPB_CUDA(cudaMemsetAsync(dx, 0, (size_t)N * H * W * C * 2, st));
  TcParams base;
  memset(&base, 0, sizeof(base));
  base.mode = MODE_DGRAD_STRIDED; base.N = C; base.BN = (C > 64) ? 128 : 64; base.a_mn = 0; base.b_mn = 1;
  base.idesc = make_idesc(base.BN, 0, 1);
  base.out_mode = OUT_BF16; base.D = dx; base.ldd = C;
  base.C = K; base.R = R; base.S = S; base.Cin_total = C; base.st = stride; base.Hfull = H; base.Wfull = W;
  CUtensorMap maps[4];
  int max_tiles = 0, nmaps = 0;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int Hp = (H - ph + stride - 1) / stride, Wp = (W - pw + stride - 1) / stride;
      const int r0 = (ph + pad_h) % stride, s0 = (pw + pad_w) % stride;
      const int nr = r0 < R ? (R - r0 + stride - 1) / stride : 0, ns = s0 < S ? (S - s0 + stride - 1) / stride : 0;
      const bool empty = (nr == 0 || ns == 0);
      if (Hp <= 0 || Wp <= 0) { if (fused) return -2; continue; }
      if (empty && !fused) continue;
      TcParams::StridedClass c;
      c.M = N * Hp * Wp; c.Hout = Hp; c.Wout = Wp; c.ph = ph; c.pw = pw; c.r0 = r0; c.s0 = s0; c.nr = nr; c.ns = ns;
      c.num_kb = empty ? 0 : nr * ns * ((K + 63) / 64);
      c.lower_h = empty ? 0 : (ph + pad_h - r0) / stride - (nr - 1);
      c.lower_w = empty ? 0 : (pw + pad_w - s0) / stride - (ns - 1);
      CUtensorMap ta;   // base pixel h' + lower_h for h' in [0, Hp): bounding box [lower, P + upper) with Hp positions
      if (!empty) {
        rc = make_map_im2col(&ta, dy, N, P, Q, K, c.lower_w, c.lower_h, Wp - Q + c.lower_w, Hp - P + c.lower_h, 1, BM);
        if (rc) return rc;
      } else {
        ta = tb;        // never dereferenced: a class without taps issues no loads
      }
      if (fused) {
        base.cls[nmaps] = c;
        maps[nmaps++] = ta;
        max_tiles = max(max_tiles, (c.M + BM - 1) / BM);
        continue;
      }
      TcParams p = base;
      p.M = c.M; p.Hout = Hp; p.Wout = Wp; p.num_kb = c.num_kb; p.kb_per_split = c.num_kb;
      p.ph = ph; p.pw = pw; p.r0 = r0; p.s0 = s0; p.nr = nr; p.ns = ns; p.lower_h = c.lower_h; p.lower_w = c.lower_w;
      dim3 grid((p.M + BM - 1) / BM, (C + p.BN - 1) / p.BN, 1);
      rc = launch(ta, tb, p, grid, st);
      if (rc) return rc;
    }
  if (!fused) return 0;
  base.nclass = nmaps;                       // 4
  base.M = base.cls[0].M; base.Hout = base.cls[0].Hout; base.Wout = base.cls[0].Wout;
  int kb_max = 1;
  for (int i = 0; i < nmaps; ++i) kb_max = max(kb_max, base.cls[i].num_kb);
  base.num_kb = kb_max; base.kb_per_split = kb_max;          // sizes the smem ring; each CTA uses its class's count
  ExtraMaps extra;
  for (int i = 1; i < 4; ++i) extra.m[i - 1] = maps[i < nmaps ? i : 0];
  base.cluster = want_cluster(max_tiles, base.BN, 1);
  dim3 grid(max_tiles, (C + base.BN - 1) / base.BN, nmaps);
  return launch(maps[0], tb, base, grid, st, &extra);
}

// dw[K,R,S,C] (fp32, pre-zeroed) += sum over pixels dy[n,p,q,k] * x[n, p*stride+r-pad, q*stride+s-pad, c]
MPB_API int mpb_tc_conv_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int K, int R, int S,
                              int P, int Q, int stride, int pad_h, int pad_w, int dil, int splitk, void* stream) {
  int rc = load_driver_fns();
  if (rc) return rc;
  if (C % 8 != 0 || K % 8 != 0) return -2;
  const long long npq = (long long)N * P * Q;
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.mode = MODE_WGRAD; p.M = K; p.N = C; p.BN = (C > 64) ? 128 : 64; p.a_mn = 1; p.b_mn = 1;
  p.num_kb = (int)((npq + BK - 1) / BK);
  if (splitk < 1) splitk = 1;
  p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
  p.idesc = make_idesc(p.BN, 1, 1);
  p.out_mode = OUT_ATOMIC_F32; p.D = dw; p.ldd = (long long)R * S * C;
  p.C = C; p.R = R; p.S = S; p.Hout = P; p.Wout = Q; p.stride = stride; p.pad_h = pad_h; p.pad_w = pad_w; p.dil = dil;
  p.tap_cols = (C + p.BN - 1) / p.BN; p.Cin_total = C;
  CUtensorMap ta, tb;
  rc = make_map_2d(&ta, dy, npq, K, K, 64, 64);
  if (rc) return rc;
  p.cluster = want_cluster((K + BM - 1) / BM, p.BN, 1);
  rc = make_map_im2col(&tb, x, N, H, W, C, -pad_w, -pad_h, pad_w - (S - 1) * dil, pad_h - (R - 1) * dil, stride, BK);
  if (rc) return rc - 1000;
  dim3 grid((K + BM - 1) / BM, R * S * p.tap_cols, (p.num_kb + p.kb_per_split - 1) / p.kb_per_split);
  return launch(ta, tb, p, grid, (cudaStream_t)stream);
}


// ------------------------------------------------------------------------------------------------------------
// Explicit im2col for convolutions whose input channel count is not a multiple of 8 (the 3-channel stems of every
// model in the zoo): rows = output pixels (n,p,q), columns = (r,s,c) padded with zeros to Kpad (multiple of 8).
// The result feeds mpb_tc_gemm for fprop (A operand, K-major) and wgrad (B operand, MN-major) -- the TMA im2col
// mode needs 16-byte pixel rows, which a 3-channel NHWC tensor does not have.
// ------------------------------------------------------------------------------------------------------------
namespace {
constexpr int IM2COL_QT = 64;   // output pixels (consecutive q of one output row) per CTA

// One CTA = IM2COL_QT consecutive output pixels of one (n, p) row.  The R input rows they touch are staged in
// shared memory with coalesced loads (each is one contiguous span of ((QT-1)*stride + S) * C elements), then every
// thread assembles 8-element (16 B) column chunks from shared memory and stores them coalesced.
__global__ void __launch_bounds__(256) im2col_nhwc_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                          int N, int H, int W, int C, int R, int S, int P, int Q,
                                                          int stride, int pad_h, int pad_w, int Kpad, int qtiles) {
  extern __shared__ __align__(16) unsigned char im2col_smem[];
  const int span_px = (IM2COL_QT - 1) * stride + S;
  const int span = span_px * C;                                   // elements per staged input row
  __nv_bfloat16* rows = reinterpret_cast<__nv_bfloat16*>(im2col_smem);            // [R][span]
  int* tab = reinterpret_cast<int*>(im2col_smem + (((size_t)R * span * 2 + 15) / 16) * 16);   // [Kpad]: r<<20 | (s*C+c)
  const int Kreal = R * S * C;
  for (int j = threadIdx.x; j < Kpad; j += blockDim.x) tab[j] = j < Kreal ? (((j / (S * C)) << 20) | (j % (S * C))) : -1;
  const int qt = blockIdx.x % qtiles;
  const int pp = (blockIdx.x / qtiles) % P;
  const int n = blockIdx.x / (qtiles * P);
  const int q0 = qt * IM2COL_QT;
  const int h0 = pp * stride - pad_h, w0 = q0 * stride - pad_w;
  const __nv_bfloat16 zero = __float2bfloat16_rn(0.f);
  // staging: each input row contributes one contiguous span; bounds become a [e_lo, e_hi) range (no per-element division)
  const int e_lo = w0 < 0 ? -w0 * C : 0;
  const int e_hi = min(span, (W - w0) * C);
  for (int r = 0; r < R; ++r) {
    const int h = h0 + r;
    const bool h_ok = h >= 0 && h < H;
    const __nv_bfloat16* src = x + (((long long)n * H + (h_ok ? h : 0)) * W + w0) * C;
    __nv_bfloat16* dst = rows + r * span;
    for (int e = threadIdx.x; e < span; e += blockDim.x) dst[e] = (h_ok && e >= e_lo && e < e_hi) ? src[e] : zero;
  }
  __syncthreads();
  const int chunks = Kpad / 8;
  const int nq = min(IM2COL_QT, Q - q0);
  // a warp handles 8 pixels x 4 chunks: table reads are broadcast / conflict-free, pixel-strided row reads hit
  // distinct banks, and the warp's stores are 8 rows x 64 contiguous bytes
  const int qgroups = (nq + 7) / 8, cgroups = (chunks + 3) / 4;
  const int lane = threadIdx.x & 31;
  for (int grp = threadIdx.x >> 5; grp < qgroups * cgroups; grp += blockDim.x >> 5) {
    const int ql = (grp % qgroups) * 8 + (lane >> 2), ch = (grp / qgroups) * 4 + (lane & 3);
    if (ql >= nq || ch >= chunks) continue;
    uint4 u;
    __nv_bfloat16* v = reinterpret_cast<__nv_bfloat16*>(&u);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int t = tab[ch * 8 + k];
      v[k] = t >= 0 ? rows[(t >> 20) * span + ql * stride * C + (t & 0xFFFFF)] : zero;
    }
    const long long m = ((long long)n * P + pp) * Q + q0 + ql;
    *reinterpret_cast<uint4*>(out + m * Kpad + ch * 8) = u;
  }
}
}  // namespace

MPB_API int mpb_im2col_nhwc(const void* x, void* out, int N, int H, int W, int C, int R, int S, int P, int Q, int stride,
                            int pad_h, int pad_w, int Kpad, void* stream) {
  if (Kpad % 8 != 0 || Kpad < R * S * C || R > 255 || S > 255 || C > 32767) return -2;
  if ((long long)N * P * Q <= 0) return 0;
  const int qtiles = (Q + IM2COL_QT - 1) / IM2COL_QT;
  const long long blocks = (long long)N * P * qtiles;
  const size_t span = (size_t)((IM2COL_QT - 1) * stride + S) * C;
  const size_t smem = (((size_t)R * span * 2 + 15) / 16) * 16 + (size_t)Kpad * sizeof(int);
  if (blocks > 2147483647ll || smem > 48 * 1024) return -3;
  im2col_nhwc_kernel<<<(unsigned)blocks, 256, smem, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)out, N, H, W, C, R, S, P, Q, stride, pad_h, pad_w, Kpad, qtiles);
  MPB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Space-to-depth for stride-2 stems with few input channels (7x7/2 on 3 channels: ResNet, DenseNet, SqueezeNet).
//   out[n][h2][j][(b*2 + b')*C + c] = x[n][2*h2 + b][2*(j - lpad) + b'][c]      (16 elements per output pixel, zero tails)
// with `lpad` zero pixels on the left and enough on the right that every window of 4 neighbouring pixels exists.
// A 7x7/2 convolution then is a 4x4/1 convolution over 4*C channels; reading 4 neighbouring 16-element pixels as ONE
// 64-element pixel (tensor map with a 16-element pixel stride) makes it a 4x1 stride-1 convolution with 64 channels,
// which the halo-tile fprop / wgrad kernels run without any im2col matrix (976 MB per step for ResNet at batch 256).
// ------------------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) s2d_stem_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                       int N, int H, int W, int C, int H2, int Wp, int lpad) {
  const long long total = (long long)N * H2 * Wp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % Wp);
    const long long t = i / Wp;
    const int h2 = (int)(t % H2);
    const int n = (int)(t / H2);
    const int px = j - lpad;
    __align__(16) __nv_bfloat16 v[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) v[e] = __float2bfloat16(0.f);
    if (px >= 0 && 2 * px + 1 < W) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const __nv_bfloat16* src = x + (((long long)n * H + 2 * h2 + b) * W + 2 * px) * C;     // 2 pixels = 2*C contiguous
        for (int e = 0; e < 2 * C; ++e) v[b * 2 * C + e] = src[e];
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + i * 16);
    dst[0] = reinterpret_cast<const uint4*>(v)[0];
    dst[1] = reinterpret_cast<const uint4*>(v)[1];
  }
}
}  // namespace

// x: NHWC bf16 [N][H][W][C] with 4*C <= 16, H and W even; out: [N][H/2][Wp][16] bf16.
MPB_API int mpb_s2d_stem(const void* x, void* out, int N, int H, int W, int C, int Wp, int lpad, void* stream) {
  if (4 * C > 16 || (H & 1) || (W & 1) || lpad + W / 2 > Wp) return -2;
  const long long total = (long long)N * (H / 2) * Wp;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)kSMs * 16) blocks = (long long)kSMs * 16;
  s2d_stem_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, N, H, W, C,
                                                                 H / 2, Wp, lpad);
  return (int)cudaGetLastError();
}
