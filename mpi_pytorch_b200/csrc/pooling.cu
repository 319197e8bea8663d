// K8 pooling (NHWC): max (with uint8 window-argmax for an exact, atomic-free backward), average
// (count_include_pad / ceil_mode), adaptive average -> any output size; plus K11 elementwise bits
// (ReLU fwd/bwd, dropout with a stateless hash mask).  Replaces the ATen CPU pooling / relu / dropout
// calls listed in SURVEY §2.5.  8 channels per thread, C contiguous.
#include "common.cuh"

using namespace mpb;

namespace {

struct PoolGeom {
  int N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw;
};

template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                          unsigned char* __restrict__ idx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    // 32-bit index math (64-bit div/mod costs ~100 instructions each and dominated this kernel)
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int ho = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    float best[8];
    unsigned char bi[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; bi[k] = 0; }
    bool first = true;
    for (int r = 0; r < g.kh; ++r) {
      const int h = ho * g.sh - g.ph + r;
      if (h < 0 || h >= g.H) continue;
      for (int s = 0; s < g.kw; ++s) {
        const int w = wo * g.sw - g.pw + s;
        if (w < 0 || w >= g.W) continue;
        float v[8];
        load8(x + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8, v);
        const unsigned char pos = (unsigned char)(r * g.kw + s);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (first || v[k] > best[k] || v[k] != v[k]) { best[k] = v[k]; bi[k] = pos; }   // torch: first max, NaN wins
        first = false;
      }
    }
    const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
    store8(y + o, best);
    if (idx) *reinterpret_cast<uint2*>(idx + o) = *reinterpret_cast<uint2*>(bi);
  }
}

// 2x2 / stride 2 / no padding (every VGG stage, models.py:56-63): windows do not overlap, so one thread owns one output
// vector and its 2x2 input quad -- four independent 16-byte loads (forward) or stores (backward) in flight, one index
// decomposition per quad instead of one per input pixel.  Argmax positions are r*2+s like the generic kernel's.
template <typename T>
__global__ void __launch_bounds__(256) maxpool2x2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                             unsigned char* __restrict__ idx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
  const long long row = (long long)g.W * g.C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int ho = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    const T* p = x + (((long long)n * g.H + 2 * ho) * g.W + 2 * wo) * g.C + cv * 8;
    float v[4][8];
    load8(p, v[0]); load8(p + g.C, v[1]); load8(p + row, v[2]); load8(p + row + g.C, v[3]);
    float m[8];
    uint2 pk = make_uint2(0u, 0u);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float best = v[0][k];
      unsigned pos = 0;
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (v[q][k] > best || v[q][k] != v[q][k]) { best = v[q][k]; pos = q; }     // first maximum wins; NaN propagates
      m[k] = best;
      if (k < 4) pk.x |= pos << (8 * k); else pk.y |= pos << (8 * (k - 4));
    }
    const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
    store8(y + o, m);
    if (idx) *reinterpret_cast<uint2*>(idx + o) = pk;
  }
}
template <typename T>
__global__ void __launch_bounds__(256) maxpool2x2_bwd_kernel(const T* __restrict__ dy, const unsigned char* __restrict__ idx,
                                                             T* __restrict__ dx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
  const long long row = (long long)g.W * g.C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int ho = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
    const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
    float gv[8];
    load8(dy + o, gv);
    float out[4][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const unsigned pos = ((k < 4 ? pk.x : pk.y) >> (8 * (k & 3))) & 0xFFu;
#pragma unroll
      for (int q = 0; q < 4; ++q) out[q][k] = (pos == (unsigned)q) ? gv[k] : 0.f;
    }
    T* p = dx + (((long long)n * g.H + 2 * ho) * g.W + 2 * wo) * g.C + cv * 8;
    store8(p, out[0]); store8(p + g.C, out[1]); store8(p + row, out[2]); store8(p + row + g.C, out[3]);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dy, const unsigned char* __restrict__ idx,
                                                          T* __restrict__ dx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.H * g.W * cvecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int w = (int)(t % (unsigned)g.W); t /= (unsigned)g.W;
    const int h = (int)(t % (unsigned)g.H);
    const int n = (int)(t / (unsigned)g.H);
    float acc[8] = {0};
    // output windows containing (h, w): ho*sh - ph <= h < ho*sh - ph + kh
    int ho_lo = (h + g.ph - g.kh + g.sh) / g.sh; if (h + g.ph - g.kh + 1 <= 0) ho_lo = 0;
    int wo_lo = (w + g.pw - g.kw + g.sw) / g.sw; if (w + g.pw - g.kw + 1 <= 0) wo_lo = 0;
    const int ho_hi = min((h + g.ph) / g.sh, g.Ho - 1), wo_hi = min((w + g.pw) / g.sw, g.Wo - 1);
    if (ho_hi - ho_lo < 2 && wo_hi - wo_lo < 2) {
      // common case (kernel <= 2*stride): at most 2x2 windows cover a pixel -- issue all candidate loads first
      uint2 pk[4];
      float gv[4][8];
      bool ok[4];
      unsigned char pos[4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int ho = ho_lo + a, wo = wo_lo + b, c4 = a * 2 + b;
          const int r = h - (ho * g.sh - g.ph), s = w - (wo * g.sw - g.pw);
          ok[c4] = ho <= ho_hi && wo <= wo_hi && r >= 0 && r < g.kh && s >= 0 && s < g.kw;
          pos[c4] = (unsigned char)(r * g.kw + s);
          if (ok[c4]) {
            const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
            pk[c4] = *reinterpret_cast<const uint2*>(idx + o);
            load8(dy + o, gv[c4]);
          }
        }
#pragma unroll
      for (int c4 = 0; c4 < 4; ++c4) {
        if (!ok[c4]) continue;
        const unsigned char* bi = reinterpret_cast<const unsigned char*>(&pk[c4]);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += (bi[k] == pos[c4]) ? gv[c4][k] : 0.f;
      }
    } else {
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
        const int r = h - (ho * g.sh - g.ph);
        if (r < 0 || r >= g.kh) continue;
        for (int wo = wo_lo; wo <= wo_hi; ++wo) {
          const int s = w - (wo * g.sw - g.pw);
          if (s < 0 || s >= g.kw) continue;
          const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
          const uint2 packed = *reinterpret_cast<const uint2*>(idx + o);
          const unsigned char* bi = reinterpret_cast<const unsigned char*>(&packed);
          float gy[8];
          load8(dy + o, gy);
          const unsigned char pos = (unsigned char)(r * g.kw + s);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += (bi[k] == pos) ? gy[k] : 0.f;
        }
      }
    }
    store8(dx + i * 8, acc);
  }
}

// 3x3 windows (every max pool in the zoo): the nine tap loads are issued before the first compare and stay packed
// (16 bytes of registers per tap for bf16) until they are compared, so each thread keeps nine requests in flight at
// <= 64 registers.
template <typename T> struct Raw8;
template <> struct Raw8<float> { float4 a, b; };
template <> struct Raw8<__nv_bfloat16> { uint4 u; };
__device__ __forceinline__ void load_raw(const float* p, Raw8<float>& r) {
  r.a = *reinterpret_cast<const float4*>(p); r.b = *reinterpret_cast<const float4*>(p + 4);
}
__device__ __forceinline__ void load_raw(const __nv_bfloat16* p, Raw8<__nv_bfloat16>& r) {
  r.u = *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ void unpack(const Raw8<float>& r, float (&v)[8]) {
  v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
}
__device__ __forceinline__ void unpack(const Raw8<__nv_bfloat16>& r, float (&v)[8]) {
  const unsigned w[4] = {r.u.x, r.u.y, r.u.z, r.u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

template <typename T>
__global__ void __launch_bounds__(256, sizeof(T) == 2 ? 4 : 2) maxpool3x3_fwd_kernel(const T* __restrict__ x,
                                                                                     T* __restrict__ y,
                                                                                     unsigned char* __restrict__ idx,
                                                                                     PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int ho = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    const int h0 = ho * g.sh - g.ph, w0 = wo * g.sw - g.pw;
    Raw8<T> raw[9];
    unsigned okmask = 0;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int h = h0 + r, w = w0 + s;
        if (h >= 0 && h < g.H && w >= 0 && w < g.W) {
          okmask |= 1u << (r * 3 + s);
          load_raw(x + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8, raw[r * 3 + s]);
        }
      }
    float best[8];
    unsigned bi_lo = 0, bi_hi = 0;                      // argmax bytes of channels 0..3 / 4..7
#pragma unroll
    for (int k = 0; k < 8; ++k) best[k] = -INFINITY;
    bool first = true;
#pragma unroll
    for (int p = 0; p < 9; ++p) {
      if (!((okmask >> p) & 1u)) continue;
      float v[8];
      unpack(raw[p], v);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (first || v[k] > best[k] || v[k] != v[k]) {   // torch: first max wins, NaN propagates
          best[k] = v[k];
          const int sft = 8 * (k & 3);
          if (k < 4) bi_lo = (bi_lo & ~(0xFFu << sft)) | ((unsigned)p << sft);
          else bi_hi = (bi_hi & ~(0xFFu << sft)) | ((unsigned)p << sft);
        }
      first = false;
    }
    const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
    store8(y + o, best);
    if (idx) *reinterpret_cast<uint2*>(idx + o) = make_uint2(bi_lo, bi_hi);
  }
}

// bf16 specialisation of the 3x3 forward: the generic kernel spends ~600 instructions per thread on per-element
// compare / NaN / byte-insert chains.  Here the maximum is taken with packed bf16x2 max (4 instructions per tap for 8
// channels) and the argmax is recovered in a second pass, taps 8..0, with one packed equality mask and one LOP3 select
// per channel pair -- "first maximum wins" as in torch.  (A NaN makes the max NaN and leaves argmax 0.)
__global__ void __launch_bounds__(256) maxpool3x3_fwd_bf16_kernel(const __nv_bfloat16* __restrict__ x,
                                                                  __nv_bfloat16* __restrict__ y,
                                                                  unsigned char* __restrict__ idx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int ho = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    const int h0 = ho * g.sh - g.ph, w0 = wo * g.sw - g.pw;
    uint4 raw[9];
    unsigned okmask = 0;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int h = h0 + r, w = w0 + s;
        raw[r * 3 + s] = make_uint4(0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u);      // -inf
        if (h >= 0 && h < g.H && w >= 0 && w < g.W) {
          okmask |= 1u << (r * 3 + s);
          raw[r * 3 + s] = *reinterpret_cast<const uint4*>(x + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8);
        }
      }
    uint4 mx = raw[0];
    __nv_bfloat162* m = reinterpret_cast<__nv_bfloat162*>(&mx);
#pragma unroll
    for (int p = 1; p < 9; ++p) {
      const __nv_bfloat162* v = reinterpret_cast<const __nv_bfloat162*>(&raw[p]);
#pragma unroll
      for (int k = 0; k < 4; ++k) m[k] = __hmax2_nan(m[k], v[k]);
    }
    unsigned id[4] = {0u, 0u, 0u, 0u};                       // two 16-bit tap indices per channel pair
#pragma unroll
    for (int p = 8; p >= 0; --p) {
      if (!((okmask >> p) & 1u)) continue;
      const __nv_bfloat162* v = reinterpret_cast<const __nv_bfloat162*>(&raw[p]);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned eq = __heq2_mask(v[k], m[k]);
        id[k] = (eq & ((unsigned)p * 0x00010001u)) | (~eq & id[k]);
      }
    }
    const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
    *reinterpret_cast<uint4*>(y + o) = mx;
    if (idx) *reinterpret_cast<uint2*>(idx + o) = make_uint2(__byte_perm(id[0], id[1], 0x6420), __byte_perm(id[2], id[3], 0x6420));
  }
}

// 3x3 / stride 2 / pad 1 on even H, W (the ResNet / DenseNet stem pool): one thread owns a 2x2 quad of input pixels.
// The quad (2a..2a+1, 2b..2b+1) is covered by exactly the four windows (a..a+1, b..b+1), so four (argmax, dy) loads
// produce four outputs -- the per-pixel formulation needs nine loads for the same four pixels.
template <typename T>
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(const T* __restrict__ dy,
                                                               const unsigned char* __restrict__ idx,
                                                               T* __restrict__ dx, PoolGeom g) {
  const int cvecs = g.C / 8;
  const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;        // Ho = H/2, Wo = W/2 quads
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cv = (int)(iu % (unsigned)cvecs);
    unsigned t = iu / (unsigned)cvecs;
    const int b = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
    const int a = (int)(t % (unsigned)g.Ho);
    const int n = (int)(t / (unsigned)g.Ho);
    uint2 pk[4];
    float gv[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ho = a + (q >> 1), wo = b + (q & 1);
      if (ho < g.Ho && wo < g.Wo) {
        const long long o = (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8;
        pk[q] = *reinterpret_cast<const uint2*>(idx + o);
        load8(dy + o, gv[q]);
      } else {
        pk[q] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);                   // position 255 never matches
#pragma unroll
        for (int k = 0; k < 8; ++k) gv[q][k] = 0.f;
      }
    }
    float o00[8], o01[8], o10[8], o11[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int sft = 8 * (k & 3);
      const unsigned p0 = ((k < 4 ? pk[0].x : pk[0].y) >> sft) & 0xFFu;
      const unsigned p1 = ((k < 4 ? pk[1].x : pk[1].y) >> sft) & 0xFFu;
      const unsigned p2 = ((k < 4 ? pk[2].x : pk[2].y) >> sft) & 0xFFu;
      const unsigned p3 = ((k < 4 ? pk[3].x : pk[3].y) >> sft) & 0xFFu;
      // window (a, b) sees the quad at taps (1..2, 1..2); (a, b+1) at (1..2, 0); (a+1, b) at (0, 1..2); (a+1, b+1) at (0, 0)
      o00[k] = (p0 == 4 ? gv[0][k] : 0.f);
      o01[k] = (p0 == 5 ? gv[0][k] : 0.f) + (p1 == 3 ? gv[1][k] : 0.f);
      o10[k] = (p0 == 7 ? gv[0][k] : 0.f) + (p2 == 1 ? gv[2][k] : 0.f);
      o11[k] = (p0 == 8 ? gv[0][k] : 0.f) + (p1 == 6 ? gv[1][k] : 0.f) + (p2 == 2 ? gv[2][k] : 0.f) +
               (p3 == 0 ? gv[3][k] : 0.f);
    }
    const long long base = (((long long)n * g.H + 2 * a) * g.W + 2 * b) * g.C + cv * 8;
    store8(dx + base, o00);
    store8(dx + base + g.C, o01);
    store8(dx + base + (long long)g.W * g.C, o10);
    store8(dx + base + (long long)g.W * g.C + g.C, o11);
  }
}

// avg pool; divisor semantics follow torch (count_include_pad with ceil_mode clipping).
template <typename T, bool BWD>
__global__ void __launch_bounds__(256) avgpool_kernel(const T* __restrict__ in, T* __restrict__ out, PoolGeom g,
                                                      int count_include_pad) {
  const int cvecs = g.C / 8;
  if (!BWD) {
    const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int cv = (int)(i % cvecs);
      long long t = i / cvecs;
      const int wo = (int)(t % g.Wo); t /= g.Wo;
      const int ho = (int)(t % g.Ho);
      const int n = (int)(t / g.Ho);
      int h0 = ho * g.sh - g.ph, w0 = wo * g.sw - g.pw;
      int h1 = min(h0 + g.kh, g.H + g.ph), w1 = min(w0 + g.kw, g.W + g.pw);
      const int pool = (h1 - h0) * (w1 - w0);
      h0 = max(h0, 0); w0 = max(w0, 0); h1 = min(h1, g.H); w1 = min(w1, g.W);
      const float div = (float)(count_include_pad ? pool : (h1 - h0) * (w1 - w0));
      float acc[8] = {0};
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) {
          float v[8];
          load8(in + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8, v);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] /= div;
      store8(out + i * 8, acc);
    }
  } else {  // in = dy [N,Ho,Wo,C], out = dx [N,H,W,C]
    const long long total = (long long)g.N * g.H * g.W * cvecs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int cv = (int)(i % cvecs);
      long long t = i / cvecs;
      const int w = (int)(t % g.W); t /= g.W;
      const int h = (int)(t % g.H);
      const int n = (int)(t / g.H);
      float acc[8] = {0};
      int ho_lo = (h + g.ph - g.kh + g.sh) / g.sh; if (h + g.ph - g.kh + 1 <= 0) ho_lo = 0;
      int wo_lo = (w + g.pw - g.kw + g.sw) / g.sw; if (w + g.pw - g.kw + 1 <= 0) wo_lo = 0;
      const int ho_hi = min((h + g.ph) / g.sh, g.Ho - 1), wo_hi = min((w + g.pw) / g.sw, g.Wo - 1);
      for (int ho = ho_lo; ho <= ho_hi; ++ho) {
        int h0 = ho * g.sh - g.ph;
        if (h < h0 || h >= h0 + g.kh) continue;
        for (int wo = wo_lo; wo <= wo_hi; ++wo) {
          int w0 = wo * g.sw - g.pw;
          if (w < w0 || w >= w0 + g.kw) continue;
          int hh0 = h0, ww0 = w0;
          int h1 = min(hh0 + g.kh, g.H + g.ph), w1 = min(ww0 + g.kw, g.W + g.pw);
          const int pool = (h1 - hh0) * (w1 - ww0);
          hh0 = max(hh0, 0); ww0 = max(ww0, 0); h1 = min(h1, g.H); w1 = min(w1, g.W);
          const float div = (float)(count_include_pad ? pool : (h1 - hh0) * (w1 - ww0));
          float gy[8];
          load8(in + (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8, gy);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += gy[k] / div;
        }
      }
      store8(out + i * 8, acc);
    }
  }
}

// Windows up to 3x3 (Inception's 3x3/1 branch pools, DenseNet's 2x2/2 transitions): 32-bit index arithmetic and all tap
// loads issued before the first add.  The generic kernel above decodes with 64-bit div/mod and walks the window with
// a dependent load per tap.
template <typename T, bool BWD>
__global__ void __launch_bounds__(256) avgpool_k3_kernel(const T* __restrict__ in, T* __restrict__ out, PoolGeom g,
                                                         int count_include_pad) {
  const int cvecs = g.C / 8;
  const unsigned total = (unsigned)((long long)g.N * (BWD ? g.H * g.W : g.Ho * g.Wo) * cvecs);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int cv = (int)(i % (unsigned)cvecs);
    unsigned t = i / (unsigned)cvecs;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (!BWD) {
      const int wo = (int)(t % (unsigned)g.Wo); t /= (unsigned)g.Wo;
      const int ho = (int)(t % (unsigned)g.Ho);
      const int n = (int)(t / (unsigned)g.Ho);
      const int h0 = ho * g.sh - g.ph, w0 = wo * g.sw - g.pw;
      const int h1 = min(h0 + g.kh, g.H + g.ph), w1 = min(w0 + g.kw, g.W + g.pw);
      const int ch0 = max(h0, 0), cw0 = max(w0, 0), ch1 = min(h1, g.H), cw1 = min(w1, g.W);
      const float inv = 1.f / (float)(count_include_pad ? (h1 - h0) * (w1 - w0) : (ch1 - ch0) * (cw1 - cw0));
      Raw8<T> raw[9];
      bool ok[9];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s2 = 0; s2 < 3; ++s2) {
          const int h = h0 + r, w = w0 + s2;
          ok[r * 3 + s2] = r < g.kh && s2 < g.kw && h >= 0 && h < g.H && w >= 0 && w < g.W;
          if (ok[r * 3 + s2]) load_raw(in + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8, raw[r * 3 + s2]);
        }
#pragma unroll
      for (int p = 0; p < 9; ++p) {
        if (!ok[p]) continue;
        float v[8];
        unpack(raw[p], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += v[k];
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] *= inv;
    } else {
      const int w = (int)(t % (unsigned)g.W); t /= (unsigned)g.W;
      const int h = (int)(t % (unsigned)g.H);
      const int n = (int)(t / (unsigned)g.H);
      int ho_lo = (h + g.ph - g.kh + g.sh) / g.sh; if (h + g.ph - g.kh + 1 <= 0) ho_lo = 0;
      int wo_lo = (w + g.pw - g.kw + g.sw) / g.sw; if (w + g.pw - g.kw + 1 <= 0) wo_lo = 0;
      const int ho_hi = min((h + g.ph) / g.sh, g.Ho - 1), wo_hi = min((w + g.pw) / g.sw, g.Wo - 1);
      Raw8<T> raw[9];
      float scale[9];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          const int ho = ho_lo + a, wo = wo_lo + b;
          const int h0 = ho * g.sh - g.ph, w0 = wo * g.sw - g.pw;
          const bool covers = ho <= ho_hi && wo <= wo_hi && h >= h0 && h < h0 + g.kh && w >= w0 && w < w0 + g.kw;
          scale[a * 3 + b] = 0.f;
          if (covers) {
            const int h1 = min(h0 + g.kh, g.H + g.ph), w1 = min(w0 + g.kw, g.W + g.pw);
            const int cnt = count_include_pad ? (h1 - h0) * (w1 - w0)
                                              : (min(h1, g.H) - max(h0, 0)) * (min(w1, g.W) - max(w0, 0));
            scale[a * 3 + b] = 1.f / (float)cnt;
            load_raw(in + (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8, raw[a * 3 + b]);
          }
        }
#pragma unroll
      for (int p = 0; p < 9; ++p) {
        if (scale[p] == 0.f) continue;
        float v[8];
        unpack(raw[p], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(v[k], scale[p], acc[k]);
      }
    }
    store8(out + (size_t)i * 8, acc);
  }
}

__device__ __forceinline__ int ada_start(int o, int in, int out) { return (int)(((long long)o * in) / out); }
__device__ __forceinline__ int ada_end(int o, int in, int out) { return (int)((((long long)(o + 1)) * in + out - 1) / out); }

template <typename T, bool BWD>
__global__ void __launch_bounds__(256) adaptive_avg_kernel(const T* __restrict__ in, T* __restrict__ out, PoolGeom g) {
  const int cvecs = g.C / 8;
  if (!BWD) {
    const long long total = (long long)g.N * g.Ho * g.Wo * cvecs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int cv = (int)(i % cvecs);
      long long t = i / cvecs;
      const int wo = (int)(t % g.Wo); t /= g.Wo;
      const int ho = (int)(t % g.Ho);
      const int n = (int)(t / g.Ho);
      const int h0 = ada_start(ho, g.H, g.Ho), h1 = ada_end(ho, g.H, g.Ho);
      const int w0 = ada_start(wo, g.W, g.Wo), w1 = ada_end(wo, g.W, g.Wo);
      float acc[8] = {0};
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) {
          float v[8];
          load8(in + (((long long)n * g.H + h) * g.W + w) * g.C + cv * 8, v);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
      const float inv = 1.f / (float)((h1 - h0) * (w1 - w0));
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] *= inv;
      store8(out + i * 8, acc);
    }
  } else {
    const long long total = (long long)g.N * g.H * g.W * cvecs;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int cv = (int)(i % cvecs);
      long long t = i / cvecs;
      const int w = (int)(t % g.W); t /= g.W;
      const int h = (int)(t % g.H);
      const int n = (int)(t / g.H);
      float acc[8] = {0};
      for (int ho = 0; ho < g.Ho; ++ho) {
        const int h0 = ada_start(ho, g.H, g.Ho), h1 = ada_end(ho, g.H, g.Ho);
        if (h < h0 || h >= h1) continue;
        for (int wo = 0; wo < g.Wo; ++wo) {
          const int w0 = ada_start(wo, g.W, g.Wo), w1 = ada_end(wo, g.W, g.Wo);
          if (w < w0 || w >= w1) continue;
          float gy[8];
          load8(in + (((long long)n * g.Ho + ho) * g.Wo + wo) * g.C + cv * 8, gy);
          const float inv = 1.f / (float)((h1 - h0) * (w1 - w0));
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += gy[k] * inv;
        }
      }
      store8(out + i * 8, acc);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) relu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n8) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
    store8(y + i * 8, v);
  }
}
template <typename T>
__global__ void __launch_bounds__(256) relu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                       T* __restrict__ dx, long long n8) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float g[8], v[8];
    load8(dy + i * 8, g);
    load8(y + i * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = v[k] > 0.f ? g[k] : 0.f;
    store8(dx + i * 8, g);
  }
}
// y = x * keep/(1-p) with keep = hash(seed, i) >= p*2^32 ; the same call with the same seed on dy is the backward.
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n8,
                                                      uint32_t seed, uint32_t thresh, float scale,
                                                      const long long* __restrict__ epoch) {
  // `epoch` (nullable): a device-resident step counter mixed into the seed, so a CUDA-graph replay of the same launch
  // draws a fresh mask every step while forward and backward of one step still agree
  if (epoch) seed ^= hash32((uint32_t)(*epoch) * 0x85EBCA6Bu + 0x27D4EB2Fu);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t h = hash32((uint32_t)(i * 8 + k) ^ seed) ^ hash32((uint32_t)((i * 8 + k) >> 32) + seed * 0x9E3779B9u);
      v[k] = (h >= thresh) ? v[k] * scale : 0.f;
    }
    store8(y + i * 8, v);
  }
}

}  // namespace

#define DISPATCH_T(dtype, ...)                               \
  if (dtype == MPB_BF16) { using T = __nv_bfloat16; __VA_ARGS__ } \
  else { using T = float; __VA_ARGS__ }

static PoolGeom geom(int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh, int sw, int ph, int pw) {
  PoolGeom g; g.N = N; g.H = H; g.W = W; g.C = C; g.Ho = Ho; g.Wo = Wo; g.kh = kh; g.kw = kw; g.sh = sh; g.sw = sw; g.ph = ph; g.pw = pw;
  return g;
}

MPB_API int mpb_maxpool_fwd(const void* x, void* y, unsigned char* idx, int N, int H, int W, int C, int Ho, int Wo,
                            int kh, int kw, int sh, int sw, int ph, int pw, int dtype, void* stream) {
  if (C % 8 != 0 || kh * kw > 255) return -1;
  if ((long long)N * H * W * (C / 8) >= (1ll << 32)) return -3;
  const PoolGeom g = geom(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw);
  const long long total = (long long)N * Ho * Wo * (C / 8);
  if (kh == 3 && kw == 3) {
    if (dtype == MPB_BF16) {
      maxpool3x3_fwd_bf16_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, idx, g);
    } else {
      maxpool3x3_fwd_kernel<float><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const float*)x, (float*)y, idx, g);
    }
  } else if (kh == 2 && kw == 2 && sh == 2 && sw == 2 && ph == 0 && pw == 0 && H % 2 == 0 && W % 2 == 0 && Ho == H / 2 &&
             Wo == W / 2) {
    DISPATCH_T(dtype, maxpool2x2_fwd_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, idx, g);)
  } else {
    DISPATCH_T(dtype, maxpool_fwd_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, idx, g);)
  }
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_maxpool_bwd(const void* dy, const unsigned char* idx, void* dx, int N, int H, int W, int C, int Ho,
                            int Wo, int kh, int kw, int sh, int sw, int ph, int pw, int dtype, void* stream) {
  if (C % 8 != 0) return -1;
  const PoolGeom g = geom(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw);
  const long long total = (long long)N * H * W * (C / 8);
  if (kh == 3 && kw == 3 && sh == 2 && sw == 2 && ph == 1 && pw == 1 && H % 2 == 0 && W % 2 == 0 && Ho == H / 2 &&
      Wo == W / 2) {
    const long long quads = (long long)N * Ho * Wo * (C / 8);
    DISPATCH_T(dtype, maxpool3x3s2_bwd_kernel<T><<<grid_for(quads, 256), 256, 0, (cudaStream_t)stream>>>((const T*)dy, idx, (T*)dx, g);)
  } else if (kh == 2 && kw == 2 && sh == 2 && sw == 2 && ph == 0 && pw == 0 && H % 2 == 0 && W % 2 == 0 && Ho == H / 2 &&
             Wo == W / 2) {
    const long long outs = (long long)N * Ho * Wo * (C / 8);
    DISPATCH_T(dtype, maxpool2x2_bwd_kernel<T><<<grid_for(outs, 256), 256, 0, (cudaStream_t)stream>>>((const T*)dy, idx, (T*)dx, g);)
  } else {
    DISPATCH_T(dtype, maxpool_bwd_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>((const T*)dy, idx, (T*)dx, g);)
  }
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_avgpool(const void* in, void* out, int N, int H, int W, int C, int Ho, int Wo, int kh, int kw, int sh,
                        int sw, int ph, int pw, int count_include_pad, int backward, int dtype, void* stream) {
  if (C % 8 != 0) return -1;
  const PoolGeom g = geom(N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw);
  const long long total = (long long)N * (backward ? (long long)H * W : (long long)Ho * Wo) * (C / 8);
  cudaStream_t st = (cudaStream_t)stream;
  // windows up to 3x3 whose backward needs at most 3x3 covering windows (kernel <= 3 * stride), 32-bit element count
  const bool k3 = kh <= 3 && kw <= 3 && kh <= 3 * sh && kw <= 3 * sw && total < (1ll << 31);
  if (k3) {
    DISPATCH_T(dtype, if (backward) avgpool_k3_kernel<T, true><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g, count_include_pad);
               else avgpool_k3_kernel<T, false><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g, count_include_pad);)
    MPB_CHECK_LAUNCH();
    return 0;
  }
  DISPATCH_T(dtype, if (backward) avgpool_kernel<T, true><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g, count_include_pad);
             else avgpool_kernel<T, false><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g, count_include_pad);)
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_adaptive_avgpool(const void* in, void* out, int N, int H, int W, int C, int Ho, int Wo, int backward,
                                 int dtype, void* stream) {
  if (C % 8 != 0) return -1;
  const PoolGeom g = geom(N, H, W, C, Ho, Wo, 0, 0, 0, 0, 0, 0);
  const long long total = (long long)N * (backward ? (long long)H * W : (long long)Ho * Wo) * (C / 8);
  cudaStream_t st = (cudaStream_t)stream;
  DISPATCH_T(dtype, if (backward) adaptive_avg_kernel<T, true><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g);
             else adaptive_avg_kernel<T, false><<<grid_for(total, 256), 256, 0, st>>>((const T*)in, (T*)out, g);)
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_relu_fwd(const void* x, void* y, long long n, int dtype, void* stream) {
  if (n % 8 != 0) return -1;
  DISPATCH_T(dtype, relu_fwd_kernel<T><<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n / 8);)
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_relu_bwd(const void* dy, const void* y, void* dx, long long n, int dtype, void* stream) {
  if (n % 8 != 0) return -1;
  DISPATCH_T(dtype, relu_bwd_kernel<T><<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const T*)dy, (const T*)y, (T*)dx, n / 8);)
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_dropout_ex(const void* x, void* y, long long n, unsigned int seed, const long long* epoch, float p,
                           int dtype, void* stream) {
  if (n % 8 != 0) return -1;
  const double t = (double)p * 4294967296.0;
  const uint32_t thresh = t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
  const float scale = p < 1.f ? 1.f / (1.f - p) : 0.f;
  DISPATCH_T(dtype, dropout_kernel<T><<<grid_for(n / 8, 256), 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n / 8, seed, thresh, scale, epoch);)
  MPB_CHECK_LAUNCH();
  return 0;
}
MPB_API int mpb_dropout(const void* x, void* y, long long n, unsigned int seed, float p, int dtype, void* stream) {
  return mpb_dropout_ex(x, y, n, seed, nullptr, p, dtype, stream);
}

// ------------------------------------------------------------------------------------------------------------
// Row-strided copy: M rows of `row_bytes` (a multiple of 16) from a source with row pitch src_pitch bytes to a
// destination with row pitch dst_pitch bytes.  This is channel concatenation (write a tensor into its channel slice
// of the NHWC output) and its inverse (gather a channel slice of a gradient into a dense NHWC tensor) at 16 bytes
// per thread -- the library's generic strided-copy kernels that torch.cat / .contiguous() use for these views run
// at a quarter of the copy rate.  (torch.cat at models.py DenseNet / Inception / Fire blocks.)
// ------------------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) copy_rows_kernel(const uint8_t* __restrict__ src, long long src_pitch,
                                                        uint8_t* __restrict__ dst, long long dst_pitch, long long M,
                                                        int vecs_per_row) {
  const long long total = M * vecs_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / vecs_per_row;
    const int v = (int)(i - row * vecs_per_row);
    const uint4 u = *reinterpret_cast<const uint4*>(src + row * src_pitch + (long long)v * 16);
    *reinterpret_cast<uint4*>(dst + row * dst_pitch + (long long)v * 16) = u;
  }
}
}  // namespace

MPB_API int mpb_copy_rows(const void* src, long long src_pitch, void* dst, long long dst_pitch, long long M,
                          long long row_bytes, void* stream) {
  if (row_bytes % 16 != 0 || src_pitch % 16 != 0 || dst_pitch % 16 != 0 || ((uintptr_t)src & 15) || ((uintptr_t)dst & 15))
    return -1;
  if (M <= 0 || row_bytes == 0) return 0;
  const int vecs = (int)(row_bytes / 16);
  copy_rows_kernel<<<grid_for(M * vecs, 256, 16), 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src, src_pitch, (uint8_t*)dst,
                                                                                  dst_pitch, M, vecs);
  MPB_CHECK_LAUNCH();
  return 0;
}


// Column gather of a small row-major matrix: dst[r][j] (+)= src[r][idx[j]] for idx[j] >= 0, else 0.  Used to move
// the 7x7x3 stem filter between its KRSC layout and the space-to-depth layout the tensor-core stem kernels read
// (forward: bf16 compute copy -> [K][4 taps][64]; backward: fp32 [K][256] filter gradient -> += the parameter's
// slot of the gradient arena), replacing torch.cat + advanced indexing (+ a zeros fill and an add on the way back).
namespace {
template <typename TI, typename TO>
__global__ void gather_cols_kernel(const TI* __restrict__ src, const long long* __restrict__ idx, TO* __restrict__ dst,
                                   int rows, int src_cols, int dst_cols, int accumulate) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= (long long)rows * dst_cols) return;
  const int r = (int)(i / dst_cols), j = (int)(i % dst_cols);
  const long long k = idx[j];
  float v = (k >= 0 && k < src_cols) ? to_f(src[(long long)r * src_cols + k]) : 0.f;
  if (accumulate) v += to_f(dst[i]);
  dst[i] = from_f<TO>(v);
}
}  // namespace

// dtype codes: 0 = fp32, 1 = bf16 (as everywhere in this library)
MPB_API int mpb_gather_cols(const void* src, const long long* idx, void* dst, int rows, int src_cols, int dst_cols,
                            int src_dtype, int dst_dtype, int accumulate, void* stream) {
  if (rows <= 0 || dst_cols <= 0) return 0;
  const long long n = (long long)rows * dst_cols;
  const int blocks = (int)((n + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (src_dtype == 1 && dst_dtype == 1)
    gather_cols_kernel<__nv_bfloat16, __nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)src, idx, (__nv_bfloat16*)dst, rows, src_cols, dst_cols, accumulate);
  else if (src_dtype == 0 && dst_dtype == 1)
    gather_cols_kernel<float, __nv_bfloat16><<<blocks, 256, 0, st>>>((const float*)src, idx, (__nv_bfloat16*)dst, rows, src_cols, dst_cols, accumulate);
  else if (src_dtype == 0 && dst_dtype == 0)
    gather_cols_kernel<float, float><<<blocks, 256, 0, st>>>((const float*)src, idx, (float*)dst, rows, src_cols, dst_cols, accumulate);
  else
    return -1;
  MPB_CHECK_LAUNCH();
  return 0;
}
