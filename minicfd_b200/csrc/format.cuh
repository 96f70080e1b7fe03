// Text of a VTK DataArray on the device: `for (v : values) vtu << v << ' ';` (mesh/vtk.hpp:600-602, 610-612, 622-624)
// with the stream's default format, i.e. printf's "%g " per value, and `vtu << i << ' '` for the integer arrays
// (connectivity, offsets, types; vtk.hpp:650-665).  At BASELINE config 2 one frame is 10 M numbers: formatting them on
// the host costs two orders of magnitude more than the Runge-Kutta steps between two frames, so the tokens are built
// here and only bytes travel.
//
// "%g" = six significant digits, correctly rounded (ties to even) from the exact binary value.  FormatG6 scales the
// value by a power of ten held to 2^-64 (a double pair) with an FMA-exact product, which decides the rounding unless
// the scaled value sits within 1e-6 of a half-integer (true ties such as 12345.25 live there) or the power of ten
// leaves the table; those few values are flagged and formatted by the host's snprintf before the tokens are packed.
#ifndef MINICFD_B200_FORMAT_CUH_
#define MINICFD_B200_FORMAT_CUH_

#include <cmath>
#include <cstdint>
#include <cstring>

#ifndef MCFD_HD
#ifdef __CUDACC__
#define MCFD_HD __host__ __device__ __forceinline__
#else
#define MCFD_HD inline
#endif
#endif

namespace mcfd {

constexpr int kPow10Min = -300, kPow10Max = 300;   // table: 10^k as hi + lo, k in [kPow10Min, kPow10Max]
constexpr int kTokenSlot = 16;                      // bytes per formatted token before packing (longest: 14)

struct Token {
  char c[kTokenSlot];
  int len;
  bool unsure;   // the host must format this one
};

// round(|v| * 10^shift) as an integer-valued double; *unsure when the fraction is too close to 1/2 to call
MCFD_HD double ScaledRound(double a, int shift, const double *pow10, bool *unsure) {
  const double ph = pow10[2 * (shift - kPow10Min)], pl = pow10[2 * (shift - kPow10Min) + 1];
  const double sh = a * ph;
  const double sl = fma(a, ph, -sh) + a * pl;
  const double r = floor(sh);
  const double f = (sh - r) + sl;
  if (fabs(f - 0.5) < 1e-6) *unsure = true;
  return f < 0.5 ? r : r + 1;
}

MCFD_HD bool SignBit(double v) {   // also of zeros and NaNs
  unsigned long long bits;
  memcpy(&bits, &v, sizeof bits);
  return (bits >> 63) != 0;
}

MCFD_HD Token FormatG6(double v, const double *pow10) {
  Token t;
  t.len = 0;
  t.unsure = false;
  int n = 0;
  if (SignBit(v)) t.c[n++] = '-';
  const double a = fabs(v);
  if (a != a) {            // glibc prints the sign of a NaN, too
    t.c[n++] = 'n'; t.c[n++] = 'a'; t.c[n++] = 'n';
  } else if (a > 1.7976931348623157e308) {
    t.c[n++] = 'i'; t.c[n++] = 'n'; t.c[n++] = 'f';
  } else if (a == 0) {
    t.c[n++] = '0';
  } else {
    int x = (int)floor(log10(a));        // decimal exponent, possibly off by one: corrected below
    if (5 - x < kPow10Min + 1 || 5 - x > kPow10Max - 1) {
      t.unsure = true;                   // beyond the table (|v| < 1e-294 or > 1e+304): the host formats it
      t.c[n++] = '0';
    } else {
      double d = ScaledRound(a, 5 - x, pow10, &t.unsure);
      if (d < 100000.0) { --x; d = ScaledRound(a, 5 - x, pow10, &t.unsure); }
      else if (d >= 1000000.0) { ++x; d = ScaledRound(a, 5 - x, pow10, &t.unsure); }
      int digits = (int)d, dg[6], nd = 6;
      for (int i = 5; i >= 0; --i) { dg[i] = digits % 10; digits /= 10; }
      while (nd > 1 && dg[nd - 1] == 0) --nd;              // %g drops trailing zeros
      if (x >= -4 && x < 6) {                               // fixed notation, precision 5 - x
        if (x >= 0) {
          for (int i = 0; i <= x; ++i) t.c[n++] = (char)('0' + dg[i]);
          if (nd > x + 1) {
            t.c[n++] = '.';
            for (int i = x + 1; i < nd; ++i) t.c[n++] = (char)('0' + dg[i]);
          }
        } else {
          t.c[n++] = '0'; t.c[n++] = '.';
          for (int i = 0; i < -x - 1; ++i) t.c[n++] = '0';
          for (int i = 0; i < nd; ++i) t.c[n++] = (char)('0' + dg[i]);
        }
      } else {                                              // d[.ddddd]e+XX
        t.c[n++] = (char)('0' + dg[0]);
        if (nd > 1) {
          t.c[n++] = '.';
          for (int i = 1; i < nd; ++i) t.c[n++] = (char)('0' + dg[i]);
        }
        t.c[n++] = 'e';
        t.c[n++] = x < 0 ? '-' : '+';
        const int e = x < 0 ? -x : x;
        if (e >= 100) t.c[n++] = (char)('0' + e / 100);
        t.c[n++] = (char)('0' + (e / 10) % 10);
        t.c[n++] = (char)('0' + e % 10);
      }
    }
  }
  t.c[n++] = ' ';
  t.len = n;
  for (int i = n; i < kTokenSlot; ++i) t.c[i] = 0;
  return t;
}

MCFD_HD Token FormatInteger(long long v) {   // operator<<(integer) and a blank
  Token t;
  t.unsure = false;
  char tmp[24];
  int m = 0, n = 0;
  unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
  do { tmp[m++] = (char)('0' + u % 10); u /= 10; } while (u && m < 14);   // mcfd_format_integers refuses 10^14 and more
  if (v < 0) t.c[n++] = '-';
  while (m) t.c[n++] = tmp[--m];
  t.c[n++] = ' ';
  t.len = n;
  for (int i = n; i < kTokenSlot; ++i) t.c[i] = 0;
  return t;
}

// 10^k as a double pair (hi + lo carries the host's 64-bit long double mantissa), k = kPow10Min..kPow10Max
inline void BuildPow10Table(double *table) {
  for (int k = kPow10Min; k <= kPow10Max; ++k) {
    const long double p = powl(10.0L, (long double)k);
    const double hi = (double)p;
    table[2 * (k - kPow10Min)] = hi;
    table[2 * (k - kPow10Min) + 1] = (double)(p - (long double)hi);
  }
}

#ifdef __CUDACC__
struct FormatArgs {
  const double *values;      // kind 0: values[i * stride]
  long long stride, n;
  long long first, step;     // kind 1: first + i * step
  int kind;
  const double *pow10;
  uint4 *slots;              // [n] 16-byte tokens
  unsigned char *lens;       // [n]
  unsigned int *unsure_count;
  unsigned int *unsure_index;  // capacity unsure_capacity
  unsigned int unsure_capacity;
};

__global__ void __launch_bounds__(256) FormatTokensKernel(FormatArgs p) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= p.n) return;
  Token t;
  if (p.kind == 0) t = FormatG6(p.values[i * p.stride], p.pow10);
  else t = FormatInteger(p.first + i * p.step);
  uint4 w;
  memcpy(&w, t.c, 16);
  p.slots[i] = w;
  p.lens[i] = (unsigned char)t.len;
  if (t.unsure) {
    const unsigned int at = atomicAdd(p.unsure_count, 1u);
    if (at < p.unsure_capacity) p.unsure_index[at] = (unsigned int)i;
  }
}

constexpr int kPackPerBlock = 1024;   // tokens per CTA of the packing kernels (256 threads x 4)

__global__ void __launch_bounds__(256) TokenBlockSumsKernel(const unsigned char *lens, long long n, unsigned int *sums) {
  __shared__ unsigned int warp_sum[8];
  const long long base = (long long)blockIdx.x * kPackPerBlock + threadIdx.x * 4;
  unsigned int s = 0;
  for (int k = 0; k < 4; ++k) if (base + k < n) s += lens[base + k];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int total = 0;
    for (int w = 0; w < 8; ++w) total += warp_sum[w];
    sums[blockIdx.x] = total;
  }
}

// exclusive scan of the block sums by one CTA (a frame has ~10^4 of them); offsets[n_blocks] = total length
__global__ void __launch_bounds__(1024) TokenScanKernel(const unsigned int *sums, int n_blocks, unsigned long long *offsets) {
  __shared__ unsigned long long part[1024];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < n_blocks; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const unsigned long long mine = i < n_blocks ? sums[i] : 0;
    part[threadIdx.x] = mine;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      const unsigned long long add = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
      __syncthreads();
      part[threadIdx.x] += add;
      __syncthreads();
    }
    if (i < n_blocks) offsets[i] = carry + part[threadIdx.x] - mine;
    __syncthreads();
    if (threadIdx.x == 1023) carry += part[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[n_blocks] = carry;
}

__global__ void __launch_bounds__(256) TokenPackKernel(const uint4 *slots, const unsigned char *lens, long long n,
    const unsigned long long *offsets, char *text) {
  __shared__ char sText[kPackPerBlock * kTokenSlot];
  __shared__ unsigned int warp_sum[8];
  const long long base = (long long)blockIdx.x * kPackPerBlock + threadIdx.x * 4;
  unsigned int len[4], mine = 0;
  for (int k = 0; k < 4; ++k) { len[k] = base + k < n ? lens[base + k] : 0; mine += len[k]; }
  unsigned int incl = mine;                        // inclusive scan over the CTA's threads
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) warp_sum[warp] = incl;
  __syncthreads();
  unsigned int before = 0, total = 0;
  for (int w = 0; w < 8; ++w) { if (w < warp) before += warp_sum[w]; total += warp_sum[w]; }
  unsigned int at = before + incl - mine;
  for (int k = 0; k < 4; ++k) {
    if (!len[k]) continue;
    uint4 w = slots[base + k];
    char c[16];
    memcpy(c, &w, 16);
    for (unsigned int j = 0; j < len[k]; ++j) sText[at + j] = c[j];
    at += len[k];
  }
  __syncthreads();
  char *dst = text + offsets[blockIdx.x];
  for (unsigned int j = threadIdx.x; j < total; j += 256) dst[j] = sText[j];
}
#endif  // __CUDACC__

}  // namespace mcfd
#endif  // MINICFD_B200_FORMAT_CUH_
