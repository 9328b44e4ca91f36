// text_parse.inl -- one text field -> binary32 bits, exactly as sscanf("%f") + "no trailing byte"
// (c++/utils/impl/StringUtils.cc:49-56).  Included by kernels_text.cu (device) and by
// tests/native/text_parse_host.cu (the same source compiled for the host, so that the CPU test
// suite can fuzz it against the C library without a GPU).  Not a CPU path of the product: the
// library only ever calls it from text_parse_kernel.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDACC__)
#define HGM_HD __host__ __device__
#else
#define HGM_HD
#endif

HGM_HD inline int hgm_clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
HGM_HD inline int hgm_clz64(unsigned long long x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return x ? __builtin_clzll(x) : 64;
#endif
}
HGM_HD inline int hgm_ctz64(unsigned long long x) {
#if defined(__CUDA_ARCH__)
  return __ffsll((long long)x) - 1;
#else
  return __builtin_ctzll(x);
#endif
}
HGM_HD inline unsigned long long hgm_d2u(double v) {
#if defined(__CUDA_ARCH__)
  return (unsigned long long)__double_as_longlong(v);
#else
  unsigned long long u;
  memcpy(&u, &v, 8);
  return u;
#endif
}
HGM_HD inline double hgm_u2d(unsigned long long u) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)u);
#else
  double v;
  memcpy(&v, &u, 8);
  return v;
#endif
}

#define HGM_POW10_LIST                                                                                        \
  1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, \
      1e20, 1e21, 1e22
#if defined(__CUDACC__)
__constant__ double hgm_pow10_dev[23] = {HGM_POW10_LIST};
#endif
static const double hgm_pow10_host[23] = {HGM_POW10_LIST};
#if defined(__CUDA_ARCH__)
#define HGM_POW10(i) hgm_pow10_dev[i]
#else
#define HGM_POW10(i) hgm_pow10_host[i]
#endif

namespace hgm_text {

// ---- exact decimal / hexadecimal -> binary32 --------------------------------------------------

struct Big {  // little-endian magnitude, 20 limbs = 640 bits (|q| <= 173 and <= 127 digits keep both sides under 450 bits)
  static constexpr int kLimbs = 20;
  uint32_t v[kLimbs];
  int n;
  bool overflow;
};
HGM_HD void big_set(Big &a, unsigned long long x) {
  a.n = 0;
  a.overflow = false;
  for (int i = 0; i < Big::kLimbs; ++i) a.v[i] = 0;
  a.v[0] = (uint32_t)x;
  a.v[1] = (uint32_t)(x >> 32);
  a.n = a.v[1] ? 2 : (a.v[0] ? 1 : 0);
}
HGM_HD void big_mul_add(Big &a, uint32_t m, uint32_t add) {
  unsigned long long carry = add;
  for (int i = 0; i < a.n; ++i) {
    const unsigned long long t = (unsigned long long)a.v[i] * m + carry;
    a.v[i] = (uint32_t)t;
    carry = t >> 32;
  }
  if (carry) {
    if (a.n < Big::kLimbs) a.v[a.n++] = (uint32_t)carry;
    else a.overflow = true;
  }
}
HGM_HD void big_shl(Big &a, int bits) {  // the caller has checked that the result fits
  if (a.n == 0 || bits <= 0) return;
  const int words = bits >> 5, rem = bits & 31;
  for (int i = Big::kLimbs - 1; i >= 0; --i) {  // limbs at and above a.n are zero by construction
    const int src = i - words;
    const uint32_t cur = src >= 0 ? a.v[src] : 0u;
    const uint32_t below = src >= 1 ? a.v[src - 1] : 0u;
    a.v[i] = rem ? ((cur << rem) | (below >> (32 - rem))) : cur;
  }
  int n = Big::kLimbs;
  while (n > 0 && a.v[n - 1] == 0) --n;
  a.n = n;
}
HGM_HD int big_bits(const Big &a) { return a.n == 0 ? 0 : 32 * (a.n - 1) + (32 - hgm_clz32(a.v[a.n - 1])); }
HGM_HD int big_cmp(const Big &a, const Big &b) {
  if (a.n != b.n) return a.n < b.n ? -1 : 1;
  for (int i = a.n - 1; i >= 0; --i)
    if (a.v[i] != b.v[i]) return a.v[i] < b.v[i] ? -1 : 1;
  return 0;
}
HGM_HD void big_mul_pow5(Big &a, int e) {
  while (e >= 13) { big_mul_add(a, 1220703125u, 0); e -= 13; }  // 5^13
  uint32_t m = 1;
  while (e-- > 0) m *= 5;
  if (m > 1) big_mul_add(a, m, 0);
}
HGM_HD void big_shl_checked(Big &a, int bits) {
  const int before = big_bits(a);
  if (before + bits > 32 * Big::kLimbs) { a.overflow = true; return; }
  big_shl(a, bits);
}



HGM_HD inline bool dig(unsigned c) { return c - '0' < 10u; }
HGM_HD inline int hexval(unsigned c) {
  if (c - '0' < 10u) return (int)(c - '0');
  const unsigned l = c | 0x20;
  if (l - 'a' < 6u) return (int)(l - 'a' + 10);
  return -1;
}

// optional sign and digits of an exponent, saturated; digits may be absent ("1e", "0x1p+": scanf
// hands them to strtof, which ignores the dangling marker)
HGM_HD int scan_exponent(const unsigned char *s, int len, int *pos) {
  int i = *pos;
  bool neg = false;
  if (i < len && (s[i] == '+' || s[i] == '-')) { neg = s[i] == '-'; ++i; }
  int e = 0;
  while (i < len && dig(s[i])) {
    if (e < 100000) e = e * 10 + (int)(s[i] - '0');
    ++i;
  }
  *pos = i;
  return neg ? -e : e;
}

// m * 2^e2 (+ sticky: a non-zero tail below m's last bit) -> binary32 magnitude bits, RN-even
HGM_HD uint32_t round_binary(unsigned long long m, int e2, bool sticky) {
  if (m == 0) return 0u;
  const int L = 64 - hgm_clz64(m);
  int E = L - 1 + e2;
  if (E > 127) return 0x7F800000u;
  if (E < -151) return 0u;  // below 2^-150 (E == -151 would need m >= 2^L, impossible)
  const bool normal = E >= -126;
  // how many low bits of m fall below the result's last place
  const int shift = normal ? L - 24 : -(e2 + 149);
  unsigned long long mant;
  if (shift <= 0) {
    mant = m << (-shift);
  } else if (shift >= 64) {
    mant = 0;  // m < 2^63 < half an ulp only if shift > L; with E >= -151 shift <= L+2 <= 62, not reached
  } else {
    mant = m >> shift;
    const unsigned long long rem = m & ((1ull << shift) - 1), half = 1ull << (shift - 1);
    if (rem > half || (rem == half && (sticky || (mant & 1)))) ++mant;
  }
  if (!normal) return (uint32_t)mant;  // mant == 2^23 is the smallest normal: the encoding agrees
  if (mant == (1ull << 24)) { mant >>= 1; ++E; }
  if (E > 127) return 0x7F800000u;
  return ((uint32_t)(E + 127) << 23) | ((uint32_t)mant & 0x7FFFFFu);
}

// exact sign of (W * 10^q - Hm * 2^t), W = all significant digits of the field
HGM_HD int exact_compare(const unsigned char *s, int len, int q, unsigned long long Hm, int t, bool *trouble) {
  Big lhs, rhs;
  big_set(lhs, 0);
  {  // W: digits in chunks of 9
    uint32_t chunk = 0, scale = 1;
    for (int i = 0; i < len; ++i) {
      const unsigned c = s[i];
      if (dig(c)) {
        chunk = chunk * 10 + (c - '0');
        scale *= 10;
        if (scale == 1000000000u) { big_mul_add(lhs, scale, chunk); chunk = 0; scale = 1; }
      } else if (c != '.') {
        break;  // exponent marker: mantissa is over
      }
    }
    if (scale > 1) big_mul_add(lhs, scale, chunk);
  }
  big_set(rhs, Hm);
  const int a = q > 0 ? q : 0, b = q < 0 ? -q : 0;
  big_mul_pow5(lhs, a);
  big_mul_pow5(rhs, b);
  int l2 = a, r2 = b + t;
  if (r2 < 0) { l2 -= r2; r2 = 0; }
  const int common = l2 < r2 ? l2 : r2;
  big_shl_checked(lhs, l2 - common);
  big_shl_checked(rhs, r2 - common);
  if (lhs.overflow || rhs.overflow) *trouble = true;
  return big_cmp(lhs, rhs);
}

// One field (no separators inside) -> binary32 bits.  false = scanf("%f%c") would not have
// returned exactly one conversion.
HGM_HD bool parse_field(const unsigned char *s, int len, uint32_t *bits_out, bool *trouble) {
  int i = 0;
  while (i < len && (s[i] == '\v' || s[i] == '\f')) ++i;  // scanf skips leading white space
  uint32_t sign = 0;
  if (i < len && (s[i] == '+' || s[i] == '-')) { sign = s[i] == '-' ? 0x80000000u : 0u; ++i; }
  if (i >= len) return false;
  const unsigned c0 = s[i] | 0x20u;
  if (c0 == 'n') {
    if (len - i == 3 && (s[i + 1] | 0x20u) == 'a' && (s[i + 2] | 0x20u) == 'n') { *bits_out = sign | 0x7FC00000u; return true; }
    return false;
  }
  if (c0 == 'i') {
    const char *w = "infinity";
    const int m = len - i;
    if (m != 3 && m != 8) return false;
    for (int j = 0; j < m; ++j)
      if ((s[i + j] | 0x20u) != (unsigned)w[j]) return false;
    *bits_out = sign | 0x7F800000u;
    return true;
  }
  if (s[i] == '0' && i + 1 < len && (s[i + 1] | 0x20u) == 'x') {  // hexadecimal literal
    i += 2;
    if (i >= len) return false;  // a bare "0x"
    unsigned long long m = 0;
    int nh = 0, e2 = 0;
    bool sticky = false, any = false, dot = false;
    for (; i < len; ++i) {
      const int h = hexval(s[i]);
      if (h < 0) break;
      any = true;
      if (m == 0 && h == 0) continue;
      if (nh < 15) { m = m * 16 + (unsigned)h; ++nh; }
      else { e2 += 4; sticky = sticky || h != 0; }
    }
    if (i < len && s[i] == '.') {
      dot = true;
      for (++i; i < len; ++i) {
        const int h = hexval(s[i]);
        if (h < 0) break;
        any = true;
        if (m == 0 && h == 0) { e2 -= 4; continue; }
        if (nh < 15) { m = m * 16 + (unsigned)h; ++nh; e2 -= 4; }
        else sticky = sticky || h != 0;
      }
    }
    if (!any && !dot) return false;
    if (any && i < len && (s[i] | 0x20u) == 'p') { ++i; e2 += scan_exponent(s, len, &i); }
    if (i != len) return false;
    *bits_out = sign | round_binary(m, e2, sticky);
    return true;
  }
  // decimal literal
  const int mant_begin = i;
  unsigned long long w = 0;
  int nd = 0, exp10 = 0, frac_digits = 0;
  bool any = false;
  for (; i < len && dig(s[i]); ++i) {
    any = true;
    const unsigned dgt = s[i] - '0';
    if (nd == 0 && dgt == 0) continue;
    if (nd < 19) { w = w * 10 + dgt; ++nd; }
    else ++exp10;
  }
  if (i < len && s[i] == '.') {
    for (++i; i < len && dig(s[i]); ++i) {
      any = true;
      ++frac_digits;
      const unsigned dgt = s[i] - '0';
      if (nd == 0 && dgt == 0) { --exp10; continue; }
      if (nd < 19) { w = w * 10 + dgt; ++nd; --exp10; }
    }
  }
  if (!any) return false;
  const int mant_end = i;
  int e_explicit = 0;
  if (i < len && (s[i] | 0x20u) == 'e') { ++i; e_explicit = scan_exponent(s, len, &i); }
  if (i != len) return false;
  if (w == 0) { *bits_out = sign; return true; }
  exp10 += e_explicit;
  const int dexp = exp10 + nd;  // value in [10^(dexp-1), 10^dexp)
  if (dexp > 39) { *bits_out = sign | 0x7F800000u; return true; }
  if (dexp < -46) { *bits_out = sign; return true; }  // < 1e-46 < 2^-150
  // approximation in binary64: relative error < 2^-49 (<= 3 roundings, <= 1e-18 from dropped digits)
  double v = (double)w;
  {
    int e = exp10;
    while (e > 0) { const int step = e > 22 ? 22 : e; v *= HGM_POW10(step); e -= step; }
    while (e < 0) { const int step = -e > 22 ? 22 : -e; v /= HGM_POW10(step); e += step; }
  }
  // the two binary32 neighbours of v, as exact doubles
  double f_lo, ulp;
  if (v >= 0x1p-126) {
    const unsigned long long vb = hgm_d2u(v);
    f_lo = hgm_u2d(vb & ~((1ull << 29) - 1));
    ulp = hgm_u2d(((vb >> 52) - 23) << 52);
  } else {
    f_lo = floor(v * 0x1p149) * 0x1p-149;
    ulp = 0x1p-149;
  }
  const double half = f_lo + 0.5 * ulp;  // exact: 25 significant bits
  int pick_hi;
  if (fabs(v - half) > v * 0x1p-46) {
    pick_hi = v > half;
  } else {
    // too close to call in binary64: decide with exact integers
    const unsigned long long hb = hgm_d2u(half);
    unsigned long long Hm = (hb & ((1ull << 52) - 1)) | (1ull << 52);
    int t = (int)(hb >> 52) - 1075;
    const int tz = hgm_ctz64(Hm);
    Hm >>= tz;
    t += tz;
    const int q_full = e_explicit - frac_digits;
    const int cmp = exact_compare(s + mant_begin, mant_end - mant_begin, q_full, Hm, t, trouble);
    if (cmp != 0) pick_hi = cmp > 0;
    else {  // a tie: the neighbour with the even significand
      const double steps = f_lo / ulp;  // exact, < 2^25
      pick_hi = ((unsigned long long)steps) & 1ull;
    }
  }
  const double f = pick_hi ? f_lo + ulp : f_lo;
  uint32_t mag;
  if (f >= 0x1p128) mag = 0x7F800000u;
  else if (f >= 0x1p-126) {
    const unsigned long long fb = hgm_d2u(f);
    mag = ((uint32_t)((fb >> 52) - 1023 + 127) << 23) | (uint32_t)((fb >> 29) & 0x7FFFFFu);
  } else {
    mag = (uint32_t)(f * 0x1p149);
  }
  *bits_out = sign | mag;
  return true;
}

}  // namespace hgm_text
