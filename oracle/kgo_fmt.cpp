// kgo_fmt.cpp — ORACLE (test infrastructure): sample-value and label encoding.
//
// Restates, for the CPU, what the reference's /metrics handler (main.go:69-71)
// reaches in third-party code that is not vendored under /root/reference:
//   * prometheus/common v0.55.0 expfmt/text_create.go writeFloat and
//     writeEscapedString (go.mod:36)                      -> rules F1, G2
//   * Go 1.23 strconv.AppendFloat(f,'g',-1,64) (go.mod:3)  -> rules F2-F5
// The shortest digits (F2) come from a Schubfach-style algorithm (R. Giulietti,
// "The Schubfach way to render doubles"): any correct shortest-round-trip
// printer yields the same digits as Go's, which is what the tests pin against
// CPython's repr().  The 128-bit powers of ten are derived here at start-up
// with a small exact big-integer, independently of the table the CUDA side
// compiles in, so the two can be cross-checked.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <mutex>
#include "kgo.h"

namespace {

struct U128 { uint64_t hi, lo; };
constexpr int kMinK = -292, kMaxK = 324;
U128 g_pow10[kMaxK - kMinK + 1];
std::once_flag g_pow10_once;

// ---- tiny exact big integer (little-endian u32 limbs) ----------------------
using Big = std::vector<uint32_t>;
void big_trim(Big& a) { while (a.size() > 1 && a.back() == 0) a.pop_back(); }
void big_mul_small(Big& a, uint32_t m) {
  uint64_t carry = 0;
  for (auto& limb : a) { uint64_t t = uint64_t(limb) * m + carry; limb = uint32_t(t); carry = t >> 32; }
  if (carry) a.push_back(uint32_t(carry));
}
int big_bitlen(const Big& a) {
  int n = int(a.size());
  while (n > 1 && a[n - 1] == 0) --n;
  uint32_t top = a[n - 1];
  int b = 0; while (top) { ++b; top >>= 1; }
  return (n - 1) * 32 + b;
}
bool big_bit(const Big& a, int i) { size_t w = size_t(i) / 32; return w < a.size() && ((a[w] >> (i % 32)) & 1); }
int big_cmp(const Big& a, const Big& b) {
  size_t n = a.size() > b.size() ? a.size() : b.size();
  for (size_t i = n; i-- > 0;) {
    uint32_t x = i < a.size() ? a[i] : 0, y = i < b.size() ? b[i] : 0;
    if (x != y) return x < y ? -1 : 1;
  }
  return 0;
}
void big_sub(Big& a, const Big& b) {  // a -= b, a >= b
  int64_t borrow = 0;
  for (size_t i = 0; i < a.size(); ++i) {
    int64_t t = int64_t(a[i]) - (i < b.size() ? b[i] : 0) - borrow;
    borrow = t < 0; if (t < 0) t += (int64_t(1) << 32);
    a[i] = uint32_t(t);
  }
}
void big_shl1(Big& a) {
  uint32_t carry = 0;
  for (auto& limb : a) { uint32_t nc = limb >> 31; limb = (limb << 1) | carry; carry = nc; }
  if (carry) a.push_back(carry);
}

// g(k) = ceil(10^k * 2^-r), r = floor(log2 10^k) - 127, so 2^127 <= g < 2^128.
U128 derive_pow10(int k) {
  U128 out{0, 0};
  if (k >= 0) {
    Big n{1};
    for (int i = 0; i < k; ++i) big_mul_small(n, 10);
    int bits = big_bitlen(n);
    bool sticky = false;
    for (int i = 0; i < 128; ++i) {  // take the top 128 bits
      int src = bits - 1 - i;
      bool b = src >= 0 && big_bit(n, src);
      if (i < 64) out.hi |= uint64_t(b) << (63 - i); else out.lo |= uint64_t(b) << (127 - i);
    }
    for (int i = 0; i < bits - 128; ++i) sticky |= big_bit(n, i);
    if (sticky) { if (++out.lo == 0) ++out.hi; }
  } else {
    Big d{1};
    for (int i = 0; i < -k; ++i) big_mul_small(d, 10);
    int s = big_bitlen(d) + 127;  // floor(2^s / d) has exactly 128 bits
    Big r{1};                      // bit s of the dividend
    // 2^s < d*2^128 and 2^s >= d*2^127: the first quotient bit appears after
    // shifting in (s - 127) bits; run the plain restoring division.
    uint64_t hi = 0, lo = 0;
    for (int i = s - 1; i >= 0; --i) {
      big_shl1(r);
      int bit = 0;
      if (big_cmp(r, d) >= 0) { big_sub(r, d); big_trim(r); bit = 1; }
      hi = (hi << 1) | (lo >> 63); lo = (lo << 1) | uint64_t(bit);
    }
    // d has a factor 5, so the division is never exact: ceil = floor + 1
    if (++lo == 0) ++hi;
    out.hi = hi; out.lo = lo;
  }
  return out;
}
void init_pow10() { for (int k = kMinK; k <= kMaxK; ++k) g_pow10[k - kMinK] = derive_pow10(k); }
const U128& pow10(int k) { std::call_once(g_pow10_once, init_pow10); return g_pow10[k - kMinK]; }

inline int floor_log2_pow10(int e) { return (e * 1741647) >> 19; }
inline int floor_log10_pow2(int e) { return (e * 1262611) >> 22; }
inline int floor_log10_three_quarters_pow2(int e) { return (e * 1262611 - 524031) >> 22; }

inline uint64_t round_to_odd(const U128& g, uint64_t cp) {
  unsigned __int128 x = (unsigned __int128)g.lo * cp;
  unsigned __int128 y = (unsigned __int128)g.hi * cp;
  uint64_t y0 = uint64_t(y), y1 = uint64_t(y >> 64), x1 = uint64_t(x >> 64);
  y0 += x1; y1 += (y0 < x1);
  return y1 | (y0 > 1);
}

// shortest decimal d * 10^k == v (v finite, > 0), d without trailing zeros
void shortest(double v, uint64_t* d_out, int* k_out) {
  uint64_t bits; std::memcpy(&bits, &v, 8);
  uint64_t frac = bits & ((uint64_t(1) << 52) - 1);
  int bexp = int((bits >> 52) & 0x7ff);
  uint64_t c; int q;
  if (bexp != 0) { c = frac | (uint64_t(1) << 52); q = bexp - 1075; } else { c = frac; q = -1074; }
  uint64_t d; int k;
  if (q <= 0 && q > -53 && (c & ((uint64_t(1) << -q) - 1)) == 0) {
    d = c >> -q; k = 0;  // an integer below 2^53: its own decimal is shortest
  } else {
    bool even = (c & 1) == 0;
    bool lower_closer = frac == 0 && bexp > 1;
    uint64_t cbl = 4 * c - 2 + lower_closer, cb = 4 * c, cbr = 4 * c + 2;
    k = lower_closer ? floor_log10_three_quarters_pow2(q) : floor_log10_pow2(q);
    int h = q + floor_log2_pow10(-k) + 1;
    const U128& g = pow10(-k);
    uint64_t vbl = round_to_odd(g, cbl << h), vb = round_to_odd(g, cb << h), vbr = round_to_odd(g, cbr << h);
    uint64_t lower = vbl + !even, upper = vbr - !even;
    uint64_t s = vb / 4;
    bool done = false;
    if (s >= 10) {
      uint64_t sp = s / 10;
      bool up_in = lower <= 40 * sp, wp_in = 40 * sp + 40 <= upper;
      if (up_in != wp_in) { d = sp + wp_in; k += 1; done = true; }
    }
    if (!done) {
      bool u_in = lower <= 4 * s, w_in = 4 * s + 4 <= upper;
      if (u_in != w_in) d = s + w_in;
      else { uint64_t mid = 4 * s + 2; d = s + (vb > mid || (vb == mid && (s & 1))); }
    }
  }
  while (d % 10 == 0) { d /= 10; ++k; }
  *d_out = d; *k_out = k;
}

}  // namespace

extern "C" void kgo_pow10_entry(int k, uint64_t* hi, uint64_t* lo) {
  const U128& g = pow10(k); *hi = g.hi; *lo = g.lo;
}

// expfmt writeFloat (F1) + strconv.AppendFloat(f,'g',-1,64) (F2-F5).
extern "C" int kgo_format_value(double v, char* out) {
  if (v == 1) { out[0] = '1'; return 1; }
  if (v == 0) { out[0] = '0'; return 1; }  // also -0
  if (v == -1) { out[0] = '-'; out[1] = '1'; return 2; }
  if (std::isnan(v)) { std::memcpy(out, "NaN", 3); return 3; }
  if (std::isinf(v)) { std::memcpy(out, v > 0 ? "+Inf" : "-Inf", 4); return 4; }
  int pos = 0;
  if (v < 0) { out[pos++] = '-'; v = -v; }
  uint64_t d; int k;
  shortest(v, &d, &k);
  char dig[20]; int n = 0;
  { char tmp[20]; while (d) { tmp[n++] = char('0' + d % 10); d /= 10; } for (int i = 0; i < n; ++i) dig[i] = tmp[n - 1 - i]; }
  int dp = n + k;   // value = 0.d1..dn * 10^dp
  int x = dp - 1;   // decimal exponent of d1
  if (x < -4 || x >= 6) {  // F3: %e form (eprec is 6 in shortest mode)
    out[pos++] = dig[0];
    if (n > 1) { out[pos++] = '.'; for (int i = 1; i < n; ++i) out[pos++] = dig[i]; }
    out[pos++] = 'e';
    out[pos++] = x < 0 ? '-' : '+';
    int ax = x < 0 ? -x : x;
    if (ax < 10) { out[pos++] = '0'; out[pos++] = char('0' + ax); }
    else if (ax < 100) { out[pos++] = char('0' + ax / 10); out[pos++] = char('0' + ax % 10); }
    else { out[pos++] = char('0' + ax / 100); out[pos++] = char('0' + (ax / 10) % 10); out[pos++] = char('0' + ax % 10); }
  } else if (dp <= 0) {  // F5
    out[pos++] = '0'; out[pos++] = '.';
    for (int i = 0; i < -dp; ++i) out[pos++] = '0';
    for (int i = 0; i < n; ++i) out[pos++] = dig[i];
  } else if (dp >= n) {
    for (int i = 0; i < n; ++i) out[pos++] = dig[i];
    for (int i = 0; i < dp - n; ++i) out[pos++] = '0';
  } else {
    for (int i = 0; i < dp; ++i) out[pos++] = dig[i];
    out[pos++] = '.';
    for (int i = dp; i < n; ++i) out[pos++] = dig[i];
  }
  return pos;
}

extern "C" void kgo_format_values(const double* v, size_t n, uint8_t* out32) {
  for (size_t i = 0; i < n; ++i) {
    uint8_t* rec = out32 + 32 * i;
    std::memset(rec, 0, 32);
    rec[31] = uint8_t(kgo_format_value(v[i], reinterpret_cast<char*>(rec)));
  }
}

// expfmt writeEscapedString: label values escape \ " \n ; help escapes \ \n (G2)
extern "C" size_t kgo_escape(const char* in, size_t n, int is_label, char* out, size_t cap) {
  size_t o = 0;
  auto put = [&](char c) { if (o < cap) out[o] = c; ++o; };
  for (size_t i = 0; i < n; ++i) {
    char c = in[i];
    if (c == '\\') { put('\\'); put('\\'); }
    else if (c == '\n') { put('\\'); put('n'); }
    else if (c == '"' && is_label) { put('\\'); put('"'); }
    else put(c);
  }
  return o;
}
