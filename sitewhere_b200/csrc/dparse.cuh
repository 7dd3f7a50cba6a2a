// dparse.cuh — decimal text -> IEEE double, bit-exact with java.lang.Double.parseDouble (correct
// rounding, round-half-even), for device and host.
//   1. Clinger fast path: <= 15 significant digits and |q| <= 22  (exact double mul/div)
//   2. Eisel-Lemire on the first 19 significant digits (128-bit truncated powers of five); if digits were
//      dropped, the answer is accepted only when w and w+1 round to the same double
//   3. exact fallback: compare the decimal with the half-way point between neighbouring doubles in
//      big-integer arithmetic and step to the correctly rounded neighbour
// No FMA contraction anywhere (the library is built with --fmad=false / -ffp-contract=off).
#pragma once
#include <stdint.h>

#include "pow5_table.cuh"

#ifdef __CUDACC__
#define SW_HD __host__ __device__
__device__ const uint64_t sw_pow5_dev[2 * 651] = {SW_POW5_ENTRIES};
#else
#define SW_HD
#endif
static const uint64_t sw_pow5_host[2 * 651] = {SW_POW5_ENTRIES};

SW_HD inline uint64_t sw_pow5(int idx) {
#ifdef __CUDA_ARCH__
    return sw_pow5_dev[idx];
#else
    return sw_pow5_host[idx];
#endif
}
SW_HD inline void sw_mul64(uint64_t a, uint64_t b, uint64_t &hi, uint64_t &lo) {
#ifdef __CUDA_ARCH__
    lo = a * b;
    hi = __umul64hi(a, b);
#else
    unsigned __int128 p = (unsigned __int128)a * b;
    lo = (uint64_t)p;
    hi = (uint64_t)(p >> 64);
#endif
}
SW_HD inline int sw_clz64(uint64_t v) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)v);
#else
    return v ? __builtin_clzll(v) : 64;
#endif
}
SW_HD inline double sw_bits2d(uint64_t b) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    double d;
    __builtin_memcpy(&d, &b, 8);
    return d;
#endif
}
SW_HD inline uint64_t sw_d2bits(double d) {
#ifdef __CUDA_ARCH__
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t b;
    __builtin_memcpy(&b, &d, 8);
    return b;
#endif
}

// ---- Eisel-Lemire: value = w * 10^q, w != 0.  Returns false when it cannot decide. -----------------
SW_HD inline bool sw_eisel_lemire(uint64_t w, int64_t q, uint64_t &bits) {
    if (q < SW_POW5_MIN_Q) {
        bits = 0;
        return true;
    }
    if (q > SW_POW5_MAX_Q) {
        bits = 0x7ff0000000000000ull;
        return true;
    }
    int lz = sw_clz64(w);
    w <<= lz;
    const int idx = 2 * (int)(q - SW_POW5_MIN_Q);
    uint64_t hi, lo;
    sw_mul64(w, sw_pow5(idx), hi, lo);
    const uint64_t precision_mask = 0xFFFFFFFFFFFFFFFFull >> 55;  // mantissa bits + 3
    if ((hi & precision_mask) == precision_mask) {
        uint64_t hi2, lo2;
        sw_mul64(w, sw_pow5(idx + 1), hi2, lo2);
        lo += hi2;
        if (hi2 > lo) hi++;
    }
    if (lo == 0xFFFFFFFFFFFFFFFFull) {
        if (!(q >= -27 && q <= 55)) return false;
    }
    const int upperbit = (int)(hi >> 63);
    uint64_t mantissa = hi >> (upperbit + 64 - 52 - 3);
    int64_t power2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 1023;
    if (power2 <= 0) {  // subnormal
        if (-power2 + 1 >= 64) {
            bits = 0;
            return true;
        }
        mantissa >>= -power2 + 1;
        mantissa += (mantissa & 1);
        mantissa >>= 1;
        power2 = (mantissa < (1ull << 52)) ? 0 : 1;
        bits = (mantissa & ~(1ull << 52)) | ((uint64_t)power2 << 52);
        return true;
    }
    if (lo <= 1 && q >= -4 && q <= 23 && ((mantissa & 3) == 1)) {
        if ((mantissa << (upperbit + 64 - 52 - 3)) == hi) mantissa &= ~1ull;  // exactly half-way: to even
    }
    mantissa += (mantissa & 1);
    mantissa >>= 1;
    if (mantissa >= (2ull << 52)) {
        mantissa = 1ull << 52;
        power2++;
    }
    mantissa &= ~(1ull << 52);
    if (power2 >= 0x7FF) {
        bits = 0x7ff0000000000000ull;
        return true;
    }
    bits = mantissa | ((uint64_t)power2 << 52);
    return true;
}

// ---- exact fallback ---------------------------------------------------------------------------------------------
#define SW_BIG_WORDS 176  // 5632 bits: 768 digits * 5^1110-scale comparisons fit
struct SwBig {
    uint32_t w[SW_BIG_WORDS];
    int n;
    bool overflow;
};
SW_HD inline void sw_big_set(SwBig &a, uint64_t v) {
    a.n = 0;
    a.overflow = false;
    while (v) {
        a.w[a.n++] = (uint32_t)v;
        v >>= 32;
    }
}
SW_HD inline void sw_big_muladd(SwBig &a, uint32_t m, uint32_t add) {
    uint64_t carry = add;
    for (int i = 0; i < a.n; i++) {
        uint64_t t = (uint64_t)a.w[i] * m + carry;
        a.w[i] = (uint32_t)t;
        carry = t >> 32;
    }
    if (carry) {
        if (a.n < SW_BIG_WORDS)
            a.w[a.n++] = (uint32_t)carry;
        else
            a.overflow = true;
    }
}
SW_HD inline void sw_big_pow5(SwBig &a, uint32_t k) {
    while (k >= 13) {
        sw_big_muladd(a, 1220703125u, 0);  // 5^13
        k -= 13;
    }
    uint32_t m = 1;
    while (k--) m *= 5;
    if (m > 1) sw_big_muladd(a, m, 0);
}
SW_HD inline void sw_big_shl(SwBig &a, uint32_t bits) {
    if (a.n == 0 || bits == 0) return;
    const uint32_t ws = bits >> 5, bs = bits & 31;
    if (a.n + (int)ws + 1 > SW_BIG_WORDS) {
        a.overflow = true;
        return;
    }
    if (bs) {
        uint32_t carry = 0;
        for (int i = 0; i < a.n; i++) {
            uint32_t v = a.w[i];
            a.w[i] = (v << bs) | carry;
            carry = v >> (32 - bs);
        }
        if (carry) a.w[a.n++] = carry;
    }
    if (ws) {
        for (int i = a.n - 1; i >= 0; i--) a.w[i + ws] = a.w[i];
        for (uint32_t i = 0; i < ws; i++) a.w[i] = 0;
        a.n += ws;
    }
}
SW_HD inline int sw_big_cmp(const SwBig &a, const SwBig &b) {
    if (a.n != b.n) return a.n > b.n ? 1 : -1;
    for (int i = a.n - 1; i >= 0; i--)
        if (a.w[i] != b.w[i]) return a.w[i] > b.w[i] ? 1 : -1;
    return 0;
}

// compare  D * 10^e10 (+ something in (0,1)*10^e10 if sticky)  with  M * 2^p2
SW_HD inline int sw_cmp_dec_bin(const SwBig &D, int64_t e10, bool sticky, uint64_t M, int64_t p2, SwBig &A, SwBig &Bn) {
    A.n = D.n;
    A.overflow = false;
    for (int i = 0; i < D.n; i++) A.w[i] = D.w[i];
    sw_big_set(Bn, M);
    if (e10 >= 0)
        sw_big_pow5(A, (uint32_t)e10);
    else
        sw_big_pow5(Bn, (uint32_t)(-e10));
    // powers of two: A * 2^e10  vs  Bn * 2^p2
    if (e10 >= p2)
        sw_big_shl(A, (uint32_t)(e10 - p2));
    else
        sw_big_shl(Bn, (uint32_t)(p2 - e10));
    if (A.overflow) return 1;   // cannot happen within the documented bounds; saturate safely
    if (Bn.overflow) return -1;
    int c = sw_big_cmp(A, Bn);
    if (c == 0 && sticky) return 1;
    return c;
}

SW_HD inline void sw_dbl_parts(uint64_t bits, uint64_t &m, int64_t &p) {
    const uint64_t e = (bits >> 52) & 0x7ff, f = bits & 0xfffffffffffffull;
    if (e == 0) {
        m = f;
        p = -1074;
    } else {
        m = f | (1ull << 52);
        p = (int64_t)e - 1075;
    }
}

// Correctly rounded positive double for D * 10^e10 (sticky: more non-zero digits followed), starting from a
// guess within a few ulps.
SW_HD inline uint64_t sw_round_exact(const SwBig &D, int64_t e10, bool sticky, uint64_t guess, SwBig &A, SwBig &Bn) {
    uint64_t z = guess > 0x7ff0000000000000ull ? 0x7fefffffffffffffull : guess;
    if (z == 0x7ff0000000000000ull) z = 0x7fefffffffffffffull;
    for (int it = 0; it < 4096; it++) {
        uint64_t m;
        int64_t p;
        sw_dbl_parts(z, m, p);
        int c = sw_cmp_dec_bin(D, e10, sticky, 2 * m + 1, p - 1, A, Bn);  // vs half-way to the next double up
        if (c > 0 || (c == 0 && (z & 1))) {
            z++;
            if (z == 0x7ff0000000000000ull) return z;  // beyond the half-way point above MAX_VALUE: infinity
            if (c == 0) return z;
            continue;
        }
        if (c == 0) return z;
        if (z == 0) return 0;
        uint64_t zp = z - 1;
        sw_dbl_parts(zp, m, p);
        c = sw_cmp_dec_bin(D, e10, sticky, 2 * m + 1, p - 1, A, Bn);  // vs half-way to the next double down
        if (c < 0 || (c == 0 && (z & 1))) {
            z = zp;
            if (c == 0) return z;
            continue;
        }
        return z;
    }
    return z;
}

// ---- decimal scanner shared by number tokens and numeric strings ----------------------------------------------------
// text = [+-]? digits [. digits]? ([eE][+-]?digits)?   (grammar already validated by the caller; the
// plain-decimal string form additionally allows ".5" and "5.")
struct SwDec {
    uint64_t w;          // first 19 significant digits
    int64_t q;           // decimal exponent of w
    bool neg, dropped_nonzero, any_digit_nonzero;
};
#define SW_MAX_SIG_DIGITS 768

template <typename B>
SW_HD inline void sw_scan_decimal(const B &b, uint32_t off, uint32_t len, SwDec &d, SwBig *big, int64_t *big_e10,
                                  bool *big_sticky) {
    uint32_t i = off;
    const uint32_t end = off + len;
    d.neg = false;
    if (i < end && (b.byte(i) == '-' || b.byte(i) == '+')) {
        d.neg = b.byte(i) == '-';
        i++;
    }
    uint64_t w = 0;
    int64_t q = 0;
    int nd = 0;
    bool after_dot = false, dropped_nonzero = false;
    // big-number accumulation (only when `big` is given): chunks of 9 digits
    uint32_t chunk = 0, chunk_n = 0;
    int big_nd = 0;
    int64_t bq = 0;
    bool bsticky = false;
    if (big) sw_big_set(*big, 0);
    for (; i < end; i++) {
        const uint32_t c = b.byte(i);
        if (c == '.') {
            after_dot = true;
            continue;
        }
        if (c < '0' || c > '9') break;
        const uint32_t dg = c - '0';
        if (nd == 0 && dg == 0) {  // leading zero: only shifts the exponent when after the point
            if (after_dot) {
                q--;
                bq--;
            }
            continue;
        }
        if (nd < 19) {
            w = w * 10 + dg;
            if (after_dot) q--;
        } else {
            if (dg) dropped_nonzero = true;
            if (!after_dot) q++;
        }
        nd++;
        if (big) {
            if (big_nd < SW_MAX_SIG_DIGITS) {
                chunk = chunk * 10 + dg;
                chunk_n++;
                big_nd++;
                if (after_dot) bq--;
                if (chunk_n == 9) {
                    sw_big_muladd(*big, 1000000000u, chunk);
                    chunk = 0;
                    chunk_n = 0;
                }
            } else {
                if (dg) bsticky = true;
                if (!after_dot) bq++;
            }
        }
    }
    if (big && chunk_n) {
        uint32_t m = 1;
        for (uint32_t k = 0; k < chunk_n; k++) m *= 10;
        sw_big_muladd(*big, m, chunk);
    }
    // exponent part
    if (i < end && (b.byte(i) == 'e' || b.byte(i) == 'E')) {
        i++;
        bool eneg = false;
        if (i < end && (b.byte(i) == '-' || b.byte(i) == '+')) {
            eneg = b.byte(i) == '-';
            i++;
        }
        int64_t e = 0;
        for (; i < end; i++) {
            const uint32_t c = b.byte(i);
            if (c < '0' || c > '9') break;
            if (e < 100000000) e = e * 10 + (c - '0');  // saturate: anything this large is 0 or infinity
        }
        if (eneg) e = -e;
        q += e;
        bq += e;
    }
    d.w = w;
    d.q = q;
    d.dropped_nonzero = dropped_nonzero;
    d.any_digit_nonzero = nd > 0;
    if (big) {
        *big_e10 = bq;
        *big_sticky = bsticky;
    }
}

SW_HD inline uint64_t sw_crude_guess(uint64_t w, int64_t q) {
    // a few ulps off at worst: the exact fallback walks to the right neighbour
    const double p10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                          1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    double v = (double)w;
    while (q > 0) {
        int s = q > 22 ? 22 : (int)q;
        v = v * p10[s];
        q -= s;
        if (v > 1.7976931348623157e308) return 0x7fefffffffffffffull;
    }
    while (q < 0) {
        int s = q < -22 ? 22 : (int)-q;
        v = v / p10[s];
        q += s;
        if (v == 0.0) return 0;
    }
    uint64_t bits = sw_d2bits(v);
    if (bits >= 0x7ff0000000000000ull) bits = 0x7fefffffffffffffull;
    return bits;
}

template <typename B>
#ifdef __CUDACC__
__host__ __device__ __noinline__
#else
__attribute__((noinline))
#endif
uint64_t sw_parse_double_slow(const B &b, uint32_t off, uint32_t len, uint64_t guess) {
    SwBig D, A, Bn;
    SwDec d;
    int64_t e10;
    bool sticky;
    sw_scan_decimal(b, off, len, d, &D, &e10, &sticky);
    if (D.n == 0) return 0;
    // magnitudes outside the double range need no big arithmetic
    return sw_round_exact(D, e10, sticky, guess, A, Bn);
}

// Magnitude and sign are returned separately so that callers can apply token-specific rules.
template <typename B>
SW_HD inline uint64_t sw_parse_double_bits(const B &b, uint32_t off, uint32_t len) {
    SwDec d;
    sw_scan_decimal(b, off, len, d, (SwBig *)nullptr, (int64_t *)nullptr, (bool *)nullptr);
    const uint64_t sign = d.neg ? 0x8000000000000000ull : 0;
    if (d.w == 0) return sign;  // all digits zero
    if (d.q < -400) return sign;                             // below half the smallest subnormal
    if (d.q > 400) return sign | 0x7ff0000000000000ull;      // w >= 1: overflow
    if (!d.dropped_nonzero && d.w <= (1ull << 53) && d.q >= -22 && d.q <= 22) {
        const double p10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                              1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
        double v = (double)d.w;
#ifdef __CUDA_ARCH__
        v = d.q < 0 ? __ddiv_rn(v, p10[-d.q]) : __dmul_rn(v, p10[d.q]);
#else
        v = d.q < 0 ? v / p10[-d.q] : v * p10[d.q];
#endif
        return sign | sw_d2bits(v);
    }
    uint64_t b0 = 0, b1 = 0;
    const bool ok0 = sw_eisel_lemire(d.w, d.q, b0);
    if (ok0 && !d.dropped_nonzero) return sign | b0;
    if (ok0 && d.dropped_nonzero) {
        const bool ok1 = sw_eisel_lemire(d.w + 1, d.q, b1);
        if (ok1 && b0 == b1) return sign | b0;
    }
    const uint64_t guess = ok0 ? b0 : sw_crude_guess(d.w, d.q);
    return sign | sw_parse_double_slow(b, off, len, guess);
}
