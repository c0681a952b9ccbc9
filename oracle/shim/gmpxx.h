/* oracle build glue (TEST INFRASTRUCTURE, not product code).
 * Minimal stand-in for <gmpxx.h>: the reference uses mpz_class only in string / signed
 * conversions (src/goldilocks/src/goldilocks_base_field_tools.hpp:118-240), never on the
 * NTT / Poseidon data path.  Values fit in a signed __int128. */
#ifndef ZKGPU_ORACLE_GMPXX_SHIM_H
#define ZKGPU_ORACLE_GMPXX_SHIM_H
#include "gmp.h"
#include <string>
#include <cstring>
#include <cstdlib>
#include <stdexcept>
#include <iostream>
class mpz_class
{
public:
    __int128 v;
    mpz_class() : v(0) {}
    mpz_class(uint64_t x) : v((__int128)x) {}
    mpz_class(unsigned long long x) : v((__int128)x) {}
    mpz_class(int64_t x) : v(x) {}
    mpz_class(long long x) : v(x) {}
    mpz_class(int x) : v(x) {}
    mpz_class(unsigned x) : v(x) {}
    mpz_class(const std::string &s, int radix = 10) : v(0)
    {
        size_t i = 0; bool neg = false;
        if (i < s.size() && (s[i] == '-' || s[i] == '+')) { neg = (s[i] == '-'); ++i; }
        if (radix == 16 && i + 1 < s.size() && s[i] == '0' && (s[i + 1] == 'x' || s[i + 1] == 'X')) i += 2;
        for (; i < s.size(); ++i)
        {
            int d; char c = s[i];
            if (c >= '0' && c <= '9') d = c - '0';
            else if (c >= 'a' && c <= 'z') d = c - 'a' + 10;
            else if (c >= 'A' && c <= 'Z') d = c - 'A' + 10;
            else throw std::invalid_argument("mpz_class shim: bad digit");
            if (d >= radix) throw std::invalid_argument("mpz_class shim: bad digit");
            v = v * radix + d;
        }
        if (neg) v = -v;
    }
    unsigned long get_ui() const { return (unsigned long)(v < 0 ? -v : v); }
    long get_si() const { return (long)v; }
    std::string get_str(int radix = 10) const
    {
        if (v == 0) return "0";
        __int128 a = v < 0 ? -v : v; std::string r;
        while (a) { int d = (int)(a % radix); r.insert(r.begin(), (char)(d < 10 ? '0' + d : 'a' + d - 10)); a /= radix; }
        if (v < 0) r.insert(r.begin(), '-');
        return r;
    }
};
static inline mpz_class zk_mk(__int128 x) { mpz_class r; r.v = x; return r; }
static inline mpz_class operator+(const mpz_class &a, const mpz_class &b) { return zk_mk(a.v + b.v); }
static inline mpz_class operator-(const mpz_class &a, const mpz_class &b) { return zk_mk(a.v - b.v); }
static inline mpz_class operator*(const mpz_class &a, const mpz_class &b) { return zk_mk(a.v * b.v); }
/* GMP's operator% truncates toward zero, like C++ */
static inline mpz_class operator%(const mpz_class &a, const mpz_class &b) { return zk_mk(a.v % b.v); }
static inline bool operator>(const mpz_class &a, const mpz_class &b) { return a.v > b.v; }
static inline bool operator<(const mpz_class &a, const mpz_class &b) { return a.v < b.v; }
static inline bool operator>=(const mpz_class &a, const mpz_class &b) { return a.v >= b.v; }
static inline bool operator<=(const mpz_class &a, const mpz_class &b) { return a.v <= b.v; }
static inline bool operator==(const mpz_class &a, const mpz_class &b) { return a.v == b.v; }
static inline bool operator!=(const mpz_class &a, const mpz_class &b) { return a.v != b.v; }
#endif
