// ntt_goldilocks.hpp — drop-in for the reference's src/goldilocks/src/ntt_goldilocks.hpp (+ ntt_goldilocks.cpp).
//
// Same class name, same public signatures and defaults (reference ntt_goldilocks.hpp:49, 153, 166-173), same include
// guard, so that putting include/zkgpu_dropin first on the include path (and dropping ntt_goldilocks.cpp from the
// build, linking -lzkgpu instead) moves Starks::ntt / Starks::nttExtended onto the GPU with no change in starkpil.
// Every method forwards to the C ABI in include/zkgpu.h; pointers may be host or device pointers.
//
// Differences a caller can observe, all documented in DESIGN.md:
//   * outputs are canonical (< p) words, i.e. Goldilocks::toU64 of what the reference writes;
//   * `buffer`, `nphase`, `nblock`, `nThreads` are accepted and ignored (CPU blocking knobs);
//   * there is no host-side roots table: root(domainPow, idx) computes the value on demand;
//   * errors throw std::runtime_error instead of assert()/undefined behaviour.
#ifndef NTT_GOLDILOCKS
#define NTT_GOLDILOCKS
#include <atomic>
#include <cstdio>
#include "goldilocks_element.hpp"
#include <sys/types.h>

#define NUM_PHASES 3
#define NUM_BLOCKS 1

class NTT_Goldilocks
{
private:
    void *h = nullptr;
    u_int32_t s = 0;
    Goldilocks::Element rootTmp;
    static u_int32_t log2(u_int64_t size)
    {
        u_int32_t res = 0;
        while (size > 1)
        {
            size >>= 1;
            res++;
        }
        return res;
    }
    static uint64_t mulmod(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) % GOLDILOCKS_PRIME); }

public:
    NTT_Goldilocks(u_int64_t maxDomainSize, u_int32_t _nThreads = 0, int extension_ = 1)
    {
        (void)_nThreads;
        if (maxDomainSize == 0) return; // ntt_goldilocks.hpp:54-55
        s = log2(maxDomainSize);
        h = zkgpu_ntt_create(maxDomainSize, extension_);
        if (!h) throw std::range_error(zkgpu_last_error()); // "Domain size too big for the curve", ntt_goldilocks.hpp:94-97
    }
    NTT_Goldilocks(const NTT_Goldilocks &) = delete;
    NTT_Goldilocks &operator=(const NTT_Goldilocks &) = delete;
    ~NTT_Goldilocks()
    {
        if (h) zkgpu_ntt_destroy(h);
    }
    // r / r_ (7^i, 7^i / N) live in device tables built on first use; kept for source compatibility
    inline void computeR(int) {}

    void NTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL,
             u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool inverse = false, bool extend = false)
    {
        (void)nphase;
        (void)nblock;
        // FRI's folding loop (friProve.cpp:44-102) runs one tiny INTT per group from inside an OpenMP region: through this class
        // every one of them is a host-pointer call (lock, copy in, launch, copy out, synchronise), serialised on the device's
        // context — correct, but far slower than the CPU was.  Say so once; zkgpu_fri_fold does a whole layer in one call.
        if (size <= 64 && ncols <= 4)
        {
            static std::atomic<unsigned> tiny{0};
            if (tiny.fetch_add(1, std::memory_order_relaxed) == 1024)
                fprintf(stderr, "[zkgpu] 1024 transforms of <= 64 rows x <= 4 columns went through NTT_Goldilocks one by one; if this is "
                                "FRIProve::prove's folding loop, call zkgpu_fri_fold per layer instead (INTEGRATION.md section 2)\n");
        }
        zkgpu_dropin::check(zkgpu_ntt(h, zkgpu_dropin::words(dst), zkgpu_dropin::words(src), size, ncols, zkgpu_dropin::words(buffer),
                                      inverse ? 1 : 0, extend ? 1 : 0),
                            "NTT_Goldilocks::NTT");
    }
    void INTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL,
              u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool extend = false)
    {
        NTT(dst, src, size, ncols, buffer, nphase, nblock, true, extend);
    }
    void reversePermutation(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t offset_cols, u_int64_t ncols,
                            u_int64_t ncols_all)
    {
        zkgpu_dropin::check(zkgpu_reverse_permutation(h, zkgpu_dropin::words(dst), zkgpu_dropin::words(src), size, offset_cols, ncols, ncols_all),
                            "NTT_Goldilocks::reversePermutation");
    }
    // roots[idx << (s - domainPow)] of the reference = W[domainPow]^idx (ntt_goldilocks.hpp:112-113, 169-172)
    inline Goldilocks::Element &root(u_int32_t domainPow, u_int64_t idx)
    {
        uint64_t w = 7277203076849721926ULL; // Goldilocks::W[32], goldilocks_base_field.cpp:11-44
        for (u_int32_t i = 32; i > domainPow; --i) w = mulmod(w, w);
        uint64_t r = 1;
        for (u_int64_t e = idx; e; e >>= 1)
        {
            if (e & 1) r = mulmod(r, w);
            w = mulmod(w, w);
        }
        rootTmp.fe = r;
        return rootTmp;
    }
    void extendPol(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t N_Extended, uint64_t N, uint64_t ncols,
                   Goldilocks::Element *buffer = NULL, u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS)
    {
        (void)nphase;
        (void)nblock;
        zkgpu_dropin::check(zkgpu_extend_pol(h, zkgpu_dropin::words(output), zkgpu_dropin::words(input), N_Extended, N, ncols,
                                             zkgpu_dropin::words(buffer)),
                            "NTT_Goldilocks::extendPol");
    }
    void *handle() const { return h; }
};
#endif
