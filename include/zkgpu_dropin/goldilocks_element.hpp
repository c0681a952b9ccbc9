// goldilocks_element.hpp — the one type the drop-in classes share with the reference.
//
// Inside zkevm-prover the reference's own src/goldilocks/src/goldilocks_base_field.hpp (include guard GOLDILOCKS_BASE,
// `class Goldilocks { typedef struct { uint64_t fe; } Element; ... }`, :20-27) is on the include path and is used
// unchanged: include it BEFORE these headers.  For stand-alone builds (tests/cpp, other hosts) this file provides
// a layout-identical stand-in with only what the drop-in classes touch.
#ifndef ZKGPU_GOLDILOCKS_ELEMENT_HPP
#define ZKGPU_GOLDILOCKS_ELEMENT_HPP
#include <stdint.h>

#ifndef GOLDILOCKS_BASE
#define GOLDILOCKS_BASE
#define GOLDILOCKS_PRIME 0xFFFFFFFF00000001ULL
class Goldilocks
{
public:
    typedef struct
    {
        uint64_t fe;
    } Element;
    // goldilocks_base_field_tools.hpp:152-163
    static inline uint64_t toU64(const Element &in1) { return in1.fe >= GOLDILOCKS_PRIME ? in1.fe - GOLDILOCKS_PRIME : in1.fe; }
    static inline Element fromU64(uint64_t in1)
    {
        Element e;
        e.fe = in1;
        return e;
    }
};
#endif

static_assert(sizeof(Goldilocks::Element) == sizeof(uint64_t), "Goldilocks::Element must be one 64-bit word");

#include <stdexcept>
#include <string>
#include "../zkgpu.h"
namespace zkgpu_dropin
{
// The reference's functions are all `void`; its convention for fatal conditions is log + exit.  The C ABI reports
// rc < 0 + message, which the wrappers turn into an exception (there is no CPU fallback to fall through to).
inline void check(int rc, const char *what)
{
    if (rc < 0) throw std::runtime_error(std::string(what) + ": " + zkgpu_last_error());
}
inline uint64_t *words(Goldilocks::Element *p) { return reinterpret_cast<uint64_t *>(p); }
inline const uint64_t *words(const Goldilocks::Element *p) { return reinterpret_cast<const uint64_t *>(p); }
} // namespace zkgpu_dropin
#endif
