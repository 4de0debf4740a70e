// Host build of align_kernel's per-thread routine (kernels.cuh: ByteStream, align_one), TEST CODE: the lines are
// taken from kernels.cuh at build time (tests/test_align.py) and compiled with the few CUDA names stubbed, one
// "lane" at a time, so the routine can be compared with oracle/twin.c on the CPU.
#include <cstdint>
#include <cstring>
#include <vector>
#define __device__
#define __forceinline__ inline
struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
template <typename T> static inline T __ldg(const T* p) { return *p; }
static struct { unsigned x; } threadIdx = {0};
static inline bool __all_sync(unsigned, bool p) { return p; }
static inline bool __any_sync(unsigned, bool p) { return p; }
static inline void __syncwarp() {}
#include "../../tksm_b200/csrc/align.cuh"
namespace tk {
#include "_build/align_one_body.inc"
}
extern "C" int align_one_emul(const uint8_t* frag, const uint8_t* slen, const uint8_t* rb, uint8_t* out, uint32_t n, uint32_t m,
                              int wide, uint32_t* eq, uint32_t* cols) {
    // padded copies: the routine reads a few bytes past the arrays (inside the padded record on the device)
    std::vector<uint8_t> f(n + 64, 0), l(n + 64, 0), r(m + 64, 0), o(m + 64, 0);
    memcpy(f.data(), frag, n); memcpy(l.data(), slen, n); memcpy(r.data(), rb, m);
    const uint32_t off = 1 + (n * 7 + m) % 23;   // the row shift a shorter read of a warp gets
    bool ok;
    if (wide) {
        std::vector<tk::AlignCell2<uint64_t>> cells((size_t)(n + 16 + off) * 32);
        ok = tk::align_one<uint64_t>(f.data(), l.data(), r.data(), o.data(), n, m, cells.data(), off, n + off % 5, true, *eq, *cols);
    } else {
        std::vector<tk::AlignCell2<uint32_t>> cells((size_t)(n + 16 + off) * 32);
        ok = tk::align_one<uint32_t>(f.data(), l.data(), r.data(), o.data(), n, m, cells.data(), off, n + off % 5, true, *eq, *cols);
    }
    memcpy(out, o.data(), m);
    return ok ? 0 : 1;
}
