// scratch_class.h — size classes of the context's scratch cache (common.cuh).  Plain C++: the CPU test tier compiles it too.
#pragma once
#include <cstddef>

namespace vsp {

// sizes are rounded up to one of eight classes per power of two (at most 1/8 of slack, 512 B at least) and a block serves
// its own class only: whether a request hits then depends on nothing but the requests made before it, so the second run
// of a workload already allocates nothing (a best-fit rule let a smaller request take the block a later one needed, and
// took several runs to settle), while nearby batch sizes still share blocks
inline size_t scratch_size_class(size_t n) {
    if (n <= 512) return 512;
    unsigned top = 63 - (unsigned)__builtin_clzll((unsigned long long)(n - 1));  // 2^top < n <= 2^(top+1)
    const size_t step = size_t(1) << (top - 3);                                  // an eighth of 2^top     
    return (n + step - 1) & ~(step - 1);
}

}  // namespace vsp
