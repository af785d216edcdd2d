// Small device helpers shared by the packed s16x2 kernels (k1_score.cu: K1 local, K2; k1_global.cu: K1 global).
#pragma once
#include <stdint.h>

__device__ __forceinline__ unsigned lds32(unsigned addr)
{
    unsigned v;
    asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// Pass-boundary words are written in pass q and read back in pass q+1 by the same thread: keep them in L2
// (evict_last) so that the 0.1 GB working set of the resident warps is not streamed through HBM.
__device__ __forceinline__ unsigned long long keep_policy()
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void st_keep(uint4 *p, uint4 v, unsigned long long pol)
{
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ uint4 ld_keep(const uint4 *p, unsigned long long pol)
{
    uint4 v;
    asm volatile("ld.global.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol) : "memory");
    return v;
}

__device__ __forceinline__ int lo16(unsigned v) { return (int)(short)(v & 0xFFFFu); }
__device__ __forceinline__ int hi16(unsigned v) { return ((int)v) >> 16; }
__device__ __forceinline__ unsigned pack2(int lo, int hi) { return ((unsigned)lo & 0xFFFFu) | ((unsigned)hi << 16); }

