// K0 -- DPX / integer-pipe micro-benchmark: the measured denominator of the GCUPS roofline
// (SURVEY.md 8d).  Every thread runs 8 independent dependency chains of the chosen instruction so
// that the issue rate, not the 4-cycle ALU latency, is the limit; all SMs are filled with 2048
// threads.  Reported: thread-instructions per second (warp-instructions * 32 / elapsed).
#include <cuda_fp16.h>

#include "swb_internal.h"

#define K0_ITERS 4096
#define K0_CHAINS 8

template <int WHICH>
__global__ void __launch_bounds__(1024) k0_kernel(unsigned *out, unsigned seed)
{
    unsigned v[K0_CHAINS], w[K0_CHAINS];
    const unsigned tid = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
    for (int c = 0; c < K0_CHAINS; ++c) { v[c] = seed * (c + 1) + tid; w[c] = seed ^ (tid * (c + 3)); }
    const unsigned k1 = seed | 0x00010001u, k2 = seed >> 3;
#pragma unroll 1
    for (int it = 0; it < K0_ITERS; ++it) {
#pragma unroll
        for (int c = 0; c < K0_CHAINS; ++c) {
            if (WHICH == 0) { v[c] = __viaddmax_s16x2(v[c], k1, w[c]); }
            else if (WHICH == 1) { v[c] = __vimax3_s16x2(v[c], k1, w[c]); }
            else if (WHICH == 2) { v[c] = __vadd2(v[c], k1); }
            else if (WHICH == 3) {      // the 4 DPX ops of one Gotoh cell-vector (+ packed add): 5 instructions
                unsigned E = __viaddmax_s16x2(w[c], k1, v[c]);
                unsigned F = __viaddmax_s16x2(v[c], k1, w[c]);
                unsigned T = __vmaxs2(E, F);
                unsigned H = __viaddmax_s16x2_relu(v[c], k2, T);
                v[c] = __vadd2(H, k1); w[c] = E;
            }
            else if (WHICH == 4) { v[c] = (unsigned)__viaddmax_s32((int)v[c], (int)k1, (int)w[c]); }
            else if (WHICH == 5) { v[c] = v[c] * k1 + w[c]; }
            else if (WHICH == 6) {      // cell mix + 3 IMAD (address / merge work of K1) on the FMA pipe
                unsigned E = __viaddmax_s16x2(w[c], k1, v[c]);
                unsigned F = __viaddmax_s16x2(v[c], k1, w[c]);
                unsigned T = __vmaxs2(E, F);
                unsigned a0 = E * k2 + F, a1 = F * k2 + T, a2 = a0 * 65536u + a1;
                unsigned H = __viaddmax_s16x2_relu(v[c], a2, T);
                v[c] = __vadd2(H, k1); w[c] = E;
            }
            else if (WHICH == 7) {      // packed fp16 max (HMNMX2): which pipe?
                __half2 a = *reinterpret_cast<__half2 *>(&v[c]), b = *reinterpret_cast<__half2 *>(&w[c]);
                a = __hmax2(a, b); v[c] = *reinterpret_cast<unsigned *>(&a);
            }
            else if (WHICH == 8) {      // packed fp16 fused multiply-add (HFMA2, FMA pipe)
                __half2 a = *reinterpret_cast<__half2 *>(&v[c]), b = *reinterpret_cast<__half2 *>(&w[c]);
                const __half2 kk = *reinterpret_cast<const __half2 *>(&k1);
                a = __hfma2(a, kk, b); v[c] = *reinterpret_cast<unsigned *>(&a);
            }
            else if (WHICH == 9) {      // one DPX op and one HMNMX2 on independent chains: do they share a pipe?
                v[c] = __viaddmax_s16x2(v[c], k1, k2);
                __half2 b = *reinterpret_cast<__half2 *>(&w[c]);
                const __half2 kk = *reinterpret_cast<const __half2 *>(&k1);
                b = __hmax2(b, kk); w[c] = *reinterpret_cast<unsigned *>(&b);
            }
            else if (WHICH == 10) {     // one DPX op and one HFMA2 on independent chains
                v[c] = __viaddmax_s16x2(v[c], k1, k2);
                __half2 b = *reinterpret_cast<__half2 *>(&w[c]);
                const __half2 kk = *reinterpret_cast<const __half2 *>(&k1);
                b = __hfma2(b, kk, kk); w[c] = *reinterpret_cast<unsigned *>(&b);
            }
            else if (WHICH == 11) {     // one DPX op and one IMAD on independent chains
                v[c] = __viaddmax_s16x2(v[c], k1, k2);
                w[c] = w[c] * k1 + k2;
            }
            else if (WHICH == 12) { v[c] = __viaddmax_s16x2(v[c], 0xFFFFFFFFu, w[c]); }   // immediate addend
            else if (WHICH == 13) { v[c] = __vadd2(v[c], 0xFFECFFECu); }                  // immediate addend
            else if (WHICH == 14) { v[c] = __viaddmax_s16x2(v[c], seed, w[c]); }          // addend in a uniform register
            else if (WHICH == 15) { v[c] = __viaddmax_s16x2(v[c], w[c], 0x00010001u); }   // immediate max operand
            else if (WHICH == 16) {     // K1's cell with immediate penalties (5 DPX), table words faked by the chains
                unsigned Fk = __viaddmax_s16x2(w[c], 0xFFFFFFFFu, v[c]);
                unsigned G = __viaddmax_s16x2_relu(v[c], k2, Fk);
                unsigned E = __viaddmax_s16x2(Fk, 0xFFFFFFFFu, G);
                unsigned Go = __vadd2(G, 0xFFECFFECu);
                v[c] = __viaddmax_s16x2(E, 0xFFECFFECu, Go); w[c] = Fk;
            }
            else if (WHICH == 17) {     // the same cell with the penalties in registers
                unsigned Fk = __viaddmax_s16x2(w[c], k1, v[c]);
                unsigned G = __viaddmax_s16x2_relu(v[c], k2, Fk);
                unsigned E = __viaddmax_s16x2(Fk, k1, G);
                unsigned Go = __vadd2(G, k2);
                v[c] = __viaddmax_s16x2(E, k2, Go); w[c] = Fk;
            }
        }
    }
    unsigned acc = 0;
#pragma unroll
    for (int c = 0; c < K0_CHAINS; ++c) acc ^= v[c] ^ w[c];
    if (acc == 0x12345678u) out[tid & 1023] = acc;      // keeps the chains alive; practically never taken
}

template <int WHICH>
static cudaError_t k0_time(int sm_count, int per_iter, double *rate)
{
    unsigned *d = nullptr;
    cudaError_t err = cudaMalloc(&d, 1024 * sizeof(unsigned));
    if (err != cudaSuccess) return err;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sm_count * 2, threads = 1024;
    double best_ms = 1e30;
    for (int rep = 0; rep < 12; ++rep) {
        cudaEventRecord(e0);
        k0_kernel<WHICH><<<blocks, threads>>>(d, 0x9E3779B9u + rep);
        cudaEventRecord(e1);
        err = cudaEventSynchronize(e1);
        if (err != cudaSuccess) break;
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 2 && ms < best_ms) best_ms = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    if (err != cudaSuccess) return err;
    const double instr = (double)blocks * threads * (double)K0_ITERS * K0_CHAINS * per_iter;
    *rate = instr / (best_ms * 1e-3);
    return cudaGetLastError();
}

cudaError_t k0_run(int which, int sm_count, double *rate)
{
    switch (which) {
    case 0: return k0_time<0>(sm_count, 1, rate);
    case 1: return k0_time<1>(sm_count, 1, rate);
    case 2: return k0_time<2>(sm_count, 1, rate);
    case 3: return k0_time<3>(sm_count, 5, rate);
    case 4: return k0_time<4>(sm_count, 1, rate);
    case 5: return k0_time<5>(sm_count, 1, rate);
    case 6: return k0_time<6>(sm_count, 8, rate);
    case 7: return k0_time<7>(sm_count, 1, rate);
    case 8: return k0_time<8>(sm_count, 1, rate);
    case 9: return k0_time<9>(sm_count, 2, rate);
    case 10: return k0_time<10>(sm_count, 2, rate);
    case 11: return k0_time<11>(sm_count, 2, rate);
    case 12: return k0_time<12>(sm_count, 1, rate);
    case 13: return k0_time<13>(sm_count, 1, rate);
    case 14: return k0_time<14>(sm_count, 1, rate);
    case 15: return k0_time<15>(sm_count, 1, rate);
    case 16: return k0_time<16>(sm_count, 5, rate);
    case 17: return k0_time<17>(sm_count, 5, rate);
    default: return cudaErrorInvalidValue;
    }
}
