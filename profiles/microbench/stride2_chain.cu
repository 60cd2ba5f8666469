// Feasibility microbenchmark for a two-byte-stride scan loop on sm_100a (profiles/microbench, not part of the library).
//
// The shipped scan does one shared-memory table lookup per input byte and is bound by the rate at which an SM accepts
// LDS instructions (lds_chain.cu).  This measures the alternative inner loops on text staged in shared memory, per 16-byte
// group and lane:
//   MODE 0  shipped shape: 16 dependent lookups (PRMT + IDP.4A + LDS.U16), word-start bit from the entry
//   MODE 1  two-byte stride over 4-bit byte classes: classes by SIMD arithmetic on the 4 words of the group, 8 dependent
//           lookups indexed by (state, class pair), word starts counted by SIMD-in-register compares (no lookups)
//   MODE 2  MODE 1 without the word count (how much of MODE 1 is the SIMD word count)
//   MODE 3  MODE 1 with the word starts summed by VABSDIFF4.ACC instead of byte-lane adds
// Output: bytes per clock per SM for 8..32 warps per SM.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stride2_chain stride2_chain.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int NST = 160;
constexpr int ROW = 516;
constexpr int WIN = 80;                 // bytes per lane and window, pitch 80: conflict-free 128-bit reads
constexpr int ITERS = 2048;             // windows per warp

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }

struct Acc { uint32_t e, words, ev, prev_ns, lanes; };

// non-space bits (bit 7 of every byte) for ASCII whitespace {9..13, 32}; bytes >= 0x80 count as non-space
__device__ __forceinline__ uint32_t nonspace_bits(uint32_t w) {
    const uint32_t w7 = w & 0x7F7F7F7Fu;
    const uint32_t ne20 = (w7 ^ 0x20202020u) + 0x7F7F7F7Fu;   // bit 7: byte != 0x20
    const uint32_t gt13 = w7 + 0x72727272u;                   // bit 7: byte > 13
    const uint32_t ge9 = w7 + 0x77777777u;                    // bit 7: byte >= 9
    return ((ne20 & (gt13 | ~ge9)) | w) & 0x80808080u;
}

template <int MODE>
__device__ __forceinline__ void group(uint32_t tbl, const uint4 q, Acc &A) {
    if (MODE == 0) {
        const uint32_t wd[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int k = 0; k < 4; k++) {
                uint32_t a = __dp4a(__byte_perm(wd[j], A.e, 0x4440u | k), 0xACACAC02u, tbl);
                A.e = lds16(a); A.words += A.e >> 15; A.ev |= A.e;
            }
    } else {
        const uint32_t wd[4] = {q.x, q.y, q.z, q.w};
        // rare bytes (>= 0x80) are handed to the exact path; here: just a flag
        A.ev |= (q.x | q.y | q.z | q.w) & 0x80808080u;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t w = wd[j];
            if (MODE != 2) {
                const uint32_t ns = nonspace_bits(w);
                const uint32_t ps = __funnelshift_l(A.prev_ns, ns, 8);   // non-space bit of the byte before
                const uint32_t starts = ns & ~ps;
                A.prev_ns = ns;
                if (MODE == 3) A.words = __vsadu4(starts, 0u) + A.words;   // 128 per start
                else A.lanes += starts >> 7;
            }
            // 4-bit classes by arithmetic: ((b & 31) * m) >> k, pairs (c0 | c1 << 4) in bytes 0 and 2
            const uint32_t h = (((w & 0x1F1F1F1Fu) * 5u) >> 3) & 0x0F0F0F0Fu;
            const uint32_t p = h + (h >> 4);
#pragma unroll
            for (int k = 0; k < 4; k += 2) {
                uint32_t a = __dp4a(__byte_perm(p, A.e, 0x4440u | k), 0xACACAC02u, tbl);
                A.e = lds16(a); A.ev |= A.e;
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(1024, 1) scan(const uint8_t *text, uint32_t *out, long long *cycles) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    uint8_t *tblp = smem;
    uint8_t *stage = smem + ((NST * ROW + 127) / 128) * 128 + warp * (32 * WIN);
    // table: mostly "fall back to a shallow state", some deeper transitions; entry = next | flag << 15
    for (uint32_t i = tid; i < NST * 256; i += blockDim.x) {
        uint32_t st = i >> 8, b = i & 255;
        uint32_t hsh = (st * 2654435761u + b * 40503u) >> 7;
        uint32_t nx = (hsh % 100 < 70) ? (b % 24) : (24 + hsh % (NST - 24));
        *reinterpret_cast<uint16_t *>(tblp + st * ROW + b * 2) = (uint16_t)(nx | ((hsh & 7) == 0 ? 0x8000u : 0u));
    }
    for (uint32_t i = lane; i < 32 * WIN / 16; i += 32)
        reinterpret_cast<uint4 *>(stage)[i] = reinterpret_cast<const uint4 *>(text)[(warp * 32 * WIN / 16 + i) % 4096];
    __syncthreads();
    const uint32_t tbl = smem_u32(tblp);
    Acc A{lane % NST, 0, 0, 0x80808080u, 0};
    const uint8_t *mine = stage + lane * WIN;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll 1
        for (int g = 0; g < WIN; g += 16) {
            const uint4 q = *reinterpret_cast<const uint4 *>(mine + g);
            group<MODE>(tbl, q, A);
        }
        if (MODE == 1) { A.words += (A.lanes * 0x01010101u) >> 24; A.lanes = 0; }
        __syncwarp();
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + tid] = A.e + A.words + A.ev + A.lanes;
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
    (void)nw;
}

template <int MODE>
static void run(const char *name, const uint8_t *d_text, uint32_t *d_out, long long *d_cyc, int sms) {
    for (int nw : {8, 12, 16, 20, 24, 28, 32}) {
        size_t smem = ((NST * ROW + 127) / 128) * 128 + (size_t)nw * 32 * WIN;
        cudaFuncSetAttribute(scan<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        scan<MODE><<<sms, nw * 32, smem>>>(d_text, d_out, d_cyc);
        scan<MODE><<<sms, nw * 32, smem>>>(d_text, d_out, d_cyc);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed: %s\n", name, cudaGetErrorString(cudaGetLastError())); return; }
        long long cyc[1024];
        cudaMemcpy(cyc, d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
        double mean = 0;
        for (int i = 0; i < sms; i++) mean += (double)cyc[i];
        mean /= sms;
        const double bytes = (double)nw * 32 * WIN * ITERS;
        printf("%-44s warps/SM %2d  bytes/clk/SM %6.2f  (at 1.965 GHz x %d SMs: %5.0f GB/s)\n", name, nw, bytes / mean, sms,
               bytes / mean * 1.965 * sms);
    }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    // English-like text: lower-case words of 2-9 letters separated by single spaces, some capitals and punctuation
    static uint8_t text[65536];
    uint64_t rng = 88172645463325252ull;
    auto rnd = [&]() { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return (uint32_t)(rng >> 11); };
    const char *freq = "eeeeeeeeeeeetttttttttaaaaaaaaooooooooiiiiiiinnnnnnnsssssshhhhhhrrrrrrddddlllluuucccmmmwwffggyyppbbvk";
    size_t n = 0;
    while (n < sizeof(text)) {
        int len = 2 + rnd() % 8;
        for (int i = 0; i < len && n < sizeof(text); i++) {
            char c = freq[rnd() % 98];
            if (i == 0 && rnd() % 10 == 0) c = (char)(c - 32);
            text[n++] = (uint8_t)c;
        }
        if (n < sizeof(text) && rnd() % 12 == 0) text[n++] = ",.!?"[rnd() % 4];
        if (n < sizeof(text)) text[n++] = ' ';
    }
    uint8_t *d_text; uint32_t *d_out; long long *d_cyc;
    cudaMalloc(&d_text, sizeof(text)); cudaMalloc(&d_out, (size_t)sms * 1024 * 4); cudaMalloc(&d_cyc, (size_t)sms * 8);
    cudaMemcpy(d_text, text, sizeof(text), cudaMemcpyHostToDevice);
    run<0>("one lookup per byte (shipped loop shape)", d_text, d_out, d_cyc, sms);
    run<1>("two-byte stride + SIMD word count", d_text, d_out, d_cyc, sms);
    run<2>("two-byte stride, no word count", d_text, d_out, d_cyc, sms);
    run<3>("two-byte stride + word count via VABSDIFF4", d_text, d_out, d_cyc, sms);
    return 0;
}
