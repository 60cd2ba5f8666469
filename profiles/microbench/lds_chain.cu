// Dependent table-lookup chains in shared memory on sm_100a: cycles per step for the address arithmetic variants the
// scan could use.  One warp per CTA measures latency, more warps measure how it scales.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_chain lds_chain.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int STEPS = 8192;
constexpr int ROW = 516;     // bytes per state row (16-bit entries)
constexpr int NST = 157;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t lds16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }

template <int MODE>
__global__ void chain(const uint32_t *text, uint32_t *out, long long *cycles, int same_byte) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t tid = threadIdx.x;
    // table: entry(st, b) -> next state pseudo-random (or stay), flags in high bits
    for (uint32_t i = tid; i < NST * 256; i += blockDim.x) {
        uint32_t st = i >> 8, b = i & 255;
        uint32_t nx = (st * 7 + b * 13 + 1) % NST;
        if (MODE == 0 || MODE == 1 || MODE == 6) *reinterpret_cast<uint16_t *>(smem + st * ROW + b * 2) = (uint16_t)(nx | ((b & 1) << 15));
        if (MODE == 2) *reinterpret_cast<uint32_t *>(smem + st * 1028 + b * 4) = nx * 1028 | ((b & 1u) << 31);     // pre-multiplied + flag
        if (MODE == 3) *reinterpret_cast<uint32_t *>(smem + st * 1028 + b * 4) = nx * 1028;                         // pure offset
        if (MODE == 4 || MODE == 5) smem[st * 260 + b] = (uint8_t)nx;
    }
    __syncthreads();
    const uint32_t base = smem_u32(smem);
    uint32_t e = tid % NST, words = 0;
    if (MODE == 2 || MODE == 3) e = (tid % NST) * 1028;
    // the lane's text: 8 words held in registers (a load from global memory inside the loop would be measured instead)
    uint32_t tw[8];
#pragma unroll
    for (int j = 0; j < 8; j++) tw[j] = same_byte ? 0x61616161u : text[(tid & 31) * 64 + j + 8 * (tid >> 5)];
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < STEPS / 32; i++) {
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const uint32_t w = tw[j];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (MODE == 0) {           // PRMT + IDP.4A + LDS.U16 (the shipped sequence)
                uint32_t a = __dp4a(__byte_perm(w, e, 0x4440u | k), 0xACACAC02u, base);
                e = lds16(a); words += e >> 15;
            } else if (MODE == 6) {    // shipped sequence, half of the lanes idle: is the limit per instruction or per lane?
                if ((tid & 1) == 0) {
                    uint32_t a = __dp4a(__byte_perm(w, e, 0x4440u | k), 0xACACAC02u, base);
                    e = lds16(a); words += e >> 15;
                }
            } else if (MODE == 1) {    // LOP + IMAD + LDS.U16
                uint32_t b2 = (w >> (8 * k) & 0xFFu) * 2u;
                uint32_t a = (e & 0xFFu) * ROW + b2 + base;
                e = lds16(a); words += e >> 15;
            } else if (MODE == 2) {    // 32-bit entries, pre-multiplied row offset + flag: LOP + IADD + LDS.32
                uint32_t b4 = (w >> (8 * k) & 0xFFu) * 4u + base;
                uint32_t a = (e & 0x7FFFFFFFu) + b4;
                e = lds32(a); words += e >> 31;
            } else if (MODE == 3) {    // 32-bit entries = pure row offset: IADD + LDS.32
                uint32_t b4 = (w >> (8 * k) & 0xFFu) * 4u + base;
                e = lds32(e + b4);
            } else if (MODE == 5) {    // 8-bit entries = pure state: IMAD(e, 260, byte + base) + LDS.U8; flags from the state id
                uint32_t pre = __dp4a(w, 1u << (8 * k), base);   // off the dependent chain
                e = lds8(e * 260u + pre);
                words = __umulhi(e, 0x1C0E070u) + words;           // + (e >= 146)
            } else if (MODE == 4) {    // 8-bit entries: PRMT + IDP + LDS.U8
                uint32_t a = __dp4a(__byte_perm(w, e, 0x4440u | k), 0x00828201u, base);
                e = lds8(a); words += e >> 7;
            }
        }
      }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + tid] = e + words;
    if (tid == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, const uint32_t *text, uint32_t *out, long long *cyc, int smem) {
    cudaFuncSetAttribute(chain<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int same = 1; same >= 0; same--)
        for (int warps : {1, 4, 8, 16, 24, 32}) {
            chain<MODE><<<148, warps * 32, smem>>>(text, out, cyc, same);
            cudaDeviceSynchronize();
            chain<MODE><<<148, warps * 32, smem>>>(text, out, cyc, same);
            cudaError_t err = cudaDeviceSynchronize();
            long long h[148];
            cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0; for (int i = 0; i < 148; i++) avg += (double)h[i]; avg /= 148;
            printf("%-34s %s warps/SM %2d  cycles/step %6.1f  lookups/clk/SM %6.2f  %s\n", name, same ? "same-byte  " : "random-byte", warps,
                   avg / STEPS, warps * 32.0 * STEPS / avg, err == cudaSuccess ? "" : cudaGetErrorString(err));
        }
}

int main() {
    uint32_t *text, *out; long long *cyc;
    cudaMalloc(&text, 32 * 64 * 4); cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    uint32_t h[32 * 64]; uint32_t s = 12345;
    for (int i = 0; i < 32 * 64; i++) { uint32_t w = 0; for (int k = 0; k < 4; k++) { s = s * 1664525u + 1013904223u; w |= (0x61u + (s >> 24) % 26u) << (8 * k); } h[i] = w; }
    cudaMemcpy(text, h, sizeof(h), cudaMemcpyHostToDevice);
    run<0>("u16 PRMT+IDP.4A (shipped)", text, out, cyc, NST * ROW + 64);
    run<1>("u16 LOP+IMAD", text, out, cyc, NST * ROW + 64);
    run<2>("u32 premultiplied+flag LOP+IADD", text, out, cyc, NST * 1028 + 64);
    run<3>("u32 pure offset IADD", text, out, cyc, NST * 1028 + 64);
    run<4>("u8 PRMT+IDP.4A", text, out, cyc, NST * 260 + 64);
    run<5>("u8 pure state IMAD", text, out, cyc, NST * 260 + 64);
    run<6>("u16 PRMT+IDP.4A, 16 of 32 lanes", text, out, cyc, NST * ROW + 64);   // lookups/clk counts all 32 lanes: halve it
    return 0;
}
