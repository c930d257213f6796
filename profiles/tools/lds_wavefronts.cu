// Shared-memory wavefront cost of the access shapes the tiled kernel uses (read with ncu:
//   ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum ./lds_wavefronts)
// One CTA of one warp per case; each case issues 1024 loads of the named shape.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CASE>
__global__ void probe(uint64_t* out) {
    __shared__ __align__(16) uint32_t sm[8192];
    const int lane = threadIdx.x;
    for (int k = lane; k < 8192; k += 32) sm[k] = k;
    __syncthreads();
    uint64_t acc = 0;
    uint32_t addr;
    uint32_t base = (uint32_t)__cvta_generic_to_shared(sm);
    // per-lane byte address by case
    if (CASE == 0) addr = base;                                  // LDS.128 broadcast
    else if (CASE == 1) addr = base + lane * 16;                 // LDS.128 distinct, contiguous
    else if (CASE == 2) addr = base;                             // LDS.64 broadcast
    else if (CASE == 3) addr = base + lane * 8;                  // LDS.64 distinct, contiguous
    else if (CASE == 4) addr = base + ((lane * 7) % 24) * 8;     // LDS.64: 24-entry row, lanes share entries
    else if (CASE == 5) addr = base + (lane % 12) * 8;           // LDS.64: 12 distinct entries
    else if (CASE == 6) addr = base;                             // LDS.32 broadcast
    else if (CASE == 7) addr = base + (lane / 16) * 16;          // LDS.128: two distinct quads (half-warps)
    else if (CASE == 8) addr = base + (lane % 8) * 8;            // LDS.64: 8 distinct entries (64 B)
    else addr = base + ((lane * 5) % 16) * 8;                    // LDS.64: 16 distinct entries = 128 B, permuted
    const uint32_t addr0 = addr;
#pragma unroll 1
    for (int it = 0; it < 1024; ++it) {
        addr = addr0 + (uint32_t)(it & 1) * 16384u;      // same bank pattern, another copy: keeps every load in the loop
        if (CASE == 0 || CASE == 1 || CASE == 7) {
            uint32_t a, b, c, d;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
            acc += a + b + c + d;
        } else if (CASE == 6) {
            uint32_t a;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(addr) : "memory");
            acc += a;
        } else {
            uint64_t a;
            asm volatile("ld.shared.u64 %0, [%1];" : "=l"(a) : "r"(addr) : "memory");
            acc += a;
        }
    }
    out[CASE * 32 + lane] = acc;
}

int main() {
    uint64_t* d;
    cudaMalloc(&d, 32 * 16 * 8);
    probe<0><<<1, 32>>>(d); probe<1><<<1, 32>>>(d); probe<2><<<1, 32>>>(d); probe<3><<<1, 32>>>(d); probe<4><<<1, 32>>>(d);
    probe<5><<<1, 32>>>(d); probe<6><<<1, 32>>>(d); probe<7><<<1, 32>>>(d); probe<8><<<1, 32>>>(d); probe<9><<<1, 32>>>(d);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
