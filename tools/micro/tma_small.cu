// Micro-benchmark: how fast can the SMs issue SMALL shared->global bulk copies (cp.async.bulk, 16-byte granular)?
// Mode 0: one 128-byte copy per thread per tile (128 copies per tile).  Mode 1: one 18 KB copy per tile (what
// k_frame_tiles does today).  Mode 2: per-thread copies of 112 bytes + 3 plain 16-byte stores (head/tail fix-ups).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_small tma_small.cu ; run: ./tma_small
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(128) k(uint8_t* out, uint32_t tiles, uint32_t* ticket) {
    extern __shared__ __align__(128) uint8_t s[];
    __shared__ uint32_t s_tile;
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < 128 * 160 / 4; i += 128) reinterpret_cast<uint32_t*>(s)[i] = i * 2654435761u;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    uint32_t it = 0;
    if (MODE == 3 && tid == 0) {
        uint64_t* bar = reinterpret_cast<uint64_t*>(s + 2 * 128 * 160);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    for (;;) {
        if (tid == 0) s_tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const uint32_t t = s_tile;
        __syncthreads();
        if (t >= tiles) break;
        uint8_t* g = out + (size_t)t * 128 * 160;
        if (MODE == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g + tid * 160), "r"(smem_u32(s + tid * 160)), "r"(128) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        } else if (MODE == 1) {
            if (tid == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s)), "r"(128 * 160) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
        } else if (MODE == 3) {
            // per-thread shared->shared bulk copy of 112 bytes (payload interior), completion on an mbarrier, then ONE tile store
            uint64_t* bar = reinterpret_cast<uint64_t*>(s + 2 * 128 * 160);
            if (tid == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(128 * 112) : "memory");
            }
            __syncthreads();
            asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(s + 128 * 160 + tid * 160 + 32)),
                         "r"(smem_u32(s + tid * 144 + 16)), "r"(112), "r"(smem_u32(bar)) : "memory");
            // header / tail pieces by the lane itself
            *reinterpret_cast<uint4*>(s + 128 * 160 + tid * 160) = make_uint4(tid, 1, 2, 3);
            *reinterpret_cast<uint4*>(s + 128 * 160 + tid * 160 + 16) = make_uint4(tid, 4, 5, 6);
            *reinterpret_cast<uint4*>(s + 128 * 160 + tid * 160 + 144) = make_uint4(tid, 7, 8, 9);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const uint32_t parity = it & 1;
            asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(smem_u32(s + 128 * 160)), "r"(128 * 160) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            it++;
        } else {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g + tid * 160 + 32), "r"(smem_u32(s + tid * 160 + 32)), "r"(112) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            const uint4 a = *reinterpret_cast<const uint4*>(s + tid * 160), b = *reinterpret_cast<const uint4*>(s + tid * 160 + 16),
                        c = *reinterpret_cast<const uint4*>(s + tid * 160 + 144);
            *reinterpret_cast<uint4*>(g + tid * 160) = a;
            *reinterpret_cast<uint4*>(g + tid * 160 + 16) = b;
            *reinterpret_cast<uint4*>(g + tid * 160 + 144) = c;
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int MODE> void run(uint8_t* out, uint32_t tiles, uint32_t* ticket, int ctas_per_sm) {
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 128 * 160 + 16);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9;
    for (int it = 0; it < 5; it++) {
        cudaMemset(ticket, 0, 4);
        cudaEventRecord(e0);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(148 * ctas_per_sm); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 2 * 128 * 160 + 16;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, k<MODE>, out, tiles, ticket);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    printf("mode %d ctas/sm %d: %.3f ms for %u tiles (%.1f M records), %.0f GB/s written, err=%s\n", MODE, ctas_per_sm, best, tiles,
           tiles * 128 / 1e6, (double)tiles * 128 * 160 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const uint32_t tiles = 78125;          // 10 M records
    uint8_t* out; uint32_t* ticket;
    cudaMalloc(&out, (size_t)tiles * 128 * 160);
    cudaMalloc(&ticket, 4);
    for (int c : {2, 5}) { run<1>(out, tiles, ticket, c); run<0>(out, tiles, ticket, c); run<3>(out, tiles, ticket, c); }
    return 0;
}
