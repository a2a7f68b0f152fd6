// Does cp.async.bulk shared::cta -> shared::cluster work (a) onto the issuing CTA itself, (b) onto the peer CTA of a 2-cluster?
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int PEER>
__global__ void __launch_bounds__(128) k(uint32_t* out) {
    __shared__ __align__(128) uint8_t src[4096];
    __shared__ __align__(128) uint8_t dst[4096];
    __shared__ __align__(8) uint64_t bar;
    cg::cluster_group cl = cg::this_cluster();
    const uint32_t tid = threadIdx.x, rank = cl.block_rank();
    for (uint32_t i = tid; i < 1024; i += 128) { reinterpret_cast<uint32_t*>(src)[i] = i + 1000 * rank; reinterpret_cast<uint32_t*>(dst)[i] = 0; }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(128 * 16) : "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    cl.sync();
    const uint32_t target = PEER ? (rank ^ 1u) : rank;
    uint32_t d_addr, b_addr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d_addr) : "r"(smem_u32(dst + tid * 32)), "r"(target));
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(b_addr) : "r"(smem_u32(&bar)), "r"(target));
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(d_addr), "r"(smem_u32(src + tid * 16)), "r"(16), "r"(b_addr) : "memory");
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    __syncthreads();
    if (blockIdx.x == 0) out[tid] = reinterpret_cast<uint32_t*>(dst)[tid * 8];
    cl.sync();
}

template <int PEER> void run(int cluster) {
    uint32_t* out; cudaMalloc(&out, 512); cudaMemset(out, 0xFF, 512);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(128);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k<PEER>, out);
    cudaError_t e2 = cudaDeviceSynchronize();
    uint32_t h[128]; cudaMemcpy(h, out, 512, cudaMemcpyDeviceToHost);
    printf("peer=%d cluster=%d launch=%s sync=%s out[0..3]=%u %u %u %u\n", PEER, cluster, cudaGetErrorString(e), cudaGetErrorString(e2), h[0], h[1], h[2], h[3]);
}
int main(int argc, char** argv) {
    int which = argc > 1 ? atoi(argv[1]) : 0;
    if (which == 0) run<0>(1);
    if (which == 1) run<0>(2);
    if (which == 2) run<1>(2);
    return 0;
}
