// bulk_ring_bench.cu -- what does a bulk-copy ring cost per CHUNK?  (question left open by the chunk-size sweep of the
// block-sparse persistent PCG kernel: 20 / 40 KB chunks -> 73 / 57 us per CG iteration.)
// One persistent CTA per SM streams its contiguous share of a 330 MB buffer through a 192 KB shared-memory ring of
// fixed-size slots with cp.async.bulk + mbarriers -- the same producer protocol as k_pcg_solve, lane 0 of one warp -- while
// NC consumer warps only wait for a chunk and release it (no math).  Prints us per pass and GB/s for every chunk size
// and NC = 1 / 8 (8 = every consumer warp visits every chunk, as in the kernel).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_ring_bench bulk_ring_bench.cu ; run: ./bulk_ring_bench
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr uint32_t RING = 192 * 1024;
constexpr int MAXS = 32;

__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(saddr(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(unsigned long long* b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(saddr(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(unsigned long long* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(saddr(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned long long* b, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                 :: "r"(saddr(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(saddr(dst)), "l"(src), "r"(bytes), "r"(saddr(bar)) : "memory");
}

// chunks_per_cta chunks of `chunk` bytes per pass and CTA; nslots = RING / chunk; nc consumer warps (warps 1 .. nc)
__global__ void __launch_bounds__(288, 1) k_ring(const unsigned char* src, int chunks_per_cta, uint32_t chunk, int nslots, int nc, int passes) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ unsigned long long full[MAXS], empty[MAXS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < nslots; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], (uint32_t)nc); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const unsigned char* mine = src + (size_t)blockIdx.x * chunks_per_cta * chunk;
    const unsigned int total = (unsigned int)chunks_per_cta * (unsigned int)passes;
    if (warp == 0) {
        if (lane == 0)
            for (unsigned int n = 0; n < total; n++) {
                const unsigned int s = n % nslots, use = n / nslots;
                if (use > 0) mbar_wait(&empty[s], (use - 1) & 1u);
                mbar_expect(&full[s], chunk);
                bulk_load(smem + (size_t)s * chunk, mine + (size_t)(n % chunks_per_cta) * chunk, chunk, &full[s]);
            }
    } else if (warp <= nc) {
        unsigned long long sink = 0;
        for (unsigned int n = 0; n < total; n++) {
            const unsigned int s = n % nslots;
            mbar_wait(&full[s], (n / nslots) & 1u);
            sink += *reinterpret_cast<volatile unsigned long long*>(smem + (size_t)s * chunk + 8 * lane);   // one touch per lane
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        if (sink == 0x123456789abcdefULL) printf("never\n");
    }
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t buf_bytes = 336u << 20;
    unsigned char* buf = nullptr;
    if (cudaMalloc((void**)&buf, buf_bytes) != cudaSuccess) { printf("cudaMalloc failed\n"); return 1; }
    cudaMemset(buf, 1, buf_bytes);
    cudaFuncSetAttribute(k_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RING);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const uint32_t chunks[] = {8192, 12288, 16384, 24576, 32768, 49152, 65536};
    const int passes = 40;
    printf("%d SMs; %d passes per launch over ~330 MB; ring %u KB\n", sms, passes, RING / 1024);
    printf("%8s %6s %4s %10s %10s %12s\n", "chunk", "slots", "nc", "us/pass", "GB/s", "us/chunk/SM");
    for (int nc : {1, 8})
        for (uint32_t chunk : chunks) {
            const int nslots = (int)(RING / chunk);
            const int cpc = (int)((330.0e6 / sms) / chunk);
            float ms = 0;
            for (int rep = 0; rep < 2; rep++) {   // first launch warms up
                cudaEventRecord(e0);
                k_ring<<<sms, 288, RING>>>(buf, cpc, chunk, nslots, nc, passes);
                cudaEventRecord(e1);
                if (cudaEventSynchronize(e1) != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                cudaEventElapsedTime(&ms, e0, e1);
            }
            const double us = ms * 1e3 / passes, bytes = (double)cpc * chunk * sms;
            printf("%8u %6d %4d %10.2f %10.1f %12.3f\n", chunk, nslots, nc, us, bytes / us * 1e-3, us / cpc);
        }
    cudaFree(buf);
    return 0;
}
