// probe 2: (a) 1-D bulk copy, (b) tensor map in GLOBAL memory, (c) tensor map as __grid_constant__ param
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t* bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar) {
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(bar)), "r"(0) : "memory");
}
__global__ void k_bulk(const double* src, double* out, int n) {
    extern __shared__ __align__(128) unsigned char sm[];
    double* F = (double*)sm; uint64_t* bar = (uint64_t*)(sm + n * 8);
    if (threadIdx.x == 0) {
        bar_init(bar);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n * 8) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(F)), "l"(src), "r"(n * 8), "r"(smem_u32(bar)) : "memory");
    }
    __syncthreads();
    bar_wait(bar);
    for (int e = threadIdx.x; e < n; e += blockDim.x) out[e] = F[e];
}
__global__ void k_tm_global(const CUtensorMap* tm, double* out, int n, int c0, int c1) {
    extern __shared__ __align__(128) unsigned char sm[];
    double* F = (double*)sm; uint64_t* bar = (uint64_t*)(sm + n * 8);
    if (threadIdx.x == 0) {
        bar_init(bar);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n * 8) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(F)), "l"(tm), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
    }
    __syncthreads();
    bar_wait(bar);
    for (int e = threadIdx.x; e < n; e += blockDim.x) out[e] = F[e];
}
__global__ void k_tm_param(const __grid_constant__ CUtensorMap tm, double* out, int n, int c0, int c1) {
    extern __shared__ __align__(128) unsigned char sm[];
    double* F = (double*)sm; uint64_t* bar = (uint64_t*)(sm + n * 8);
    if (threadIdx.x == 0) {
        bar_init(bar);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n * 8) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(F)), "l"(&tm), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
    }
    __syncthreads();
    bar_wait(bar);
    for (int e = threadIdx.x; e < n; e += blockDim.x) out[e] = F[e];
}
int main(int argc, char** argv) {
    int which = argc > 1 ? atoi(argv[1]) : 0;
    const int pitch = 64, rows = 36, pw = 16, ph = 16;
    std::vector<double> h((size_t)pitch * rows);
    for (size_t i = 0; i < h.size(); i++) h[i] = (double)i;
    double *d, *o;
    cudaMalloc(&d, h.size() * 8); cudaMalloc(&o, pw * ph * 8);
    cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t ge = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    printf("entry point: %s q=%d fn=%p\n", cudaGetErrorString(ge), (int)q, fn);
    alignas(64) CUtensorMap tm;
    cuuint64_t gdim[2] = { pitch, rows }; cuuint64_t gstr[1] = { pitch * 8 };
    cuuint32_t box[2] = { pw, ph }; cuuint32_t es[2] = { 1, 1 };
    CUresult r = ((encode_fn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d; map words:", (int)r);
    for (int i = 0; i < 16; i++) printf(" %016llx", (unsigned long long)((uint64_t*)&tm)[i]);
    printf("\n");
    int n = pw * ph;
    size_t smem = n * 8 + 16;
    if (which == 0) k_bulk<<<1, 128, smem>>>(d + 16, o, n);
    if (which == 1) { CUtensorMap* dtm; cudaMalloc(&dtm, sizeof tm); cudaMemcpy(dtm, &tm, sizeof tm, cudaMemcpyHostToDevice); k_tm_global<<<1, 128, smem>>>(dtm, o, n, 16, 2); }
    if (which == 2) k_tm_param<<<1, 128, smem>>>(tm, o, n, argc > 2 ? atoi(argv[2]) : 16, 2);
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d sync: %s\n", which, cudaGetErrorString(e));
    if (e == cudaSuccess) {
        std::vector<double> g(n);
        cudaMemcpy(g.data(), o, n * 8, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int j = 0; j < ph; j++) for (int i = 0; i < pw; i++) {
            double want = which == 0 ? h[16 + j * pw + i] : h[(size_t)(2 + j) * pitch + (argc > 2 ? atoi(argv[2]) : 16) + i];
            if (g[j * pw + i] != want) bad++;
        }
        printf("variant %d mismatches %d\n", which, bad);
    }
    return 0;
}
