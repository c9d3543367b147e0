// tools/tma_probe.cu -- standalone probe: 3-D FP64 TMA box load exactly as push_deposit_kernel issues it.
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdio.h>
#include <stdint.h>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int MODE>
__global__ void probe(const __grid_constant__ CUtensorMap tmap, double* out, int pw, int ph, int np, int c0, int c1) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* F = (double*)smem_raw;
    uint64_t* bar = (uint64_t*)(smem_raw + ((np * ph * pw * 8 + 127) & ~127));
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(np * ph * pw * 8) : "memory");
        if (MODE == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(smem_u32(F)), "l"(&tmap), "r"(c0), "r"(c1), "r"(0), "r"(smem_u32(bar)) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(smem_u32(F)), "l"(&tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(bar)), "r"(0) : "memory");
    for (int e = threadIdx.x; e < np * ph * pw; e += blockDim.x) out[e] = F[e];
}
#include <stdlib.h>
int main(int argc, char** argv) {
    const int pitch = 64, rows = 36, planes = 6, pw = 18, ph = 18;
    std::vector<double> h((size_t)pitch * rows * planes);
    for (size_t i = 0; i < h.size(); i++) h[i] = (double)i;
    double *d, *o;
    cudaMalloc(&d, h.size() * 8); cudaMalloc(&o, planes * ph * pw * 8);
    cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    struct V { int rank; int pw; CUtensorMapDataType dt; int esz; const char* name; };
    V vs[] = { {2, 18, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, "2d f64 18"}, {2, 16, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, "2d f64 16"},
               {3, 16, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, "3d f64 16"}, {2, 18, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, "2d u64 18"},
               {2, 36, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, "2d f32 36"}, {3, 36, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, "3d f32 36"},
               {3, 18, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, "3d u64 18"}, {3, 18, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 8, "3d f64 18"} };
    int which = argc > 1 ? atoi(argv[1]) : 0;
    V v = vs[which];
    {
        CUtensorMap tm;
        int sc = 8 / v.esz;
        cuuint64_t gdim[3] = { (cuuint64_t)pitch * sc, rows, planes }; cuuint64_t gstr[2] = { pitch * 8, (cuuint64_t)pitch * rows * 8 };
        cuuint32_t box[3] = { (cuuint32_t)v.pw, ph, planes }; cuuint32_t es[3] = { 1, 1, 1 };
        int np = v.rank == 3 ? planes : 1;
        int pwd = v.pw / sc;     // doubles per row
        CUresult r = ((encode_fn)fn)(&tm, v.dt, v.rank, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("%s encode rc=%d\n", v.name, (int)r);
        size_t smem = ((np * ph * pwd * 8 + 127) & ~127) + 16;
        if (v.rank == 3) probe<3><<<1, 128, smem>>>(tm, o, pwd, ph, np, 15 * sc, 1); else probe<2><<<1, 128, smem>>>(tm, o, pwd, ph, np, 15 * sc, 1);
        cudaError_t e = cudaDeviceSynchronize();
        printf("%s sync: %s\n", v.name, cudaGetErrorString(e));
        if (e == cudaSuccess) {
            std::vector<double> g(np * ph * pwd);
            cudaMemcpy(g.data(), o, g.size() * 8, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int p = 0; p < np; p++) for (int j = 0; j < ph; j++) for (int i = 0; i < pwd; i++)
                if (g[(p * ph + j) * pwd + i] != h[(size_t)p * pitch * rows + (size_t)(1 + j) * pitch + 15 + i]) bad++;
            printf("%s mismatches: %d\n", v.name, bad);
        } else return 1;
    }
    return 0;
}
