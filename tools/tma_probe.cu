// Debug probe: exercises cp.async.bulk.tensor tile loads of rank 2 and 4 (not part of the library).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int RANK>
__global__ void probe(const __grid_constant__ CUtensorMap tmap, float* out, int nelem, int c3) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* dst = reinterpret_cast<float*>(smem);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + nelem * 4);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(nelem * 4) : "memory");
        if (RANK == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(dst)), "l"(&tmap), "r"(smem_u32(bar)), "r"(0), "r"(c3) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                         ::"r"(smem_u32(dst)), "l"(&tmap), "r"(smem_u32(bar)), "r"(0), "r"(0), "r"(0), "r"(c3) : "memory");
    }
    uint32_t ok = 0;
    for (int spin = 0; spin < (1 << 22) && !ok; spin++)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(0) : "memory");
    if (!ok) { if (threadIdx.x == 0) printf("probe rank %d: timeout\n", RANK); return; }
    for (int i = threadIdx.x; i < nelem; i += blockDim.x) out[i] = dst[i];
}
int main() {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)p;
    printf("encode fn %p q=%d\n", p, (int)q);
    const int nres_pad = 128, kpad = 12, nf = 5;
    size_t n = (size_t)nf * 3 * kpad * nres_pad;
    std::vector<float> h(n); for (size_t i = 0; i < n; i++) h[i] = (float)i;
    float *d, *out; cudaMalloc(&d, n * 4); cudaMalloc(&out, 64 * kpad * 3 * 4);
    cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice);
    {   // rank 2: [rows = nf*3*kpad][cols = nres_pad], box 64 x 12
        CUtensorMap m; cuuint64_t gd[2] = {nres_pad, (cuuint64_t)nf * 3 * kpad}; cuuint64_t gs[1] = {nres_pad * 4};
        cuuint32_t box[2] = {64, 12}, es[2] = {1, 1};
        CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("rank2 encode %d\n", (int)r);
        probe<2><<<1, 128, 64 * 12 * 4 + 16>>>(m, out, 64 * 12, 12);
        cudaError_t e = cudaDeviceSynchronize(); printf("rank2 run: %s\n", cudaGetErrorString(e));
        std::vector<float> o(64 * 12); cudaMemcpy(o.data(), out, o.size() * 4, cudaMemcpyDeviceToHost);
        printf("rank2 out[0]=%g (want %g) out[64]=%g (want %g)\n", o[0], h[12 * nres_pad], o[64], h[13 * nres_pad]);
    }
    {   // rank 4 as the library builds it
        CUtensorMap m; cuuint64_t gd[4] = {nres_pad, kpad, 3, nf};
        cuuint64_t gs[3] = {nres_pad * 4, (cuuint64_t)kpad * nres_pad * 4, (cuuint64_t)3 * kpad * nres_pad * 4};
        cuuint32_t box[4] = {64, kpad, 3, 1}, es[4] = {1, 1, 1, 1};
        CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, d, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("rank4 encode %d\n", (int)r);
        probe<4><<<1, 128, 64 * kpad * 3 * 4 + 16>>>(m, out, 64 * kpad * 3, 2);
        cudaError_t e = cudaDeviceSynchronize(); printf("rank4 run: %s\n", cudaGetErrorString(e));
        std::vector<float> o(64 * kpad * 3); cudaMemcpy(o.data(), out, o.size() * 4, cudaMemcpyDeviceToHost);
        size_t base = (size_t)2 * 3 * kpad * nres_pad;
        printf("rank4 out[0]=%g (want %g) out[64]=%g (want %g) out[64*12]=%g (want %g)\n", o[0], h[base], o[64], h[base + nres_pad],
               o[64 * 12], h[base + (size_t)kpad * nres_pad]);
    }
    return 0;
}
