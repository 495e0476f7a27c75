// tma_probe.cu -- development probe (not part of the product): which TMA box / coordinate
// configurations for FP64 tiles work on sm_100a.  usage: tma_probe <case>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int BW, int BH>
__global__ void probe_kernel(const __grid_constant__ CUtensorMap tm, int x, int y, double *out)
{
    extern __shared__ unsigned char raw[];
    unsigned char *smem = (unsigned char *)(((uintptr_t)raw + 127) & ~(uintptr_t)127);
    uint64_t *bar = (uint64_t *)(smem + ((BW * BH * 8 + 127) & ~127));
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(BW * BH * 8) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(smem)), "l"((uint64_t)&tm), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}\n"
                 ::"r"(smem_u32(bar)), "r"(0) : "memory");
    const double *s = (const double *)smem;
    for (int t = threadIdx.x; t < BW * BH; t += blockDim.x) out[t] = s[t];
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int BW, int BH>
int run(int nx, int ny, int x, int y, CUtensorMapL2promotion l2)
{
    std::vector<double> h((size_t)nx * ny);
    for (size_t k = 0; k < h.size(); ++k) h[k] = (double)k;
    double *d, *o;
    cudaMalloc(&d, h.size() * 8);
    cudaMalloc(&o, BW * BH * 8);
    cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    void *fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    cuuint64_t gdim[2] = {(cuuint64_t)nx, (cuuint64_t)ny}, gstr[1] = {(cuuint64_t)nx * 8};
    cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
    CUtensorMap tm;
    CUresult r = ((EncodeFn)fp)(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode=%d ", (int)r);
    if (r != CUDA_SUCCESS) return 2;
    int smem = BW * BH * 8 + 512;
    cudaFuncSetAttribute(probe_kernel<BW, BH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe_kernel<BW, BH><<<1, 128, smem>>>(tm, x, y, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("run=%s ", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<double> got(BW * BH);
    cudaMemcpy(got.data(), o, got.size() * 8, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int jj = 0; jj < BH; ++jj)
        for (int ii = 0; ii < BW; ++ii) {
            long gx = x + ii, gy = y + jj;
            double want = (gx >= 0 && gx < nx && gy >= 0 && gy < ny) ? (double)(gy * nx + gx) : 0.0;
            if (got[jj * BW + ii] != want) ++bad;
        }
    printf("mismatches=%d ", bad);
    return bad ? 3 : 0;
}

int main(int argc, char **argv)
{
    int c = argc > 1 ? atoi(argv[1]) : 0;
    int rc = -1;
    printf("case %d: ", c);
    switch (c) {
    case 0: rc = run<128, 4>(1024, 64, 0, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;     // aligned, pow2 box
    case 1: rc = run<130, 4>(1024, 64, 0, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;     // 130-wide box
    case 2: rc = run<130, 4>(1024, 64, 1, 3, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;     // odd x
    case 3: rc = run<128, 4>(1024, 64, 0, -1, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;    // negative y
    case 4: rc = run<130, 4>(70, 48, 0, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;       // box wider than tensor
    case 5: rc = run<130, 4>(1024, 64, 0, 0, CU_TENSOR_MAP_L2_PROMOTION_NONE); break;
    case 6: rc = run<128, 4>(1024, 64, 1, 0, CU_TENSOR_MAP_L2_PROMOTION_NONE); break;        // odd x, pow2 box
    case 7: rc = run<130, 4>(1024, 64, 2, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;     // even x (16B aligned)
    case 8: rc = run<132, 4>(1024, 64, 0, 0, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break;     // 32B-multiple box row
    case 9: rc = run<64, 8>(16448, 64, 16380, 60, CU_TENSOR_MAP_L2_PROMOTION_L2_256B); break; // OOB right/top
    }
    printf("rc=%d\n", rc);
    return rc;
}
