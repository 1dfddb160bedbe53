// Stand-alone probe of the TMA path the KLT kernel uses: one warp, one 48 B x 32 row box of a padded u8 plane.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <vector>
struct Maps { CUtensorMap m[4]; };
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int VARIANT>
__global__ void probe(const __grid_constant__ Maps maps, int level, int x, int y, unsigned* out)
{
    __shared__ __align__(128) unsigned tile[4][704];
    __shared__ __align__(8) unsigned long long bars[4];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long* bar = &bars[w];
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%1], %0;" ::"r"(1), "r"(smem_u32(bar)) : "memory");
        if (VARIANT & 1) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(1536), "r"(smem_u32(bar)) : "memory");
        if (VARIANT & 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(tile[w])),
                         "l"(reinterpret_cast<unsigned long long>(&maps.m[level])), "r"(smem_u32(bar)), "r"(x), "r"(y) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(tile[w])),
                         "l"(reinterpret_cast<unsigned long long>(&maps.m[level])), "r"(smem_u32(bar)), "r"(x), "r"(y) : "memory");
    }
    asm volatile("{\n\t.reg .pred P1;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(0) : "memory");
    __syncwarp();
    for (int i = lane; i < 384; i += 32) out[w * 384 + i] = tile[w][i];
}
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main()
{
    const int stride = 1344, rows = 440;
    std::vector<unsigned char> h((size_t)stride * rows);
    for (size_t i = 0; i < h.size(); i++) h[i] = (unsigned char)((i * 7 + i / stride) & 255);
    unsigned char* d; cudaMalloc(&d, h.size()); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry point: %s q=%d p=%p\n", cudaGetErrorString(e), (int)q, p);
    Maps maps; memset(&maps, 0, sizeof(maps));
    const cuuint64_t gdim[2] = { (cuuint64_t)stride, (cuuint64_t)rows }; const cuuint64_t gstr[1] = { (cuuint64_t)stride };
    const cuuint32_t box[2] = { 48, 32 }, es[2] = { 1, 1 };
    CUresult r = ((EncodeTiledFn)p)(&maps.m[1], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d, sizeof(Maps)=%zu alignof=%zu\n", (int)r, sizeof(Maps), alignof(Maps));
    unsigned* out; cudaMalloc(&out, 4 * 384 * 4);
    for (int variant = 0; variant < 4; variant++) {
        cudaMemset(out, 0, 4 * 384 * 4);
        if (variant == 0) probe<0><<<1, 128>>>(maps, 1, 96, 37, out);
        if (variant == 1) probe<1><<<1, 128>>>(maps, 1, 96, 37, out);
        if (variant == 2) probe<2><<<1, 128>>>(maps, 1, 96, 37, out);
        if (variant == 3) probe<3><<<1, 128>>>(maps, 1, 96, 37, out);
        e = cudaDeviceSynchronize();
        std::vector<unsigned> o(4 * 384);
        cudaMemcpy(o.data(), out, o.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int w = 0; w < 4; w++) for (int rr = 0; rr < 32; rr++) for (int b = 0; b < 48; b++) {
            const unsigned char got = ((unsigned char*)&o[w * 384])[rr * 48 + b];
            if (got != h[(size_t)(37 + rr) * stride + 96 + b]) bad++;
        }
        printf("variant %d (fence_init=%d, dst=%s): %s, mismatching bytes %d\n", variant, variant & 1, (variant & 2) ? "shared::cta" : "shared::cluster",
               cudaGetErrorString(e), bad);
        if (e != cudaSuccess) break;
    }
    return 0;
}
