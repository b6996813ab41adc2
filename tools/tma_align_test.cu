// Does cp.async.bulk.tensor (tiled, u8, no interleave) accept an innermost box coordinate that is not a multiple of 16 bytes?
// (k_orient_describe wants 48 x 37 boxes at 4-byte-aligned x.)  Prints one line per tried x offset.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap map, int x, int y, uint8_t* out) {
    __shared__ __align__(128) uint8_t buf[48 * 37];
    __shared__ uint64_t bar;
    const unsigned sb = (unsigned)__cvta_generic_to_shared(&bar), sd = (unsigned)__cvta_generic_to_shared(buf);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sb) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sb), "r"(48 * 37) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(sd), "l"(&map), "r"(sb), "r"(x), "r"(y), "r"(0) : "memory");
    }
    __syncthreads();
    unsigned ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(sb), "r"(0) : "memory");
    } while (!ok);
    for (int i = threadIdx.x; i < 48 * 37; i += blockDim.x) out[i] = buf[i];
}
int main() {
    const int W = 1241, H = 376, P = 1280;
    std::vector<uint8_t> h((size_t)P * H);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)((i * 2654435761u) >> 24);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, 48 * 37);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    CUtensorMap m;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, 1}, strides[2] = {(cuuint64_t)P, (cuuint64_t)P * H};
    const cuuint32_t box[3] = {48, 37, 1}, es[3] = {1, 1, 1};
    CUresult r = ((EncodeTiledFn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d\n", (int)r);
    const int xs[] = {64, 68, 100, 101, 7, 1200};
    for (int x : xs) {
        k<<<1, 128>>>(m, x, 50, o);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("x=%d: %s\n", x, cudaGetErrorString(e)); return 1; }
        std::vector<uint8_t> g(48 * 37);
        cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int yy = 0; yy < 37; ++yy) for (int xx = 0; xx < 48; ++xx) {
            const uint8_t want = (x + xx < W) ? h[(size_t)(50 + yy) * P + x + xx] : 0;
            bad += g[yy * 48 + xx] != want;
        }
        printf("x=%d: ok, %d mismatches\n", x, bad);
    }
    return 0;
}
