#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
struct Maps { CUtensorMap m[2]; };
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const __grid_constant__ Maps maps, int c0, int c1, int c2, int bw, int bh, uint8_t* out, int which) {
    __shared__ __align__(128) uint8_t buf[16384];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bw * bh) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_u32(buf)), "l"(reinterpret_cast<uint64_t>(&maps.m[which])), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(&bar)) : "memory");
    }
    __syncthreads();
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = buf[i];
}
int main(int argc, char** argv) {
    int c0 = argc > 1 ? atoi(argv[1]) : 0, c1 = argc > 2 ? atoi(argv[2]) : 0, bw = argc > 3 ? atoi(argv[3]) : 144, bh = argc > 4 ? atoi(argv[4]) : 46;
    const int pitch = 672, rows = 486, frames = 2;
    uint8_t* d; cudaMalloc(&d, (size_t)pitch * rows * frames);
    uint8_t* hbuf = (uint8_t*)malloc((size_t)pitch * rows * frames);
    for (size_t i = 0; i < (size_t)pitch * rows * frames; ++i) hbuf[i] = (uint8_t)(i * 7 + i / pitch);
    cudaMemcpy(d, hbuf, (size_t)pitch * rows * frames, cudaMemcpyHostToDevice);
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    printf("entry: %d %p q=%d\n", (int)e, fn, (int)q);
    Maps maps; memset(&maps, 0, sizeof(maps));
    cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)rows, (cuuint64_t)frames};
    cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)pitch * rows};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = ((EncodeFn)fn)(&maps.m[1], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode: %d\n", (int)r);
    uint8_t* o; cudaMalloc(&o, 16384);
    k<<<1, 128>>>(maps, c0, c1, 1, bw, bh, o, 1);
    e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    uint8_t ho[16384]; cudaMemcpy(ho, o, 16384, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int y = 0; y < bh; ++y) for (int x = 0; x < bw; ++x) {
        int gx = c0 + x, gy = c1 + y;
        uint8_t exp = (gx >= 0 && gx < pitch && gy >= 0 && gy < rows) ? hbuf[(size_t)pitch * rows + (size_t)gy * pitch + gx] : 0;
        if (ho[y * bw + x] != exp) ++bad;
    }
    printf("c0=%d c1=%d box=%dx%d mismatches=%d\n", c0, c1, bw, bh, bad);
    return 0;
}
