// H2D bandwidth: flat copy vs the strided 3-D copy upload_frames uses (640-byte rows into 672-byte pitch planes)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdint>
int main() {
    const int W = 640, H = 480, B = 128, pitch = 672, rows = 486;
    uint8_t *h, *d, *hout;
    cudaHostAlloc(&h, (size_t)W * H * B, cudaHostAllocDefault);
    cudaHostAlloc(&hout, (size_t)11 << 20, cudaHostAllocDefault);
    memset(h, 1, (size_t)W * H * B);
    cudaMalloc(&d, (size_t)pitch * rows * B + (11 << 20));
    cudaStream_t s, s2; cudaStreamCreate(&s); cudaStreamCreate(&s2);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode = 0; mode < 4; ++mode) {
        float best = 1e9;
        for (int it = 0; it < 8; ++it) {
            cudaEventRecord(a, s);
            for (int rep = 0; rep < 4; ++rep) {
                if (mode == 0 || mode == 2) cudaMemcpyAsync(d, h, (size_t)W * H * B, cudaMemcpyHostToDevice, s);
                else {
                    cudaMemcpy3DParms p; memset(&p, 0, sizeof(p));
                    p.srcPtr = make_cudaPitchedPtr(h, W, W, H);
                    p.dstPtr = make_cudaPitchedPtr(d, pitch, W, rows);
                    p.extent = make_cudaExtent(W, H, B);
                    p.kind = cudaMemcpyHostToDevice;
                    cudaMemcpy3DAsync(&p, s);
                }
                if (mode >= 2) cudaMemcpyAsync(hout, d + (size_t)pitch * rows * B, (size_t)10765328, cudaMemcpyDeviceToHost, s2);
            }
            cudaEventRecord(b, s);
            cudaStreamSynchronize(s); cudaStreamSynchronize(s2);
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (ms < best) best = ms;
        }
        printf("mode %d (%s%s): %.3f ms per 128 frames, %.1f GB/s\n", mode, (mode & 1) ? "3D strided" : "flat", mode >= 2 ? " + concurrent D2H" : "",
               best / 4, (double)W * H * B / (best / 4 * 1e-3) / 1e9);
    }
    return 0;
}
