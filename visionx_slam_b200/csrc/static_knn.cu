// static_knn.cu — vxorb_knn2_static: the split brute-force Hamming 2-NN of match.cu for callers that may not
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  vxorb_knn2 keeps its device
// staging (descriptor rows, row counts, result rows) and the per-split partial top-2 keys of launch_knn2 in buffers a
// handle cudaMallocs at create time and regrows on demand.  This entry point takes no handle and keeps all of that
// in ONE file-scope __device__ array of a fixed, stated size (kStaticBytes, ~1.7 MiB; every GPU that loads the
// library holds it until the process exits), so the call itself issues only copies, two kernel launches and a
// stream synchronize — no cudaMalloc, no cudaFree, no allocating library routine.
//
// The array's name is a host-side stand-in in host code, not a device address: the device address is looked up once
// per device with cudaGetSymbolAddress (that device current) and only that pointer is handed to copies and kernels.
// All calls share the array and it keeps the last call's contents.  Nothing has to start at zero: the row counts are
// rewritten by every call, and every partial-key slot and result row that a call reads is written earlier in the same
// call (knn2_kernel writes one slot per live query and split, empty splits included).  Calls that overlap on one GPU
// (two host threads) race on the array: serialise them.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "vxorb_internal.h"

namespace {

constexpr int kMaxQ = VXORB_STATIC_MAX_QUERY;
constexpr int kMaxT = VXORB_STATIC_MAX_TRAIN;
// layout of the array, every region 256-byte aligned
constexpr size_t kOffCounts = 0;                                   // int q_n, int t_n
constexpr size_t kOffQuery = 256;                                  // kMaxQ x 32 B
constexpr size_t kOffTrain = kOffQuery + (size_t)kMaxQ * 32;       // kMaxT x 32 B
constexpr size_t kOffKnn = kOffTrain + (size_t)kMaxT * 32;         // kMaxQ x vxorb_knn2_t
constexpr size_t kOffPart = kOffKnn + (size_t)kMaxQ * sizeof(vxorb_knn2_t);  // kMaxTrainSplits x kMaxQ x uint2
constexpr size_t kStaticBytes = kOffPart + (size_t)vxorb::kMaxTrainSplits * kMaxQ * sizeof(uint2);
static_assert(kOffTrain % 256 == 0 && kOffKnn % 256 == 0 && kOffPart % 256 == 0, "regions must stay aligned");
static_assert(kStaticBytes <= (2u << 20), "keep the resident scratch modest");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mutex;                          // guards the table below
unsigned char* g_scratch_ptr[kMaxDevices];   // device address of g_static_scratch on device i, or nullptr

int fail(char* err, size_t err_cap, int code, const char* fmt, ...) {
    if (err && err_cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Device address of the array on the CURRENT device; resolved once per device.
cudaError_t scratch_address(unsigned char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    return cudaSuccess;
}

}  // namespace

extern "C" int vxorb_knn2_static(const uint8_t* query, int nq, const uint8_t* train, int nt, vxorb_knn2_t* knn_out,
                                 char* err, size_t err_cap) {
    if (err && err_cap) err[0] = '\0';
    if (nq < 0 || nt < 0) return fail(err, err_cap, VXORB_E_INVALID, "negative row count (nq = %d, nt = %d)", nq, nt);
    if (nq > kMaxQ || nt > kMaxT)
        return fail(err, err_cap, VXORB_E_INVALID,
                    "vxorb_knn2_static holds at most %d query and %d train rows in its static scratch (got %d x %d); "
                    "use vxorb_knn2 for larger sets",
                    kMaxQ, kMaxT, nq, nt);
    if ((nq > 0 && (!query || !knn_out)) || (nt > 0 && !train))
        return fail(err, err_cap, VXORB_E_INVALID, "null argument");
    if (nq == 0) return VXORB_OK;

    unsigned char* base = nullptr;
    cudaError_t e = scratch_address(&base);
    if (e != cudaSuccess)
        return fail(err, err_cap, VXORB_E_CUDA, "cudaGetSymbolAddress(static scratch): %s", cudaGetErrorString(e));

    cudaStream_t s = cudaStreamLegacy;  // the default stream: always there, nothing to create
    const int counts[2] = {nq, nt};
    if ((e = cudaMemcpyAsync(base + kOffCounts, counts, sizeof(counts), cudaMemcpyHostToDevice, s)) != cudaSuccess ||
        (e = cudaMemcpyAsync(base + kOffQuery, query, (size_t)nq * 32, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
        (nt > 0 &&
         (e = cudaMemcpyAsync(base + kOffTrain, train, (size_t)nt * 32, cudaMemcpyHostToDevice, s)) != cudaSuccess))
        return fail(err, err_cap, VXORB_E_CUDA, "upload: %s", cudaGetErrorString(e));

    vxorb::MatchArgs a{};
    a.q_desc = base + kOffQuery;
    a.t_desc = base + kOffTrain;
    a.q_xy = a.t_xy = nullptr;
    a.q_n = reinterpret_cast<const int*>(base + kOffCounts);
    a.t_n = a.q_n + 1;
    a.n_stride = 1;
    a.q_pair_stride = kMaxQ;
    a.t_pair_stride = kMaxT;
    a.max_q = nq;
    a.max_t = nt;
    a.pairs = 1;
    a.radius = 0.0f;
    a.knn = reinterpret_cast<vxorb_knn2_t*>(base + kOffKnn);
    a.knn_stride = kMaxQ;
    a.part = reinterpret_cast<uint2*>(base + kOffPart);  // kMaxTrainSplits x max_q slots, max_q <= kMaxQ
    vxorb::launch_knn2(a, s);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return fail(err, err_cap, VXORB_E_CUDA, "launch: %s", cudaGetErrorString(e));
    if ((e = cudaMemcpyAsync(knn_out, base + kOffKnn, (size_t)nq * sizeof(vxorb_knn2_t), cudaMemcpyDeviceToHost, s)) !=
            cudaSuccess ||
        (e = cudaStreamSynchronize(s)) != cudaSuccess)
        return fail(err, err_cap, VXORB_E_CUDA, "read-back: %s", cudaGetErrorString(e));
    return VXORB_OK;
}
