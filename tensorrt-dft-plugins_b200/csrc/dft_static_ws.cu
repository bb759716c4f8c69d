// dft_static_ws.cu -- trt_dft_exec_static_ws: run a plan that needs workspace without anybody
// allocating that workspace.
//
// Who needs this.  trt_dft_exec expects the caller to bring the plan's workspace.  Some callers
// have no way to get one at the moment they must launch:
//   * code that runs where cudaMalloc is forbidden or harmful -- inside a plugin's enqueue, in
//     the middle of a latency-critical serving loop (cudaMalloc synchronises the device), or on
//     a thread that is not allowed to touch the allocator at all;
//   * hosts that wire our kernels in behind an interface with no workspace argument (an inference
//     runtime that gave the op zero scratch bytes, a custom-op ABI with only inputs and outputs).
// For them the library keeps one fixed scratch region per GPU: a __device__ array, so it costs
// TRT_DFT_STATIC_WS_BYTES of device memory on every GPU on which this entry point has run -- from
// its first call there (with CUDA's lazy module loading the array's module is loaded by that call's
// address lookup, not when the library is opened) until the process exits -- and this entry point
// hands that region to the ordinary executor.  Nothing is allocated or freed here, by us or by any
// library we call.
//
// Rules that follow from sharing one region:
//   * plans whose workspace exceeds TRT_DFT_STATIC_WS_BYTES are refused (status
//     TRT_DFT_ERR_UNSUPPORTED and a message on stderr) -- use trt_dft_exec with your own buffer;
//   * two calls whose GPU work can overlap (different streams, or one stream per thread) race on
//     the region: keep static-workspace calls of one GPU on one stream, or order them with events;
//   * the region keeps whatever the previous call left in it.  No kernel of ours depends on its
//     initial contents: each call's first pass writes every byte its later passes read, so there
//     is nothing to clear between calls.
//
// Set-up.  A host function cannot use the array's name as a pointer: on the host side that name
// is only a handle for the runtime, not a device address.  The device address is looked up with
// cudaGetSymbolAddress the first time this entry point is called on a given GPU, with that GPU
// current, and cached per device ordinal.  That first call therefore does a little extra host
// work (the lookup also forces the module to be loaded on the device); make it before starting
// any stream capture on that GPU.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>

#include "../../include/trt_dft.h"

namespace {

// The scratch region itself.  Only device code and the symbol lookup below may name it.
static __device__ __align__(256) unsigned char g_static_ws[TRT_DFT_STATIC_WS_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_ws_addr[kMaxDevices] = {};      // device address of g_static_ws, per device ordinal
bool g_ws_ready[kMaxDevices] = {};

// Device address of this GPU's scratch region; nullptr (and *status set) on failure.
void* static_ws_for_current_device(int* status)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        cudaGetLastError();
        *status = TRT_DFT_ERR_CUDA;
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_ws_ready[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_ws);
        if (e != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            fprintf(stderr, "trt_dft_exec_static_ws: cudaGetSymbolAddress failed on device %d: %s\n", dev,
                    cudaGetErrorString(e));
            *status = TRT_DFT_ERR_CUDA;
            return nullptr;
        }
        g_ws_addr[dev] = p;
        g_ws_ready[dev] = true;
    }
    return g_ws_addr[dev];
}

}  // namespace

extern "C" int trt_dft_exec_static_ws(trt_dft_plan_t plan, const void* in, void* out, void* stream)
{
    if (!plan) return TRT_DFT_ERR_INVALID_ARG;
    const size_t need = trt_dft_workspace_bytes(plan);
    if (need > (size_t)TRT_DFT_STATIC_WS_BYTES) {
        fprintf(stderr,
                "trt_dft_exec_static_ws: this plan needs %zu bytes of workspace but the built-in scratch region "
                "holds %zu; call trt_dft_exec with a buffer of your own instead\n",
                need, (size_t)TRT_DFT_STATIC_WS_BYTES);
        return TRT_DFT_ERR_UNSUPPORTED;
    }
    void* ws = nullptr;
    if (need) {
        int status = TRT_DFT_OK;
        ws = static_ws_for_current_device(&status);
        if (!ws) return status;
    }
    // the ordinary executor: launches only, no allocation, no synchronisation
    return trt_dft_exec(plan, in, out, ws, stream);
}
