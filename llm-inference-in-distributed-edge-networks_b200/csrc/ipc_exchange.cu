// Payload hop between two PROCESSES (one per GPU) without any SM: the receiver's copy engine pulls
// the partner's payload through a CUDA-IPC mapping, and the two sides order themselves with
// stream memory operations on 32-bit flags (cuStreamWriteValue32 / cuStreamWaitValue32).
//
// Why: the fused scoring kernel is persistent and fills every SM's registers and shared memory; a
// send/recv KERNEL (NCCL's) has to wait for the gap between two of its launches, which showed as
// +7 % step time on 2 GPUs.  A DMA transfer plus stream memory operations needs no SM at all.
//
// No reference counterpart: the reference only simulates the hop in place
// (Experiments/Pythia-70M/pythia_model.py:176-188).
#include <cuda.h>
#include <string.h>

#include "common.cuh"

namespace splitq {
namespace {

typedef CUresult (*MemGetAddressRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
typedef CUresult (*StreamValue32Fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

template <typename Fn>
Fn driver_fn(const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
        return nullptr;
    return (Fn)p;
}

}  // namespace
}  // namespace splitq

using namespace splitq;

static_assert(sizeof(cudaIpcMemHandle_t) == SPLITQ_IPC_HANDLE_BYTES, "IPC handle size");

extern "C" int splitq_ipc_export(const void* dev_ptr, void* handle_out, size_t* offset_out) {
    SPLITQ_REQUIRE(dev_ptr && handle_out && offset_out, SPLITQ_ERR_ARG, "ipc_export: NULL argument");
    static MemGetAddressRangeFn range = driver_fn<MemGetAddressRangeFn>("cuMemGetAddressRange");
    SPLITQ_REQUIRE(range != nullptr, SPLITQ_ERR_CUDA, "ipc_export: cuMemGetAddressRange is unavailable");
    CUdeviceptr base = 0;
    size_t size = 0;
    CUresult r = range(&base, &size, (CUdeviceptr)dev_ptr);
    SPLITQ_REQUIRE(r == CUDA_SUCCESS, SPLITQ_ERR_CUDA, "ipc_export: cuMemGetAddressRange failed (%d)", (int)r);
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, (void*)base);
    if (e != cudaSuccess) (void)cudaGetLastError();  // reported below; keep it out of the caller's next CUDA call
    SPLITQ_REQUIRE(e == cudaSuccess, SPLITQ_ERR_CUDA,
                   "ipc_export: cudaIpcGetMemHandle failed: %s (the buffer must come from cudaMalloc, not from a "
                   "virtual-memory allocator such as PyTorch's expandable segments)",
                   cudaGetErrorString(e));
    memcpy(handle_out, &h, sizeof(h));
    *offset_out = (size_t)((CUdeviceptr)dev_ptr - base);
    return SPLITQ_OK;
}

extern "C" int splitq_ipc_open(const void* handle, size_t offset, void** dev_ptr_out) {
    SPLITQ_REQUIRE(handle && dev_ptr_out, SPLITQ_ERR_ARG, "ipc_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* base = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) (void)cudaGetLastError();
    SPLITQ_REQUIRE(e == cudaSuccess, SPLITQ_ERR_CUDA, "ipc_open: cudaIpcOpenMemHandle failed: %s (a handle cannot be "
                   "opened by the process that exported it, and the two devices need peer access)",
                   cudaGetErrorString(e));
    *dev_ptr_out = (char*)base + offset;
    return SPLITQ_OK;
}

extern "C" int splitq_ipc_close(void* dev_ptr, size_t offset) {
    SPLITQ_REQUIRE(dev_ptr, SPLITQ_ERR_ARG, "ipc_close: NULL pointer");
    SPLITQ_CUDA(cudaIpcCloseMemHandle((char*)dev_ptr - offset));
    return SPLITQ_OK;
}

extern "C" int splitq_stream_write_u32(void* dev_addr, uint32_t value, splitq_stream_t stream) {
    SPLITQ_REQUIRE(dev_addr && (reinterpret_cast<uintptr_t>(dev_addr) & 3u) == 0, SPLITQ_ERR_ALIGN,
                   "stream_write_u32: address must be a 4-byte aligned device address");
    static StreamValue32Fn fn = driver_fn<StreamValue32Fn>("cuStreamWriteValue32");
    SPLITQ_REQUIRE(fn != nullptr, SPLITQ_ERR_CUDA, "stream_write_u32: cuStreamWriteValue32 is unavailable");
    // default flags: earlier work of the stream (the pack kernel's stores) is visible before the flag is
    CUresult r = fn((CUstream)stream, (CUdeviceptr)dev_addr, value, CU_STREAM_WRITE_VALUE_DEFAULT);
    SPLITQ_REQUIRE(r == CUDA_SUCCESS, SPLITQ_ERR_CUDA, "stream_write_u32: cuStreamWriteValue32 failed (%d)", (int)r);
    return SPLITQ_OK;
}

extern "C" int splitq_stream_wait_geq_u32(const void* dev_addr, uint32_t value, splitq_stream_t stream) {
    SPLITQ_REQUIRE(dev_addr && (reinterpret_cast<uintptr_t>(dev_addr) & 3u) == 0, SPLITQ_ERR_ALIGN,
                   "stream_wait_geq_u32: address must be a 4-byte aligned device address");
    static StreamValue32Fn fn = driver_fn<StreamValue32Fn>("cuStreamWaitValue32");
    SPLITQ_REQUIRE(fn != nullptr, SPLITQ_ERR_CUDA, "stream_wait_geq_u32: cuStreamWaitValue32 is unavailable");
    // (int32)(*addr - value) >= 0: the step counters may wrap
    CUresult r = fn((CUstream)stream, (CUdeviceptr)dev_addr, value, CU_STREAM_WAIT_VALUE_GEQ);
    SPLITQ_REQUIRE(r == CUDA_SUCCESS, SPLITQ_ERR_CUDA, "stream_wait_geq_u32: cuStreamWaitValue32 failed (%d)", (int)r);
    return SPLITQ_OK;
}

extern "C" int splitq_copy_async(void* dst, const void* src, size_t bytes, splitq_stream_t stream) {
    SPLITQ_REQUIRE(dst && src, SPLITQ_ERR_ARG, "copy_async: NULL buffer");
    if (bytes == 0) return SPLITQ_OK;
    // unified addressing resolves local, peer and IPC-mapped pointers; the copy engines do the work
    SPLITQ_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
    return SPLITQ_OK;
}
