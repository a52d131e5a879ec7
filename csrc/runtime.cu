// Library runtime: error slot, device queries, TMA descriptor encoding, raw device allocations that
// can be exported to peer processes (CUDA IPC) for the NVLink comm kernels.
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "../include/b200_trainer.h"
#include "common/host.h"

namespace b200 {

static thread_local char g_err[512] = "";

const char* last_error() { return g_err; }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 0;
  return dev;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int encode_2d(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base, uint64_t inner,
                     uint64_t rows, uint64_t ld, uint32_t box_inner, uint32_t box_rows, CUtensorMapSwizzle sw) {
  EncodeTiledFn fn = encode_fn();
  B200_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base address must be 16-byte aligned");
  B200_REQUIRE((ld * elem_bytes) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes (ld=%llu)",
               (unsigned long long)ld);
  B200_REQUIRE(box_inner <= 256 && box_rows <= 256, "TMA box dims must be <= 256");
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {ld * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d (inner=%llu rows=%llu ld=%llu box=%ux%u)",
               (int)r, (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)ld, box_inner,
               box_rows);
  return 0;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t ld,
                      uint32_t box_inner, uint32_t box_rows) {
  B200_REQUIRE(box_inner == 64, "128B-swizzled bf16 boxes are 64 elements wide");
  return encode_2d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, inner, rows, ld, box_inner, box_rows,
                   CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_2d_sw128(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t ld,
                       uint32_t box_rows) {
  B200_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "sw128 tensor map: element size must be 2 or 4 bytes");
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  return encode_2d(out, dt, elem_bytes, base, inner, rows, ld, 128 / elem_bytes, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_2d_plain(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t ld,
                       uint32_t box_inner, uint32_t box_rows) {
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  return encode_2d(out, dt, elem_bytes, base, inner, rows, ld, box_inner, box_rows, CU_TENSOR_MAP_SWIZZLE_NONE);
}

}  // namespace b200

extern "C" {

const char* b200_last_error(void) { return b200::last_error(); }

int b200_version(void) { return 100; }

int b200_sm_count(void) { return b200::sm_count(); }

int b200_device_malloc(void** ptr, size_t bytes) {
  B200_CHECK_CUDA(cudaMalloc(ptr, bytes));
  B200_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
  return 0;
}

// Stream-ordered copy between any two mapped device addresses (own or peer allocations): the copy
// engines move it, no SM is involved -- the all-gathers of the sharded engine use this so that parameter
// traffic never competes with the persistent GEMM / attention CTAs for SM resources.
int b200_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  if (bytes == 0) return 0;
  B200_CHECK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int b200_device_free(void* ptr) {
  B200_CHECK_CUDA(cudaFree(ptr));
  return 0;
}

int b200_ipc_export(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  B200_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}

int b200_ipc_open(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int b200_ipc_close(void* ptr) {
  B200_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int b200_enable_peer_access(int peer_device) {
  int dev = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev == peer_device) return 0;
  int can = 0;
  B200_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  B200_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  B200_CHECK_CUDA(e);
  return 0;
}

}  // extern "C"
