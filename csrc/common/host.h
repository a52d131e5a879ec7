// Host-side helpers shared by every translation unit: error handling for the C ABI, the driver entry
// point for cuTensorMapEncodeTiled (resolved at run time so the library has no link-time libcuda
// dependency and can be built on a GPU-less box), and tensor-map builders.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace b200 {

// Every extern "C" entry point returns 0 on success or a negative code; the message is kept here.
const char* last_error();
void set_error(const char* fmt, ...);
int sm_count();
// Per-device one-time configuration (cudaFuncSetAttribute is per device/context): launchers keep a
// `static bool configured[kMaxDevices]` indexed by current_device().
constexpr int kMaxDevices = 64;
int current_device();   // cudaGetDevice clamped into [0, kMaxDevices)

#define B200_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return -1;                                                                           \
    }                                                                                      \
  } while (0)

#define B200_REQUIRE(cond, ...)        \
  do {                                 \
    if (!(cond)) {                     \
      b200::set_error(__VA_ARGS__);    \
      return -2;                       \
    }                                  \
  } while (0)

// 2-D bf16 tensor map: `inner` contiguous elements per row, `rows` rows, row pitch `ld` elements.
// box = {box_inner, box_rows}; 128-byte swizzle (box_inner must be 64 bf16).
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t rows, uint64_t ld,
                      uint32_t box_inner, uint32_t box_rows);
// 128B-swizzled box of 128 bytes per row: 64 bf16 (elem_bytes 2) or 32 fp32 (elem_bytes 4) wide.
int make_tmap_2d_sw128(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t ld,
                       uint32_t box_rows);
// Same without swizzle (plain row-major box), any 2-byte or 4-byte element.
int make_tmap_2d_plain(CUtensorMap* out, const void* base, int elem_bytes, uint64_t inner, uint64_t rows, uint64_t ld,
                       uint32_t box_inner, uint32_t box_rows);

}  // namespace b200
