// Shared helpers for the hg_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <math.h>

#include "../../include/hg_b200.h"

namespace hg {

// one half of a split operand: the bit pattern of an FP16 value (split.cu, knn_tc.cuh); buffers of them are only ever
// written by the split kernel and read by TMA
typedef unsigned short op16_t;

// ---- error plumbing: no exceptions cross the C ABI -------------------------
void set_error(const char* fmt, ...);
const char* get_error();
void count_launches(int n);              // kernels launched by this library (bench.py gpu_launches)
void prof_begin(cudaStream_t st);         // optional CUDA-event bracket around the fused kernel
void prof_end(cudaStream_t st);

#define HG_CHECK_CUDA(expr)                                                          \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            hg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,              \
                          cudaGetErrorString(_e));                                   \
            return HG_ERR_CUDA;                                                      \
        }                                                                            \
    } while (0)

#define HG_CHECK_LAUNCH() HG_CHECK_CUDA(cudaGetLastError())

#define HG_REQUIRE(cond, code, ...)                                                  \
    do {                                                                             \
        if (!(cond)) {                                                               \
            hg::set_error(__VA_ARGS__);                                              \
            return (code);                                                           \
        }                                                                            \
    } while (0)

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline int64_t min64(int64_t a, int64_t b) { return a < b ? a : b; }
__host__ __device__ inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

// workspace carving: every sub-buffer 256-byte aligned
struct Carver {
    char* base;
    size_t off;
    explicit Carver(void* p) : base(static_cast<char*>(p)), off(0) {}
    template <typename T>
    T* take(size_t count) {
        T* p = reinterpret_cast<T*>(base + off);
        off += (count * sizeof(T) + 255) & ~size_t(255);
        return p;
    }
};

// ---- geometry of the tensor-core operand buffer ------------------------------
constexpr int kRowPad = 256;   // operand rows padded to a multiple of the widest TMA box
constexpr int kColPad = 64;    // feature dim padded to one 128-byte swizzle atom of 16-bit values

inline int64_t ops_rows(int64_t n) { return round_up(n, kRowPad); }
inline int64_t ops_cols(int64_t d) { return round_up(d, kColPad); }
// 16-bit elements from one operand row to the next: 2 * d_pad (32 hi | 32 lo per k-block) plus a pad
// that keeps the row pitch off multiples of 2 KB (split.cu)
int64_t ops_row_elems(int64_t d_pad);

// ---- internal launchers shared across translation units ----------------------
// per-column epilogue parameters of the fused kernel: score = a_j + b_j * dot
struct ColParam { float a, b; };

int launch_split(const float* feats, int64_t n, int64_t d, op16_t* ops, float* sq_norms,
                 cudaStream_t st, bool dense = false, float* rho_max = nullptr);

}  // namespace hg
