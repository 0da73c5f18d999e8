/* TEST INFRASTRUCTURE ONLY (oracle/_ref build shim).
 * Stands in for the llama.cpp ggml headers that the reference's
 * extern/BitNet/src/ggml-bitnet-mad.cpp includes but that are absent from the
 * reference tree (.SUBMODULES.json: extern/BitNet/3rdparty/llama.cpp "missing").
 * Only the three names that translation unit needs are declared. */
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
enum ggml_type { GGML_TYPE_I2_S = 36 };
static inline size_t ggml_row_size(enum ggml_type t, int64_t ne) { (void)t; return (size_t)ne; }
#ifdef __cplusplus
extern "C" {
#endif
size_t quantize_i2_s(const float* src, void* dst, int64_t nrow, int64_t n_per_row, const float* quant_weights);
void ggml_vec_dot_i2_i8_s(int n, float* s, size_t bs, const void* vx, size_t bx, const void* vy, size_t by, int nrc);
#ifdef __cplusplus
}
#endif
