/* TEST INFRASTRUCTURE, force-included when oracle/build_ref.py compiles the reference's CUDA
 * lattice in place.  permutohedral.cu:433-435,444 spell `cudaMemcpyToSymbol((char*)&sym, ...)`,
 * the CUDA 3 idiom.  With today's headers that argument binds the runtime's
 * `template<class T> cudaMemcpyToSymbol(const T& symbol, ...)` with T = char*, which takes the
 * address of the TEMPORARY pointer -> cudaErrorInvalidSymbol at run time.  An exact-match
 * non-template overload restores the intended meaning (the pointer is the symbol's address)
 * without touching the reference source. */
#pragma once
#ifdef __cplusplus
#include <cuda_runtime.h>
static inline cudaError_t cudaMemcpyToSymbol(char* symbol, const void* src, size_t count,
                                             size_t offset = 0,
                                             cudaMemcpyKind kind = cudaMemcpyHostToDevice) {
  return ::cudaMemcpyToSymbol((const void*)symbol, src, count, offset, kind);
}
#endif
