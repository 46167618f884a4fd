// Block-wide bitonic sort (descending) of 64-bit keys, shared by the NMS and VOC-AP kernels.
#pragma once
#include "odk_common.cuh"

// One merge pass (k, j) over s[0..m): element `local` has global index base+local (direction bit).
__device__ __forceinline__ void odk_bitonic_pass(unsigned long long* s, int m, int base, int k, int j) {
  for (int t = threadIdx.x; t < (m >> 1); t += blockDim.x) {
    const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
    const int l = i | j;
    const unsigned long long x = s[i], y = s[l];
    const bool desc = (((base + i) & k) == 0);
    if (desc ? (x < y) : (x > y)) { s[i] = y; s[l] = x; }
  }
}

__device__ __forceinline__ void bitonic_sort_shared(unsigned long long* s, int m, int base, int k_from, int k_to) {
  for (int k = k_from; k <= k_to; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      odk_bitonic_pass(s, m, base, k, j);
      __syncthreads();
    }
}

// Sorts g[0..cnt) descending with the whole CTA.  `s` is a shared-memory scratch of `chunk` keys
// (a power of two).  Up to `chunk` keys the sort runs entirely in shared memory; above that, chunks
// are sorted in shared memory and the long-range merge passes run in global memory (L2).  g must have
// room for next_pow2(cnt) keys (the tail is used as zero padding).  Ends with a __syncthreads().
__device__ __forceinline__ void odk_block_sort_desc(unsigned long long* g, int cnt, unsigned long long* s, int chunk) {
  int n_pad = 2;
  while (n_pad < cnt) n_pad <<= 1;
  if (cnt > 1) {
    if (n_pad <= chunk) {
      for (int i = threadIdx.x; i < n_pad; i += blockDim.x) s[i] = (i < cnt) ? g[i] : 0ull;
      __syncthreads();
      bitonic_sort_shared(s, n_pad, 0, 2, n_pad);
      for (int i = threadIdx.x; i < cnt; i += blockDim.x) g[i] = s[i];
    } else {
      for (int i = cnt + threadIdx.x; i < n_pad; i += blockDim.x) g[i] = 0ull;
      __syncthreads();
      for (int base = 0; base < n_pad; base += chunk) {
        for (int i = threadIdx.x; i < chunk; i += blockDim.x) s[i] = g[base + i];
        __syncthreads();
        bitonic_sort_shared(s, chunk, base, 2, chunk);
        for (int i = threadIdx.x; i < chunk; i += blockDim.x) g[base + i] = s[i];
        __syncthreads();
      }
      for (int k = chunk << 1; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j >= chunk; j >>= 1) {
          odk_bitonic_pass(g, n_pad, 0, k, j);
          __syncthreads();
        }
        for (int base = 0; base < n_pad; base += chunk) {
          for (int i = threadIdx.x; i < chunk; i += blockDim.x) s[i] = g[base + i];
          __syncthreads();
          for (int j = chunk >> 1; j > 0; j >>= 1) {
            odk_bitonic_pass(s, chunk, base, k, j);
            __syncthreads();
          }
          for (int i = threadIdx.x; i < chunk; i += blockDim.x) g[base + i] = s[i];
          __syncthreads();
        }
      }
    }
  }
  __syncthreads();
}
