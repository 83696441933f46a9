// TMA (cp.async.bulk.tensor) helpers shared by the Gram and Jacobian kernels: 2-D tensor maps over row-major FP64
// matrices, box = 16 columns x 32 rows (4 KB), CU_TENSOR_MAP_SWIZZLE_128B.  In a swizzled box the element
// (row r, column cc) sits at byte  r * 128 + ((cc / 2) ^ (r & 7)) * 16 + (cc & 1) * 8 ; boxes need 1024-byte alignment.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace dvf {

constexpr int TMA_BOX_COLS = 16, TMA_BOX_ROWS = 32, TMA_BOX_BYTES = TMA_BOX_COLS * TMA_BOX_ROWS * 8;

// host: tensor map over base[rows][cols] (row pitch `pitch_bytes`, multiple of 16), box 16 x 32, 128B swizzle, zero
// fill out of bounds.  The driver entry point is fetched through the runtime (no libcuda link dependency).
int make_tmap_2d(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes);
// host: tensor map over a rank-3 FP64 tensor (dims d0 fastest; byte strides of d1, d2), box 16 x 1 x 32 (one 4 KB
// swizzled box of 32 "rows" along d2), zero fill out of bounds
int make_tmap_3d(CUtensorMap* map, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                 uint64_t stride2_bytes);
// host: map over a vector of n doubles, box = 32 elements, no swizzle
int make_tmap_vec(CUtensorMap* map, const void* base, uint64_t n);

__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* map, int c0, int c1, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, int c0, int c1, int c2, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}
// consumer side of a TMA-fed ring: generic-proxy reads of a stage must be ordered before the producer's next
// async-proxy (TMA) write to it.  Without this fence B200 shows stale / torn stages a few times per thousand steps.
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void named_barrier_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// scratch from the device's stream-ordered pool; the pool's release threshold is raised once so that buffers are not
// handed back to the driver at every synchronisation (convenience entry points without a caller workspace)
int pool_alloc(void** p, size_t bytes, cudaStream_t st);

}  // namespace dvf
