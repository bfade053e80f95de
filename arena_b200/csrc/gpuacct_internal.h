// Hooks between the translation units of libarena_gpuacct.so (not part of the C ABI).
#pragma once
#include "../../include/arena_gpuacct.h"

#include <cuda_runtime.h>

namespace arena_internal {

// arena_acct_init() done?  Selects the device and returns the library stream.
int ready(cudaStream_t *stream, int *sm_count);
void add_launches(unsigned long long n);
// Builds a snapshot handle around device columns that live in `block` (one cudaMalloc'ed
// allocation, owned by the handle from here on, also when this fails).
int adopt_device_block(void *block, size_t block_bytes, const arena_snapshot_view &device_view, arena_snapshot **out);

// cudaMalloc / cudaFree through a small cache of freed blocks (see gpuacct.cu).
cudaError_t device_alloc(void **out, size_t bytes);
void device_release(void *p, size_t bytes);
void device_pool_clear();
// Host -> device copy, enqueued on `st`: a large PAGEABLE source goes through page-locked slots filled
// by several host threads (json_device.cu), anything else is one cudaMemcpyAsync.
int copy_to_device(void *d_dst, const void *src, size_t bytes, cudaStream_t st);
// json_device.cu keeps its device workspace between calls; arena_acct_shutdown() drops it.
void free_loader_workspace();

}  // namespace arena_internal
