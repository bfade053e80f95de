// Hooks between the translation units of libarena_gpuacct.so (not part of the C ABI).
#pragma once
#include "../../include/arena_gpuacct.h"

#include <cuda_runtime.h>

namespace arena_internal {

// The library keeps one device context per GPU ("slot"; slot 0 = arena_acct_init's device) and every
// entry point works on the calling thread's current slot.  The multi-device driver (multi.cu) selects
// a slot around the calls it makes for that device; entry points that take a snapshot handle select
// the handle's own slot themselves.
int current_slot();
void set_slot(int slot);
struct UseSlot {
    int prev;
    explicit UseSlot(int slot) : prev(current_slot()) { set_slot(slot); }
    ~UseSlot() { set_slot(prev); }
    UseSlot(const UseSlot &) = delete;
    UseSlot &operator=(const UseSlot &) = delete;
};
// arena_acct_init for the current slot (used by arena_acct_init_devices).
int init_slot(int device);
// In-process totals exchange between the first n slots (peer access instead of CUDA IPC).
int peer_exchange_in_process(int n);
// arena_acct_shutdown() tells the multi-device driver (multi.cu) that its slots are gone.
void multi_reset();

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
