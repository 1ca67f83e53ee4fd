// Shared between unique.cu (caller-provided workspace) and its static-workspace variant.
#pragma once

#include "engine.cuh"

namespace dfm {

constexpr int kUniqueHeaderInts = 128;  // 512 bytes: the StepState the claim kernel counts in, padded
static_assert(sizeof(StepState) <= kUniqueHeaderInts * sizeof(int), "workspace header too small");

// Header cleared, id -> slot map set to -1.  Stream-ordered, no allocation.
int unique_workspace_init(void* workspace, long long V, cudaStream_t stream);
// Claim + per-position lookup + map restore.  `workspace` must have been initialised once.
// Ids outside [0, V) are treated as the nearest valid id (0 or V-1): nothing is read or written outside the map.
int unique_slots_run(const int* ids, int n, long long V, void* workspace, int* uniq_out, int* pos_slot_out,
                     int* count_out, cudaStream_t stream);

}  // namespace dfm
