// Device side of programmatic dependent launch (host side + the rules: common.cuh, launch_pdl).
//   pdl_trigger(): griddepcontrol.launch_dependents — the next kernel on the stream may be scheduled once every CTA of this
//                  grid has executed it (or exited) and an SM has room; first statement of every PDL kernel.
//   pdl_wait():    griddepcontrol.wait — blocks until the predecessor grid has completed and its memory is visible; placed
//                  after the prologue (barrier init, TMEM allocation, tensor-map prefetch) and before ANY global access.
// Both are no-ops in a kernel launched without the programmatic-serialisation attribute.
#pragma once

namespace peft {
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
}  // namespace peft
