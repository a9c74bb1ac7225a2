// emu_extra.h — TEST INFRASTRUCTURE ONLY (see emu_cuda.h): helpers the rewritten sources refer to.
#pragma once
[[noreturn]] static inline void emu_unsupported(const char* what) {
  std::fprintf(stderr, "emu: this instruction is not emulated: %s (run fused models with NHSTEP_TILE_SWEEP=1)\n", what);
  std::abort();
}

#include <sched.h>
// between two polls of another rank's flag: let that rank's process run when the host has fewer cores than ranks
static inline void emu_relax() { sched_yield(); }
