// emu_extra.h — TEST INFRASTRUCTURE ONLY (see emu_cuda.h): helpers the rewritten sources refer to.
#pragma once
[[noreturn]] static inline void emu_unsupported(const char* what) {
  std::fprintf(stderr, "emu: this instruction is not emulated: %s (run fused models with NHSTEP_TILE_SWEEP=1)\n", what);
  std::abort();
}
