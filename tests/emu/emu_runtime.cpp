// emu_runtime.cpp — TEST INFRASTRUCTURE ONLY (see emu_cuda.h): the block / thread scheduler of the CUDA stand-in.
// A launch runs block after block on the calling host thread.  The threads of a block are fibers (a minimal x86-64 stack switch) resumed round-robin;
// __syncthreads() suspends a fiber until every fiber of the block that has not finished has reached a barrier, __syncwarp() the same
// within the 32 consecutive threads of a warp.  A kernel whose first thread finishes without ever reaching a barrier is taken to be
// barrier-free, and the remaining threads of that launch run as plain calls (no fiber switch).
#include "emu_cuda.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

namespace emu {
uint3 thread_idx, block_idx;
dim3 block_dim, grid_dim;

namespace {
const size_t kStack = 256 * 1024;
// callee-saved registers on the fiber's own stack, the stack pointer in the context: the System V x86-64 switch without the two
// signal-mask system calls of swapcontext (the kernels never touch the floating-point control words)
struct Context { void* sp = nullptr; };
extern "C" void emu_switch(Context* from, Context* to);
asm(".text\n.globl emu_switch\n.type emu_switch,@function\nemu_switch:\n"
    "  pushq %rbp\n  pushq %rbx\n  pushq %r12\n  pushq %r13\n  pushq %r14\n  pushq %r15\n"
    "  movq %rsp, (%rdi)\n  movq (%rsi), %rsp\n"
    "  popq %r15\n  popq %r14\n  popq %r13\n  popq %r12\n  popq %rbx\n  popq %rbp\n  ret\n"
    ".size emu_switch, .-emu_switch\n");

struct Fiber {
  Context ctx;
  char* stack = nullptr;
  uint3 tid;
  int state = 0;          // 0 = runnable, 1 = at a block barrier, 2 = at a warp barrier, 3 = finished
};
std::vector<Fiber> fibers;
Context scheduler;
int current = -1;
const std::function<void()>* body = nullptr;
std::vector<unsigned char> dyn_shared;
unsigned long long n_launches = 0;
bool saw_barrier = false;

void entry() {
  (*body)();
  fibers[current].state = 3;
  emu_switch(&fibers[current].ctx, &scheduler);
  std::abort();                                                  // a finished fiber is never resumed
}

void yield_with(int state) {
  saw_barrier = true;
  if (current < 0) { std::fprintf(stderr, "emu: barrier in a kernel that was classified barrier-free\n"); std::abort(); }
  fibers[current].state = state;
  emu_switch(&fibers[current].ctx, &scheduler);
  thread_idx = fibers[current].tid;
}

// NHSTEP_EMU_ORDER=reverse resumes the threads of a block (and walks the blocks of a grid) in the opposite order: a kernel whose
// result depends on the order in which threads or blocks run between barriers — a race — then gives different bits.
// NHSTEP_EMU_ORDER=random:<seed> reshuffles the order of the threads on every scheduler pass (a different interleaving around every
// barrier and every mbarrier poll), with a fixed seed for reproducibility.
unsigned long long rng_state = 0;
int order_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* o = std::getenv("NHSTEP_EMU_ORDER");
    mode = !o ? 0 : (std::strcmp(o, "reverse") == 0 ? 1 : (std::strncmp(o, "random:", 7) == 0 ? 2 : 0));
    if (mode == 2) rng_state = std::strtoull(o + 7, nullptr, 10) * 2654435761ull + 88172645463325252ull;
  }
  return mode;
}
unsigned next_random() {                                          // xorshift64
  rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
  return (unsigned)(rng_state >> 32);
}

void ensure_fibers(size_t n) {
  while (fibers.size() < n) {
    Fiber f;
    f.stack = (char*)mmap(nullptr, kStack, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (f.stack == MAP_FAILED) { std::perror("emu: mmap"); std::abort(); }
    fibers.push_back(f);
  }
}

void run_block_with_fibers(unsigned nthreads, int first) {
  // `first` threads have already run to completion as plain calls (only when first == 0 do we get here before any ran)
  ensure_fibers(nthreads);
  for (unsigned t = first; t < nthreads; ++t) {
    Fiber& f = fibers[t];
    f.tid = uint3{t % block_dim.x, (t / block_dim.x) % block_dim.y, t / (block_dim.x * block_dim.y)};
    f.state = 0;
    // first switch into the fiber: six zeroed callee-saved registers, then `ret` into entry() with the stack as after a call
    void** sp = (void**)(f.stack + kStack) - 8;
    for (int r = 0; r < 6; ++r) sp[r] = nullptr;
    sp[6] = (void*)&entry;
    sp[7] = nullptr;
    f.ctx.sp = sp;
  }
  for (unsigned t = 0; t < (unsigned)first; ++t) fibers[t].state = 3;
  unsigned done = first;
  const bool reversed = order_mode() == 1, shuffled = order_mode() == 2;
  std::vector<unsigned> perm(nthreads);
  for (unsigned q = 0; q < nthreads; ++q) perm[q] = q;
  while (done < nthreads) {
    bool progressed = false;
    if (shuffled) for (unsigned q = nthreads - 1; q > 0; --q) std::swap(perm[q], perm[next_random() % (q + 1)]);
    for (unsigned q = 0; q < nthreads; ++q) {
      const unsigned t = shuffled ? perm[q] : (reversed ? nthreads - 1 - q : q);
      Fiber& f = fibers[t];
      if (f.state != 0) continue;
      current = (int)t;
      thread_idx = f.tid;
      emu_switch(&scheduler, &f.ctx);
      progressed = true;
      if (f.state == 3) ++done;
      if (f.state == 4) f.state = 0;                              // polling an mbarrier: runnable again on the next pass
    }
    current = -1;
    // release a warp whose live threads are all at a warp barrier
    for (unsigned w = 0; w * 32 < nthreads; ++w) {
      bool all = true, any = false;
      for (unsigned t = w * 32; t < std::min(nthreads, w * 32 + 32); ++t) {
        if (fibers[t].state == 2) any = true;
        else if (fibers[t].state != 3) all = false;
      }
      if (any && all) { for (unsigned t = w * 32; t < std::min(nthreads, w * 32 + 32); ++t) if (fibers[t].state == 2) fibers[t].state = 0; progressed = true; }
    }
    // release the block when every live thread is at the block barrier
    bool all = true, any = false;
    for (unsigned t = 0; t < nthreads; ++t) {
      if (fibers[t].state == 1) any = true;
      else if (fibers[t].state != 3) all = false;
    }
    if (any && all) { for (unsigned t = 0; t < nthreads; ++t) if (fibers[t].state == 1) fibers[t].state = 0; progressed = true; }
    if (!progressed) { std::fprintf(stderr, "emu: deadlock (threads wait at different barriers)\n"); std::abort(); }
  }
}
}  // namespace

// ---- mbarrier + bulk tensor copy (the TMA sweep) ----------------------------------------------------------------------------
struct MBar { int expected = 0, pending = 0; long long tx = 0; unsigned phase = 0; };
static std::unordered_map<unsigned, MBar> mbars;

static void mbar_settle(MBar& b) {
  if (b.pending == 0 && b.tx == 0) { b.phase ^= 1u; b.pending = b.expected; }
}
static MBar& mbar_at(unsigned bar) {
  auto it = mbars.find(bar);
  if (it == mbars.end()) { std::fprintf(stderr, "emu: mbarrier %#x used before mbarrier.init\n", bar); std::abort(); }
  return it->second;
}
void mbar_init(unsigned bar, int count) { MBar b; b.expected = b.pending = count; mbars[bar] = b; }
void mbar_arrive(unsigned bar) { MBar& b = mbar_at(bar); b.pending -= 1; mbar_settle(b); }
void mbar_expect_tx(unsigned bar, unsigned bytes) { MBar& b = mbar_at(bar); b.tx += bytes; b.pending -= 1; mbar_settle(b); }
void mbar_wait(unsigned bar, unsigned parity) {
  // try_wait.parity succeeds once the phase of that parity has completed, i.e. the barrier's current phase has the other parity
  unsigned long long spins = 0;
  while (mbar_at(bar).phase == (parity & 1u)) {
    if (++spins > 5000000ull) { std::fprintf(stderr, "emu: mbarrier wait never satisfied (deadlock)\n"); std::abort(); }
    yield_with(4);
  }
}
void tma_load_3d(unsigned dst, const void* tensor_map, unsigned bar, int c0, int c1, int c2) {
  const unsigned long long* m = ((const CUtensorMap*)tensor_map)->opaque;
  const char* base = (const char*)m[0];
  const size_t es = m[1];
  const long long dims[3] = {(long long)m[2], (long long)m[3], (long long)m[4]};
  const size_t stride1 = m[5], stride2 = m[6];
  const unsigned box[3] = {(unsigned)m[7], (unsigned)m[8], (unsigned)m[9]};
  unsigned char* smem = dyn_shared.data();
  unsigned char* out = smem + (dst - (unsigned)(uintptr_t)smem);
  if (out < smem || out + (size_t)box[0] * box[1] * box[2] * es > smem + dyn_shared.size()) {
    std::fprintf(stderr, "emu: bulk tensor copy outside the launch's shared memory\n"); std::abort();
  }
  if ((c0 * (long long)es) % 16 != 0) { std::fprintf(stderr, "emu: bulk tensor copy whose box does not start on 16 bytes (illegal on sm_100)\n"); std::abort(); }
  for (unsigned z = 0; z < box[2]; ++z) for (unsigned y = 0; y < box[1]; ++y) for (unsigned x = 0; x < box[0]; ++x) {
    const long long i = c0 + (long long)x, j = c1 + (long long)y, k = c2 + (long long)z;
    unsigned char* o = out + (((size_t)z * box[1] + y) * box[0] + x) * es;
    if (i < 0 || j < 0 || k < 0 || i >= dims[0] || j >= dims[1] || k >= dims[2]) std::memset(o, 0, es);   // out-of-bounds elements read as zero
    else std::memcpy(o, base + (size_t)i * es + (size_t)j * stride1 + (size_t)k * stride2, es);
  }
  MBar& b = mbar_at(bar);
  b.tx -= (long long)box[0] * box[1] * box[2] * es;
  mbar_settle(b);
}

void syncthreads() { yield_with(1); }
void syncwarp() { yield_with(2); }
unsigned char* dynamic_shared() { return dyn_shared.data(); }
unsigned long long launches() { return n_launches; }

static std::unordered_map<const void*, int> max_dynamic_shared;
void set_max_dynamic_shared(const void* kernel, int bytes) { max_dynamic_shared[kernel] = bytes; }

void launch(dim3 grid, dim3 block, size_t shared_bytes, const std::function<void()>& thread_body, const void* kernel, const char* name) {
  ++n_launches;
  // the launch-configuration errors a GPU reports and a CPU would not: more than 48 KB of dynamic shared memory needs the opt-in
  // of cudaFuncSetAttribute on THIS instantiation, and grid extents are limited
  if (shared_bytes > 48 * 1024) {
    auto it = max_dynamic_shared.find(kernel);
    if (it == max_dynamic_shared.end() || (size_t)it->second < shared_bytes) {
      std::fprintf(stderr, "emu: %s launched with %zu bytes of dynamic shared memory without a sufficient cudaFuncSetAttribute\n", name, shared_bytes);
      std::abort();
    }
  }
  if (grid.y > 65535 || grid.z > 65535) { std::fprintf(stderr, "emu: %s: grid.y / grid.z above 65535\n", name); std::abort(); }
  const unsigned nthreads = block.x * block.y * block.z;
  if (nthreads == 0 || nthreads > 1024) { std::fprintf(stderr, "emu: invalid block of %u threads\n", nthreads); std::abort(); }
  if ((unsigned long long)grid.x * grid.y * grid.z == 0) { std::fprintf(stderr, "emu: empty grid\n"); std::abort(); }
  body = &thread_body;
  block_dim = block; grid_dim = grid;
  if (dyn_shared.size() < shared_bytes + 64) dyn_shared.resize(shared_bytes + 64);
  bool barrier_free = false;
  const unsigned long long nblocks = (unsigned long long)grid.x * grid.y * grid.z;
  for (unsigned long long q = 0; q < nblocks; ++q) {
    const unsigned long long b = order_mode() == 1 ? nblocks - 1 - q : q;
    const unsigned bx = (unsigned)(b % grid.x), by = (unsigned)((b / grid.x) % grid.y), bz = (unsigned)(b / ((unsigned long long)grid.x * grid.y));
    block_idx = uint3{bx, by, bz};
    std::memset(dyn_shared.data(), 0xFF, shared_bytes);          // uninitialised shared memory reads show up as NaN
    mbars.clear();
    if (barrier_free) {
      for (unsigned q2 = 0; q2 < nthreads; ++q2) {
        const unsigned t = order_mode() == 1 ? nthreads - 1 - q2 : q2;
        thread_idx = uint3{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
        thread_body();
      }
      continue;
    }
    saw_barrier = false;
    run_block_with_fibers(nthreads, 0);
    if (!saw_barrier) barrier_free = true;                        // same code for every block of the launch
  }
  body = nullptr;
}
}  // namespace emu

// device memory is host memory; fresh allocations are zeroed (what a fresh CUDA context hands out in practice).
// NHSTEP_EMU_GUARD=end | start puts every allocation against an inaccessible page on that side (the other side is only page-aligned),
// so a kernel that reads or writes past an allocation faults immediately — the out-of-bounds check of a memory checker, one side per run.
namespace {
struct Guarded { void* map; size_t map_bytes; };
std::unordered_map<void*, Guarded> guarded;
int guard_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* g = std::getenv("NHSTEP_EMU_GUARD");
    mode = !g ? 0 : (std::strcmp(g, "start") == 0 ? 1 : 2);
  }
  return mode;
}
}  // namespace

// ---- NHSTEP_EMU_SHM=1: allocations other processes can open (the multi-rank tests: one process per rank, peer stores, flag barriers)
namespace {
struct Shared { std::string name; size_t bytes; };
std::map<char*, Shared> shared_allocs;                 // own allocations by base address
std::unordered_map<void*, size_t> opened_maps;         // peers' allocations mapped here
bool shm_mode() { static const bool on = std::getenv("NHSTEP_EMU_SHM") != nullptr; return on; }
void unlink_all() { for (auto& a : shared_allocs) unlink(a.second.name.c_str()); }
}  // namespace

cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t* h, void* p) {
  std::memset(h, 0, sizeof(*h));
  if (!shm_mode()) { std::memcpy(h, &p, sizeof(p)); return cudaSuccess; }
  auto it = shared_allocs.upper_bound((char*)p);
  if (it == shared_allocs.begin()) return cudaErrorInvalidValue;
  --it;
  const size_t off = (char*)p - it->first;
  if (off >= it->second.bytes || it->second.name.size() > 47) return cudaErrorInvalidValue;
  std::memcpy(h->reserved, it->second.name.c_str(), it->second.name.size());
  std::memcpy(h->reserved + 48, &off, 8);
  std::memcpy(h->reserved + 56, &it->second.bytes, 8);
  return cudaSuccess;
}
cudaError_t cudaIpcOpenMemHandle(void** p, cudaIpcMemHandle_t h, unsigned) {
  if (!shm_mode()) { std::memcpy(p, &h, sizeof(*p)); return cudaSuccess; }
  size_t off, bytes;
  std::memcpy(&off, h.reserved + 48, 8);
  std::memcpy(&bytes, h.reserved + 56, 8);
  h.reserved[47] = 0;
  const int fd = open(h.reserved, O_RDWR);
  if (fd < 0) return cudaErrorInvalidValue;
  void* m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (m == MAP_FAILED) return cudaErrorInvalidValue;
  opened_maps[(char*)m + off] = bytes;
  *p = (char*)m + off;
  return cudaSuccess;
}
cudaError_t cudaIpcCloseMemHandle(void* p) {
  auto it = opened_maps.find(p);
  if (it != opened_maps.end()) opened_maps.erase(it);      // the mapping itself is left to process exit (offset unknown here)
  return cudaSuccess;
}

cudaError_t cudaMalloc(void** p, size_t bytes) {
  const size_t want = bytes ? bytes : 256;
  if (shm_mode()) {
    static int counter = 0;
    static bool registered = false;
    if (!registered) { std::atexit(unlink_all); registered = true; }
    char name[64];
    std::snprintf(name, sizeof name, "/dev/shm/nhstep_emu.%d.%d", (int)getpid(), counter++);
    const int fd = open(name, O_RDWR | O_CREAT | O_EXCL, 0600);
    if (fd < 0) return cudaErrorInvalidValue;
    if (ftruncate(fd, (off_t)want) != 0) { close(fd); unlink(name); return cudaErrorInvalidValue; }
    void* m = mmap(nullptr, want, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);      // fresh pages of the file read as zero
    close(fd);
    if (m == MAP_FAILED) { unlink(name); return cudaErrorInvalidValue; }
    if (std::getenv("NHSTEP_EMU_POISON")) std::memset(m, 0xFF, want);
    shared_allocs[(char*)m] = Shared{name, want};
    *p = m;
    return cudaSuccess;
  }
  static const bool poison = std::getenv("NHSTEP_EMU_POISON") != nullptr;     // fresh device memory = NaN bit patterns instead of zeros
  if (guard_mode() == 0) {
    void* q = nullptr;
    if (posix_memalign(&q, 256, want) != 0) return cudaErrorInvalidValue;
    std::memset(q, poison ? 0xFF : 0, want);
    *p = q;
    return cudaSuccess;
  }
  const size_t page = 4096, body = (want + page - 1) / page * page;
  char* map = (char*)mmap(nullptr, body + 2 * page, PROT_NONE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  if (map == MAP_FAILED) return cudaErrorInvalidValue;
  if (mprotect(map + page, body, PROT_READ | PROT_WRITE) != 0) return cudaErrorInvalidValue;
  // end mode: the last byte of the allocation is the last accessible byte (start then 16-byte aligned at most: sizes here are
  // multiples of 8, element accesses stay aligned); start mode: the first byte follows the guard page directly
  char* q = guard_mode() == 1 ? map + page : map + page + (body - want) / 16 * 16;
  guarded[q] = Guarded{map, body + 2 * page};
  *p = q;
  return cudaSuccess;
}
cudaError_t cudaFree(void* p) {
  if (!p) return cudaSuccess;
  auto sh = shared_allocs.find((char*)p);
  if (sh != shared_allocs.end()) {
    munmap(p, sh->second.bytes);
    unlink(sh->second.name.c_str());
    shared_allocs.erase(sh);
    return cudaSuccess;
  }
  auto it = guarded.find(p);
  if (it == guarded.end()) { std::free(p); return cudaSuccess; }
  munmap(it->second.map, it->second.map_bytes);
  guarded.erase(it);
  return cudaSuccess;
}

CUresult emu_cuTensorMapEncodeTiled(CUtensorMap* out, CUtensorMapDataType type, cuuint32_t rank, void* address, const cuuint64_t* dims,
                                    const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* element_strides,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill) {
  if (rank != 3 || element_strides[0] != 1 || element_strides[1] != 1 || element_strides[2] != 1) return 1;
  const size_t es = type == CU_TENSOR_MAP_DATA_TYPE_FLOAT64 ? 8 : 4;
  if ((uintptr_t)address % 16 != 0 || strides[0] % 16 != 0 || strides[1] % 16 != 0) return 1;     // the driver's alignment rules
  if (box[0] > 256 || box[1] > 256 || box[2] > 256 || (box[0] * es) % 16 != 0) return 1;
  unsigned long long* m = out->opaque;
  m[0] = (unsigned long long)(uintptr_t)address; m[1] = es;
  m[2] = dims[0]; m[3] = dims[1]; m[4] = dims[2];
  m[5] = strides[0]; m[6] = strides[1];
  m[7] = box[0]; m[8] = box[1]; m[9] = box[2];
  return CUDA_SUCCESS;
}
