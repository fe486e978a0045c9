// cusim.cpp — TEST INFRASTRUCTURE: fiber scheduler of the CPU emulator in cusim.h.
#include "cusim.h"

#include <sys/mman.h>

namespace cusim {
thread_local Fiber *cur = nullptr;
thread_local Block blk;
static thread_local void *sched_sp = nullptr;
static thread_local char *stack_pool = nullptr;
static thread_local Fiber *fiber_pool = nullptr;
static thread_local WarpSlot *warp_pool = nullptr;
static const size_t kStack = 128 * 1024;
static const int kMaxThreads = 1024;

extern "C" void cusim_switch(void **save_sp, void *load_sp);
asm(R"(
.text
.globl cusim_switch
.type cusim_switch,@function
cusim_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size cusim_switch,.-cusim_switch
)");

void yield_to_scheduler() { cusim_switch(&cur->sp, sched_sp); }

static void fiber_entry() {
    (*blk.body)();
    Fiber *f = cur;
    f->status = 3;
    blk.nlive--;
    if (blk.bar_arrived > 0 && blk.bar_arrived == blk.nlive) {  // exited threads do not take part in barriers
        blk.bar_arrived = 0;
        blk.bar_gen++;
    }
    yield_to_scheduler();
    abort();
}

static void init_fiber(Fiber *f, int i) {
    char *top = stack_pool + (size_t)(i + 1) * kStack;
    void **sp = (void **)(top - 64);
    for (int k = 0; k < 6; k++) sp[k] = nullptr;
    sp[6] = (void *)fiber_entry;
    sp[7] = nullptr;
    f->sp = sp;
    f->status = 0;
}

void launch(dim3 grid, dim3 block, const std::function<void()> &body) {
    int nthreads = (int)(block.x * block.y * block.z);
    if (nthreads < 1 || nthreads > kMaxThreads) { fprintf(stderr, "cusim: bad block size %d\n", nthreads); abort(); }
    if (!stack_pool) {
        stack_pool = (char *)mmap(nullptr, kStack * kMaxThreads, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (stack_pool == MAP_FAILED) { perror("cusim mmap"); abort(); }
        fiber_pool = new Fiber[kMaxThreads];
        warp_pool = new WarpSlot[kMaxThreads / 32];
    }
    static const char *seed_env = getenv("CUSIM_SEED");
    static thread_local unsigned long long rng = seed_env ? strtoull(seed_env, nullptr, 10) * 2654435761ULL + 1 : 0;
    const bool shuffle = seed_env != nullptr;
    Block saved = blk;
    Fiber *saved_cur = cur;
    blk.fib = fiber_pool;
    blk.warps = warp_pool;
    blk.bdim = block;
    blk.gdim = grid;
    blk.body = &body;
    blk.nthreads = nthreads;
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                blk.bid = uint3_{bx, by, bz};
                blk.nlive = nthreads;
                blk.bar_arrived = 0;
                blk.bar_gen = 0;
                int nwarps = (nthreads + 31) / 32;
                for (int w = 0; w < nwarps; w++) { warp_pool[w].arrived = 0; warp_pool[w].gen = 0; }
                for (int i = 0; i < nthreads; i++) {
                    Fiber *f = &fiber_pool[i];
                    init_fiber(f, i);
                    f->tid = uint3_{(unsigned)i % block.x, ((unsigned)i / block.x) % block.y, (unsigned)i / (block.x * block.y)};
                    f->lane = i & 31;
                    f->warp = i >> 5;
                }
                while (blk.nlive > 0) {
                    bool progress = false;
                    // CUSIM_SEED set: visit the fibers in a different pseudo-random order on every pass, so that
                    // code relying on threads running in index order (a data race on the GPU) fails here too
                    unsigned start = 0, stride = 1;
                    if (shuffle && (nthreads & (nthreads - 1)) == 0) {
                        rng = rng * 6364136223846793005ULL + 1442695040888963407ULL;
                        start = (unsigned)(rng >> 33) & (nthreads - 1);
                        stride = ((unsigned)(rng >> 20) & (nthreads - 1)) | 1u;
                    }
                    for (int ii = 0; ii < nthreads; ii++) {
                        int i = (int)((start + (unsigned)ii * stride) & (unsigned)(nthreads - 1));
                        if (stride == 1 && start == 0) i = ii;
                        Fiber *f = &fiber_pool[i];
                        if (f->status == 3) continue;
                        if (f->status == 1 && f->bar_gen == blk.bar_gen) continue;
                        if (f->status == 2 && f->warp_gen == warp_pool[f->warp].gen) continue;
                        f->status = 0;
                        cur = f;
                        cusim_switch(&sched_sp, f->sp);
                        progress = true;
                    }
                    if (!progress && blk.nlive > 0) {
                        fprintf(stderr, "cusim: deadlock in block (%u,%u,%u): %d live threads all waiting\n", bx, by, bz, blk.nlive);
                        int nb = 0, nw = 0;
                        for (int i = 0; i < nthreads; i++) { nb += fiber_pool[i].status == 1; nw += fiber_pool[i].status == 2; }
                        fprintf(stderr, "  %d at the block barrier, %d at a warp collective; warp-waiters:", nb, nw);
                        for (int i = 0, c = 0; i < nthreads && c < 40; i++) if (fiber_pool[i].status == 2) { fprintf(stderr, " %d", i); c++; }
                        fprintf(stderr, "\n  barrier-waiters:");
                        for (int i = 0, c = 0; i < nthreads && c < 40; i++) if (fiber_pool[i].status == 1) { fprintf(stderr, " %d", i); c++; }
                        fprintf(stderr, "\n");
                        abort();
                    }
                }
            }
    blk = saved;
    cur = saved_cur;
}
}  // namespace cusim
