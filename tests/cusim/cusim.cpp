// cusim.cpp — TEST INFRASTRUCTURE: fiber scheduler of the CPU emulator in cusim.h.
#include "cusim.h"

#include <mutex>
#include <sys/mman.h>

namespace cusim {
static std::mutex launch_mu;   // one kernel at a time across host threads: __shared__ variables are function-local statics here
thread_local Fiber *cur = nullptr;
thread_local Block *blkp = nullptr;
static thread_local void *sched_sp = nullptr;
static thread_local char *stack_pool = nullptr;
static thread_local Fiber *fiber_pool = nullptr;
static thread_local WarpSlot *warp_pool = nullptr;
static thread_local Block *block_pool = nullptr;
static const size_t kStack = 128 * 1024;
static const int kMaxThreads = 1024;       // per block
static const int kMaxFibers = 16384;       // all resident blocks of a cooperative launch
static const int kMaxBlocks = 256;
static thread_local int grid_arrived = 0, grid_blocks = 0;
static thread_local unsigned grid_gen = 0;
static thread_local int cl_arrived[kMaxBlocks];
static thread_local unsigned cl_gen[kMaxBlocks];

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

unsigned wait_change(volatile unsigned *addr, unsigned old) {
    Fiber *f = cur;
    while (*addr == old) {
        f->status = 5; f->wait_addr = addr; f->wait_old = old;
        yield_to_scheduler();
    }
    return *addr;
}

void sync_grid() {
    Fiber *f = cur;
    if (++grid_arrived == grid_blocks) { grid_arrived = 0; grid_gen++; return; }
    f->status = 4; f->bar_gen = grid_gen;
    yield_to_scheduler();
}

void *map_shared_rank(void *p, int rank) {
    Block &b = *blkp;
    Block &o = block_pool[b.cluster * b.csize + rank];
    return (char *)o.smem + ((char *)p - (char *)b.smem);
}

void sync_cluster() {
    Fiber *f = cur;
    Block &b = *blkp;
    if (++cl_arrived[b.cluster] == b.csize) { cl_arrived[b.cluster] = 0; cl_gen[b.cluster]++; return; }
    f->status = 6; f->bar_gen = cl_gen[b.cluster];
    yield_to_scheduler();
}

static void fiber_entry() {
    (*blkp->body)();
    Fiber *f = cur;
    Block &b = *blkp;
    f->status = 3;
    b.nlive--;
    if (b.bar_arrived > 0 && b.bar_arrived == b.nlive) {  // exited threads do not take part in barriers
        b.bar_arrived = 0;
        b.bar_gen++;
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

static void ensure_pools() {
    if (stack_pool) return;
    stack_pool = (char *)mmap(nullptr, kStack * kMaxFibers, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (stack_pool == MAP_FAILED) { perror("cusim mmap"); abort(); }
    fiber_pool = new Fiber[kMaxFibers];
    warp_pool = new WarpSlot[kMaxFibers / 32 + kMaxBlocks];
    block_pool = new Block[kMaxBlocks];
}

// Runs blocks [0, nblocks) of block_pool to completion, all resident at once.
static void run_resident(int nblocks, int nthreads) {
    static const char *seed_env = getenv("CUSIM_SEED");
    static thread_local unsigned long long rng = seed_env ? strtoull(seed_env, nullptr, 10) * 2654435761ULL + 1 : 0;
    const bool shuffle = seed_env != nullptr;
    const int total = nblocks * nthreads;
    const bool pow2 = (total & (total - 1)) == 0;
    int live = total;
    grid_arrived = 0; grid_blocks = nblocks; grid_gen = 0;
    while (live > 0) {
        bool progress = false;
        // CUSIM_SEED set: visit the fibers in a different pseudo-random order on every pass, so that
        // code relying on threads running in index order (a data race on the GPU) fails here too
        unsigned start = 0, stride = 1;
        if (shuffle && pow2) {
            rng = rng * 6364136223846793005ULL + 1442695040888963407ULL;
            start = (unsigned)(rng >> 33) & (total - 1);
            stride = ((unsigned)(rng >> 20) & (total - 1)) | 1u;
        }
        for (int ii = 0; ii < total; ii++) {
            int i = (stride == 1 && start == 0) ? ii : (int)((start + (unsigned)ii * stride) & (unsigned)(total - 1));
            Fiber *f = &fiber_pool[i];
            Block *b = &block_pool[f->block];
            if (f->status == 3) continue;
            if (f->status == 1 && f->bar_gen == b->bar_gen) continue;
            if (f->status == 2 && f->warp_gen == b->warps[f->warp].gen) continue;
            if (f->status == 4 && f->bar_gen == grid_gen) continue;
            if (f->status == 5 && *f->wait_addr == f->wait_old) continue;
            if (f->status == 6 && f->bar_gen == cl_gen[b->cluster]) continue;
            f->status = 0;
            cur = f;
            blkp = b;
            cusim_switch(&sched_sp, f->sp);
            if (f->status == 3) live--;
            progress = true;
        }
        if (!progress && live > 0) {
            fprintf(stderr, "cusim: deadlock: %d live threads all waiting (%d blocks)\n", live, nblocks);
            int nb = 0, nw = 0, ng = 0;
            for (int i = 0; i < total; i++) { nb += fiber_pool[i].status == 1; nw += fiber_pool[i].status == 2; ng += fiber_pool[i].status == 4; }
            fprintf(stderr, "  %d at a block barrier, %d at a warp collective, %d at the grid barrier; warp-waiters:", nb, nw, ng);
            for (int i = 0, c = 0; i < total && c < 40; i++) if (fiber_pool[i].status == 2) { fprintf(stderr, " %d", i); c++; }
            fprintf(stderr, "\n  block-barrier-waiters:");
            for (int i = 0, c = 0; i < total && c < 40; i++) if (fiber_pool[i].status == 1) { fprintf(stderr, " %d", i); c++; }
            fprintf(stderr, "\n");
            {   // who waits where: (status, source line of the rendezvous) -> number of fibers
                int seen = 0;
                for (int i = 0; i < total && seen < 12; i++) {
                    if (fiber_pool[i].status == 3) continue;
                    bool first = true;
                    for (int q = 0; q < i; q++)
                        if (fiber_pool[q].status == fiber_pool[i].status && fiber_pool[q].line == fiber_pool[i].line) { first = false; break; }
                    if (!first) continue;
                    int cnt = 0;
                    for (int q = i; q < total; q++) cnt += fiber_pool[q].status == fiber_pool[i].status && fiber_pool[q].line == fiber_pool[i].line;
                    fprintf(stderr, "  %d fiber(s) with status %d at source line %d (first: fiber %d)\n", cnt, fiber_pool[i].status, fiber_pool[i].line, i);
                    seen++;
                }
            }
            abort();
        }
    }
}

static void setup_block(Block *b, int bi, int first_fiber, uint3_ bid, dim3 grid, dim3 block, const std::function<void()> *body, void *smem) {
    int nthreads = (int)(block.x * block.y * block.z);
    int nwarps = (nthreads + 31) / 32;
    b->fib = fiber_pool + first_fiber;
    b->warps = warp_pool + (first_fiber / 32) + bi;   // disjoint even when nthreads is not a multiple of 32
    b->bdim = block; b->gdim = grid; b->body = body; b->nthreads = nthreads;
    b->bid = bid; b->nlive = nthreads; b->bar_arrived = 0; b->bar_gen = 0; b->smem = smem;
    b->cluster = bi; b->crank = 0; b->csize = 1;
    for (int w = 0; w < nwarps; w++) { b->warps[w].arrived = 0; b->warps[w].gen = 0; }
    for (int i = 0; i < nthreads; i++) {
        Fiber *f = &fiber_pool[first_fiber + i];
        init_fiber(f, first_fiber + i);
        f->block = bi;
        f->tid = uint3_{(unsigned)i % block.x, ((unsigned)i / block.x) % block.y, (unsigned)i / (block.x * block.y)};
        f->lane = i & 31;
        f->warp = i >> 5;
    }
}

void launch(dim3 grid, dim3 block, const std::function<void()> &body) {
    std::lock_guard<std::mutex> guard(launch_mu);
    int nthreads = (int)(block.x * block.y * block.z);
    if (nthreads < 1 || nthreads > kMaxThreads) { fprintf(stderr, "cusim: bad block size %d\n", nthreads); abort(); }
    ensure_pools();
    Block *saved_blk = blkp;
    Fiber *saved_cur = cur;
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                setup_block(&block_pool[0], 0, 0, uint3_{bx, by, bz}, grid, block, &body, nullptr);
                run_resident(1, nthreads);
            }
    blkp = saved_blk;
    cur = saved_cur;
}

void launch_cluster(dim3 grid, dim3 block, int cl, size_t smem_bytes, const std::function<void()> &body) {
    std::lock_guard<std::mutex> guard(launch_mu);
    int nthreads = (int)(block.x * block.y * block.z);
    int nblocks = (int)grid.x;
    if (nthreads < 1 || nthreads > kMaxThreads || nblocks < 1 || nblocks > kMaxBlocks || nblocks * nthreads > kMaxFibers || cl < 1 ||
        nblocks % cl || grid.y != 1 || grid.z != 1) {
        fprintf(stderr, "cusim: bad cluster launch %d x %d, cluster %d\n", nblocks, nthreads, cl);
        abort();
    }
    ensure_pools();
    Block *saved_blk = blkp;
    Fiber *saved_cur = cur;
    char *smem = (char *)calloc((size_t)nblocks, smem_bytes + 64);
    for (int b = 0; b < nblocks; b++) {
        setup_block(&block_pool[b], b, b * nthreads, uint3_{(unsigned)b, 0, 0}, grid, block, &body, smem + (size_t)b * (smem_bytes + 64));
        block_pool[b].cluster = b / cl; block_pool[b].crank = b % cl; block_pool[b].csize = cl;
        cl_arrived[b] = 0; cl_gen[b] = 0;
    }
    run_resident(nblocks, nthreads);
    free(smem);
    blkp = saved_blk;
    cur = saved_cur;
}

void launch_coop(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()> &body) {
    std::lock_guard<std::mutex> guard(launch_mu);
    int nthreads = (int)(block.x * block.y * block.z);
    int nblocks = (int)(grid.x * grid.y * grid.z);
    if (nthreads < 1 || nthreads > kMaxThreads || nblocks < 1 || nblocks > kMaxBlocks || nblocks * nthreads > kMaxFibers) {
        fprintf(stderr, "cusim: bad cooperative launch %d x %d\n", nblocks, nthreads);
        abort();
    }
    ensure_pools();
    Block *saved_blk = blkp;
    Fiber *saved_cur = cur;
    char *smem = (char *)calloc((size_t)nblocks, smem_bytes + 64);
    for (int b = 0; b < nblocks; b++)
        setup_block(&block_pool[b], b, b * nthreads, uint3_{(unsigned)b % grid.x, ((unsigned)b / grid.x) % grid.y, (unsigned)b / (grid.x * grid.y)},
                    grid, block, &body, smem + (size_t)b * (smem_bytes + 64));
    run_resident(nblocks, nthreads);
    free(smem);
    blkp = saved_blk;
    cur = saved_cur;
}
}  // namespace cusim
