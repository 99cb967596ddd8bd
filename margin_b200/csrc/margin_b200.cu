// margin_b200.cu — host side of the C ABI in include/margin_b200.h.
//
// One context = one CUDA device + one stream + grow-only device / pinned buffers, meant to be owned by one host
// thread (the reference runs one chunk per OpenMP thread, polish.c:476; thread t uses device t % nGPU).
// There is no CPU fallback anywhere in this file: without a usable device every entry point fails.
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <functional>
#include <iterator>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <unordered_set>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "../../include/margin_b200.h"
#include "pairhmm_tiles_f32.cuh"

namespace {

thread_local std::string g_err;

int fail(const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

constexpr int kMaxDevices = 64;
std::atomic<int> g_live_contexts[kMaxDevices];  // top-level contexts per device (sub-contexts of a pipeline excluded)
std::mutex g_compute_token[kMaxDevices];        // see margin_b200_posteriors_batch

#define CU(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

constexpr int kFwdThreads = 128;
constexpr int kBwdThreads = 128;
constexpr int kTileThreads = 128;

// ---- block caches
// cudaFree and cudaFreeHost synchronise the whole device, and cudaMallocHost pins pages under a process-wide lock: with
// `margin polish -t 16` every thread owns a context whose buffers are still growing during its first chunks, and each
// growth stalled all the other threads' kernels and copies (measured on the 2 Mb contig: `aligner` 0.06 s alone, 1.4-5.7 s
// with 16 threads; batch assembly in pinned memory 0.3 -> 3 s).  Device and pinned blocks therefore come in size classes
// (2^k x {1, 1.25, 1.5, 1.75}) and go back to a process-wide cache instead of to the driver; a block is handed to the
// cache only after the streams of the releasing thread's context are idle, so its next owner may use it at once.  The
// caches are trimmed (largest blocks first) only beyond a byte limit or when an allocation fails.
struct CachedBlock {
    void *p;
    bool part;  // a piece of a larger driver allocation (small size classes are allocated 4-16 at a time): never freed
};
struct BlockCache {
    std::mutex mu;
    std::multimap<size_t, CachedBlock> blocks;
    size_t bytes = 0;
};
// Never destroyed: a host program may still hand blocks back while the process is exiting (after static destructors ran).
BlockCache *const g_dev_cache = new BlockCache[kMaxDevices];
BlockCache &g_pin_cache = *new BlockCache;
constexpr size_t kDevCacheLimit = (size_t) 24 << 30, kPinCacheLimit = (size_t) 8 << 30;
void quiesce_own_streams();  // waits for the streams of the context this thread last worked for (defined below the context)

inline size_t block_class(size_t bytes) {
    size_t c = 4096;
    while (c < bytes) c <<= 1;
    if (c <= 65536) return c;
    const size_t h = c >> 1;  // h < bytes <= c
    for (int j = 1; j <= 3; j++)
        if (h + (h >> 2) * j >= bytes) return h + (h >> 2) * j;
    return c;
}

// Blocks of a small class come 4-16 to a driver call: sixteen threads' contexts, each with some seventy mostly small
// buffers, otherwise queue up in the driver's allocator during their first batches.
inline int blocks_per_allocation(size_t cls) {
    if (cls > ((size_t) 1 << 20)) return 1;
    return (int) std::max<size_t>(4, std::min<size_t>(16, ((size_t) 4 << 20) / cls));
}

inline bool cache_take(BlockCache &cache, size_t cls, CachedBlock &out) {
    std::lock_guard<std::mutex> lock(cache.mu);
    auto it = cache.blocks.find(cls);
    if (it == cache.blocks.end()) return false;
    out = it->second;
    cache.blocks.erase(it);
    cache.bytes -= cls;
    return true;
}

template <typename FreeFn>
inline void cache_trim(BlockCache &cache, size_t keep, FreeFn free_fn) {  // caller holds no lock
    std::vector<void *> victims;
    {
        std::lock_guard<std::mutex> lock(cache.mu);
        for (auto it = cache.blocks.end(); cache.bytes > keep && it != cache.blocks.begin();) {
            --it;
            if (it->second.part) continue;
            victims.push_back(it->second.p);
            cache.bytes -= it->first;
            it = cache.blocks.erase(it);
        }
    }
    for (void *p : victims) free_fn(p);
}

template <typename FreeFn>
inline void cache_put(BlockCache &cache, void *p, bool part, size_t cls, size_t limit, FreeFn free_fn) {
    quiesce_own_streams();
    bool over;
    {
        std::lock_guard<std::mutex> lock(cache.mu);
        cache.blocks.emplace(cls, CachedBlock{p, part});
        cache.bytes += cls;
        over = cache.bytes > limit;
    }
    if (over) cache_trim(cache, limit / 2, free_fn);
}

// A block of class `cls` from the cache, or from the driver (alloc_fn(&ptr, bytes) -> cudaError_t); the surplus blocks of
// a multi-block allocation go straight into the cache.
template <typename AllocFn, typename FreeFn>
inline bool block_take(BlockCache &cache, size_t cls, CachedBlock &out, AllocFn alloc_fn, FreeFn free_fn) {
    if (cache_take(cache, cls, out)) return true;
    const int n = blocks_per_allocation(cls);
    void *base = nullptr;
    cudaError_t e = alloc_fn(&base, cls * (size_t) n);
    if (e != cudaSuccess) {  // give the cached blocks back to the driver and try once more, for one block
        cudaGetLastError();
        cache_trim(cache, 0, free_fn);
        e = alloc_fn(&base, cls);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        out = CachedBlock{base, false};
        return true;
    }
    out = CachedBlock{base, n > 1};
    if (n > 1) {
        std::lock_guard<std::mutex> lock(cache.mu);
        for (int k = 1; k < n; k++) cache.blocks.emplace(cls, CachedBlock{static_cast<char *>(base) + (size_t) k * cls, true});
        cache.bytes += cls * (size_t) (n - 1);
    }
    return true;
}

// Grow-only device buffer.
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int dev = 0;
    bool part = false;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        release();
        if (cudaGetDevice(&dev) != cudaSuccess) return fail("cudaGetDevice failed");
        dev &= kMaxDevices - 1;
        // geometric growth: a slightly larger chunk must not re-allocate GBs
        const size_t want = block_class(bytes + bytes / 8 + 256);
        CachedBlock b;
        if (!block_take(g_dev_cache[dev], want, b, [](void **q, size_t n) { return cudaMalloc(q, n); }, [](void *q) { cudaFree(q); }))
            return fail("cudaMalloc(%zu bytes) failed", want);
        p = b.p;
        part = b.part;
        cap = want;
        return 0;
    }
    void release() {
        if (p) cache_put(g_dev_cache[dev], p, part, cap, kDevCacheLimit, [](void *q) { cudaFree(q); });
        p = nullptr;
        cap = 0;
    }
    template <typename T> T *as() { return reinterpret_cast<T *>(p); }
};

inline bool pinned_take(size_t want, CachedBlock &out) {
    return block_take(g_pin_cache, want, out, [](void **q, size_t n) { return cudaMallocHost(q, n); }, [](void *q) { cudaFreeHost(q); });
}

// Grow-only pinned host buffer.
struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    bool part = false;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        release();
        const size_t want = block_class(bytes + bytes / 8 + 256);
        CachedBlock b;
        if (!pinned_take(want, b)) return fail("cudaMallocHost(%zu bytes) failed", want);
        p = b.p;
        part = b.part;
        cap = want;
        return 0;
    }
    void release() {
        if (p) cache_put(g_pin_cache, p, part, cap, kPinCacheLimit, [](void *q) { cudaFreeHost(q); });
        p = nullptr;
        cap = 0;
    }
    template <typename T> T *as() { return reinterpret_cast<T *>(p); }
};

// The large scratch arrays of a posterior sweep — the forward values, the staged tuples — are only live while a context
// holds its device's compute token (first DP launch to final synchronisation), so the contexts of one device share
// ONE set: `margin polish -t 48` on one GPU needs the memory of one sweep, not of 48.  A context that runs without the
// token (sub-contexts of a pipelined call, the *_resident entry points) uses buffers of its own.
struct ScratchPool {
    DevBuf F, F32, shiftF, Fpriv, stage, stage2;
};
ScratchPool g_scratch[kMaxDevices];

}  // namespace

struct margin_b200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream_sym = nullptr;  // symbols travel and are packed beside the anchor checks and the band phase
    cudaEvent_t ev_sym = nullptr;       // ... and the DP kernels wait for this
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool have_model[2] = {false, false};
    bool have_params = false;
    margin_b200_params params{};
    mb::DevModel host_models[2];
    DevBuf d_models, d_rsm[2];

    // the uploaded batch
    int64_t n_aln = 0;
    std::vector<mb::Job> jobs;
    std::vector<int32_t> aln_first_slot;
    std::vector<int32_t> aln_first_job;
    int64_t total_diag = 0;  // entries in kl / cell_off
    int seg_slots = 0;       // capacity of the segment table (upper bound from the sequence lengths)
    int total_slots = 0;     // tiles of the current batch
    // tiling (pairhmm_tiles.cuh): tile length in diagonals (multiple of 10), warm-up diagonals on either side
    bool mode_sequential = false;
    int tile_len = 1600, warmup = 192, force_group = 0, tile_len_used = 0, bwd_slots = 0;
    bool tile_len_fixed = false;  // tile_len % 10 == 0: tiles start on total-probability samples
    std::vector<mb::Seg> segs;
    std::vector<mb::Tile> tiles;
    std::vector<int32_t> job_nseg, job_first_tile;
    DevBuf d_geo, d_tiles, d_bnd, d_slot_bnd_off, d_emit5[2], d_px, d_py, d_stage2;
    DevBuf d_chunks, d_job_first_chunk, d_chunk_sum, d_chunk_maxw, d_totals, d_band_out;
    DevBuf d_arena, d_offs, d_status;
    PinBuf h_arena;
    mb::Tile *p_tiles = nullptr;
    int64_t *p_ring_off = nullptr, *p_bnd_off = nullptr;
    int32_t *p_aln_first_slot = nullptr, *p_ids = nullptr;
    std::vector<uint64_t> keys[4], key_tmp;
    std::vector<int32_t> span_tmp;
    std::vector<uint8_t> tiny_mask;  // forward-only: jobs swept by tiny_forward_kernel (skipped by build_bins)
    PinBuf h_tiny_ids;
    DevBuf d_sym_err, d_tiny_ids, d_chunks_fwd, d_job_first_chunk_fwd, d_fwd_flags, d_fwd_counts;
    margin_b200_ctx *fwd_sub = nullptr;  // the few matrices too large for the register-resident kernel (fast forward path)
    std::vector<mb::BandChunk> chunks_fwd;
    std::vector<int32_t> job_first_chunk_fwd;
    // ---- pipelined batches (posteriors_pipelined): sub-contexts with their own stream and arenas
    std::vector<margin_b200_ctx *> subs;
    std::unique_ptr<margin_b200_model> user_models[2];  // as given to set_model, replayed into new sub-contexts
    int pipeline_k = 1;                                 // MARGIN_B200_PIPELINE: in-call pipeline of K runs (opt-in; 1 = off)
    bool pipeline_force = false;                        // MARGIN_B200_PIPELINE_FORCE: pipeline even beside other contexts
    int pipe_fill = 2;                                  // MARGIN_B200_PIPE_FILL: SM share a run's tiles are cut for, in 1/K
    int64_t pipeline_min_bytes = 8 << 20;               // MARGIN_B200_PIPELINE_MIN_BYTES of host input
    bool is_sub = false, last_pipelined = false, counted = false;
    bool take_token = false;                            // posteriors_resident: hold the device's compute token while the DP kernels run
    bool kernel_timing = false;                         // margin_b200_set_kernel_timing / MARGIN_B200_KERNEL_TIMING
    int pair_format = MARGIN_B200_PAIRS_TRIPLES;        // what the caller asked for
    bool packed_now = false;                            // ... and what this batch gets (sequence lengths permitting)
    int fmt_now = MARGIN_B200_PAIRS_TRIPLES;            // MARGIN_B200_PAIRS_* of this batch
    DevBuf d_esc_counts, d_esc_off, d_esc, d_esc_off_out;  // DELTA32: escapes per [list][slot], their scan, the escapes, per-alignment offsets
    PinBuf h_esc, h_esc_off;
    int64_t esc_total = 0, esc_list[3] = {0, 0, 0};
    DevBuf d_anchors_delta;
    int last_k = 0;
    PinBuf h_pipe_out[3], h_pipe_off;
    std::function<void()> after_h2d_enqueue;            // fired once the batch's host->device copies are queued
    int32_t *bins_ids_h = nullptr;
    size_t bins_ids_off = 0, arena_o_seams = 0, arena_o_flags = 0;
    int bins_n_ids = 0;
    std::vector<uint32_t> key_cnt;
    std::vector<mb::BandChunk> chunks;
    std::vector<int32_t> job_first_chunk;
    double v_thr = 0.0;
    bool have_switch = false;
    DevBuf d_symx, d_symy, d_anchors, d_jobs, d_kl, d_kr, d_cell_off, d_job_cells, d_job_maxw, d_job_nseg, d_err;
    DevBuf d_segs, d_seg_counts, d_seg_off, d_scan_tmp, d_slot_ring_off, d_ring, d_ids, d_aln_first_slot;
    DevBuf d_F, d_Fpriv, d_stage, d_out, d_off_out, d_fwd_out;
    // per-tile classification and the seam guard (build_tiles, guard_check_kernel)
    std::vector<uint32_t> chunk_base;    // first cell of every band chunk inside its job
    std::vector<int32_t> chunk_maxw;     // widest diagonal of every band chunk
    std::vector<int32_t> tile_maxw;      // widest diagonal of [f0, D] per tile as first cut (sizes its boundary sections)
    std::vector<int32_t> tile_binw;      // ... of its current [f0, D] (picks the kernel; grows when a repair extends the warm-up)
    std::vector<int32_t> tile_seg;       // index into segs of the traceback segment the tile belongs to
    std::vector<uint8_t> tile_rerun;     // repair pass: the tiles that run again (empty = all)
    std::vector<int32_t> job_gap_first;  // first entry of c->gaps per job (job_gap_first[nj] = gaps.size())
    int repair_factor = 8;               // a failed seam re-runs its tile with this many times the warm-up
    std::vector<uint8_t> job_force_seq;  // jobs a failed seam sent to the sequential sweep
    std::vector<mb::Seam> seams;
    mb::Seam *p_seams = nullptr;
    int32_t *p_guard_flags = nullptr;
    bool guard_on = true;                // MARGIN_B200_GUARD=0 turns the seam guard off
    double guard_tol = 1e-9, guard_floor = 100.0;
    int64_t wide_cells = 0, seq_cells = 0, priv_cells = 0, stage_cells = 0;
    int guard_failures = 0, guard_reruns = 0;
    bool guard_stats = false;            // MARGIN_B200_GUARD_STATS=1: histogram of the seam differences on stderr
    DevBuf d_guard_worst, d_class_cells;
    // FP32 fast mode (pairhmm_tiles_f32.cuh): lane-group tiles in FP32 with renormalisation, F as float4
    bool f32 = false;                    // margin_b200_set_precision / MARGIN_B200_PRECISION=f32
    bool warmup_env = false;             // MARGIN_B200_WARMUP pins the warm-up for both precisions
    double guard_tol32 = 1e-3;
    float v_thr_f = 0.f;
    DevBuf d_F32, d_shiftF, d_bnd32, d_emit_f[2], d_emit_f_ptrs;  // d_bnd32: boundary sections of FP32 tiles (float4 cells)
    PinBuf h_stage_in, h_small, h_out, h_off, h_gaps;
    // anchor-free stretches reported by the band kernel (sorted by job, start); see build_tiles
    std::vector<mb::AnchorGap> gaps;
    int gap_min = 32;
    bool gaps_overflow = false;
    int64_t n_anchors_total = 0;
    // after band_kernel
    std::vector<int64_t> job_cells;
    std::vector<int32_t> job_maxw;
    int64_t total_cells = 0;
    int64_t total_tuples = 0;
    int64_t list_tuples[3] = {0, 0, 0};
    int launches = 0;
    // per-kernel-class device times of the last call (CUDA events on the stream around each launch)
    std::vector<cudaEvent_t> ev_pool;
    std::vector<int> ev_class;  // class of the i-th (start, stop) pair
    int ev_used = 0;
};

namespace {
// Contexts whose streams exist.  A thread remembers the context it last worked for; the block caches wait for that
// context's streams before a block changes hands, unless the context has been destroyed in the meantime.
std::shared_mutex &g_live_mu = *new std::shared_mutex;  // leaked on purpose, like the caches
std::unordered_set<const margin_b200_ctx *> &g_live = *new std::unordered_set<const margin_b200_ctx *>;
thread_local margin_b200_ctx *tls_ctx = nullptr;

// Makes the context's device current and tells the block caches which streams this thread's work runs on.
inline cudaError_t enter_ctx(margin_b200_ctx *c) {
    tls_ctx = c;
    return cudaSetDevice(c->device);
}

void quiesce_own_streams() {
    if (!tls_ctx) return;
    std::shared_lock<std::shared_mutex> lock(g_live_mu);
    if (!g_live.count(tls_ctx)) {
        tls_ctx = nullptr;
        return;
    }
    if (tls_ctx->stream) cudaStreamSynchronize(tls_ctx->stream);
    if (tls_ctx->stream_sym) cudaStreamSynchronize(tls_ctx->stream_sym);
}
}  // namespace

namespace {

enum { K_BAND = 0, K_SCHEDULE, K_FORWARD, K_BACKWARD, K_SCAN, K_GATHER, K_OFFSETS, K_FINAL, K_GUARD, K_NCLASS };

// Records an event pair around one launch; the pairs are resolved by margin_b200_last_kernel_ms.
struct KTimer {
    margin_b200_ctx *c;
    int idx;
    KTimer(margin_b200_ctx *ctx, int cls) : c(ctx), idx(-1) {
        if (!c->kernel_timing) return;  // two event records per launch are driver calls the product path can do without
        if ((size_t) (2 * c->ev_used + 2) > c->ev_pool.size()) {
            cudaEvent_t a, b;
            cudaEventCreate(&a);
            cudaEventCreate(&b);
            c->ev_pool.push_back(a);
            c->ev_pool.push_back(b);
            c->ev_class.push_back(cls);
        }
        idx = c->ev_used++;
        c->ev_class[idx] = cls;
        cudaEventRecord(c->ev_pool[2 * idx], c->stream);
    }
    ~KTimer() {
        if (idx >= 0) cudaEventRecord(c->ev_pool[2 * idx + 1], c->stream);
    }
};

// Host-side phase trace (MARGIN_B200_TRACE=1): wall-clock microseconds between marks, printed to stderr.
struct Trace {
    bool on;
    std::chrono::steady_clock::time_point t0;
    std::string line;
    const void *tag;
    static std::chrono::steady_clock::time_point epoch() {
        static const std::chrono::steady_clock::time_point e = std::chrono::steady_clock::now();
        return e;
    }
    explicit Trace(const void *who = nullptr)
        : on(getenv("MARGIN_B200_TRACE") != nullptr), t0(std::chrono::steady_clock::now()), tag(who) {
        if (on) {
            char buf[64];
            snprintf(buf, sizeof(buf), " ctx=%04x t=%.0fus:", (unsigned) ((uintptr_t) who >> 4) & 0xffff,
                     std::chrono::duration<double, std::micro>(t0 - epoch()).count());
            line = buf;
        }
    }
    void mark(const char *what) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        char buf[64];
        snprintf(buf, sizeof(buf), " %s=%.0fus", what, std::chrono::duration<double, std::micro>(t1 - t0).count());
        line += buf;
        t0 = t1;
    }
    ~Trace() {
        if (on) fprintf(stderr, "[margin_b200 trace]%s\n", line.c_str());
    }
};

// ---- control transfers
// Small host<->device transfers (job tables, band counters, tile lists, offsets) do NOT go through cudaMemcpyAsync: a
// copy engine serves its queue in order, so a 30 KB read-back queued behind another context's (or another pipelined
// run's) 100 MB list transfer waits for all of it and stalls the kernels that depend on it (measured: band phase
// 0.26 -> 1.85 ms with a second context downloading).  A copy kernel moves them over the same PCIe link directly
// between device memory and pinned host memory (cudaMallocHost memory is device-accessible under UVA).
__global__ void ctl_copy_kernel(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src, size_t n_words) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    const size_t n4 = n_words / 4;
    if (i < n4) {
        if (((uintptr_t) dst | (uintptr_t) src) % 16 == 0) {
            reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(src)[i];
        } else {
            for (int k = 0; k < 4; k++) dst[4 * i + k] = src[4 * i + k];
        }
    }
    if (i < n_words - 4 * n4) dst[4 * n4 + i] = src[4 * n4 + i];
}

// bytes must be a multiple of 4; one of dst / src is pinned host memory, the other device memory
int ctl_copy(cudaStream_t stream, void *dst, const void *src, size_t bytes) {
    if (bytes == 0) return 0;
    if (bytes % 4) return fail("internal error: control copy of %zu bytes", bytes);
    const size_t words = bytes / 4, threads = std::max<size_t>(1, (words + 3) / 4);
    ctl_copy_kernel<<<(unsigned) ((threads + 255) / 256), 256, 0, stream>>>(reinterpret_cast<uint32_t *>(dst),
                                                                         reinterpret_cast<const uint32_t *>(src), words);
    CU(cudaGetLastError());
    return 0;
}

__global__ void ctl_fill_status_kernel(int32_t *status, int64_t n) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        status[2 * i] = 0x7fffffff;
        status[2 * i + 1] = 0;
    }
}

// bounds[l] = seg_off[l * S] for l = 0..3, written straight to pinned host memory
__global__ void ctl_bounds_kernel(const int64_t *__restrict__ seg_off, int S, int64_t *__restrict__ bounds) {
    if (threadIdx.x < 4) bounds[threadIdx.x] = seg_off[(size_t) threadIdx.x * S];
}

int check_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail("margin_b200: no usable CUDA device (%s); there is no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail("margin_b200: device %d out of range (0..%d)", device, count - 1);
    return 0;
}

}  // namespace

extern "C" {

const char *margin_b200_last_error(void) { return g_err.c_str(); }

int margin_b200_version(void) { return 100; }

int margin_b200_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    return count;
}

// Page-locked host memory for the caller's batch arrays: cudaMemcpyAsync from pageable memory is staged by the driver
// (synchronous, a few GB/s); a host program that assembles its batch in memory from here gets true asynchronous copies.
void *margin_b200_pinned_alloc(size_t bytes) {
    // a 64-byte header in front of the caller's memory remembers the block's size class for margin_b200_pinned_free
    const size_t want = block_class(bytes + 64);
    CachedBlock b;
    if (!pinned_take(want, b)) {
        fail("margin_b200_pinned_alloc: cudaMallocHost(%zu bytes) failed", want);
        return nullptr;
    }
    unsigned char *p = static_cast<unsigned char *>(b.p);
    reinterpret_cast<size_t *>(p)[0] = want;
    reinterpret_cast<size_t *>(p)[1] = b.part ? 1 : 0;
    return p + 64;
}

void margin_b200_pinned_free(void *p) {
    if (!p) return;
    unsigned char *base = static_cast<unsigned char *>(p) - 64;
    // the caller is done with the memory, and every library call that read or wrote it has synchronised its stream
    cache_put(g_pin_cache, base, reinterpret_cast<size_t *>(base)[1] != 0, reinterpret_cast<size_t *>(base)[0], kPinCacheLimit,
              [](void *q) { cudaFreeHost(q); });
}

int margin_b200_prewarm(void) { return margin_b200_prewarm_devices(kMaxDevices); }

int margin_b200_prewarm_devices(int n) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    for (int d = 0; d < count && d < n; d++) {
        if (cudaSetDevice(d) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) continue;
        // forces the kernel images in (lazy module loading would otherwise do it inside the first batch)
        cudaFuncAttributes fa;
        cudaFuncGetAttributes(&fa, mb::backward_tile_kernel<8, kTileThreads, true, false, true>);
        cudaFuncGetAttributes(&fa, mb::forward_tile_kernel<8, kTileThreads, false>);
        cudaFuncGetAttributes(&fa, mb::tiny_forward_kernel<16, kTileThreads, false>);
        cudaFuncGetAttributes(&fa, mb::backward_tile_kernel_f<8, kTileThreads, true, false, true>);
        cudaFuncGetAttributes(&fa, mb::forward_tile_kernel_f<8, kTileThreads, false>);
    }
    cudaGetLastError();
    return count;
}

int margin_b200_create(margin_b200_ctx **out, int device) {
    if (!out) return fail("margin_b200_create: null output pointer");
    *out = nullptr;
    if (check_device(device)) return 1;
    CU(cudaSetDevice(device));
    margin_b200_ctx *c = new margin_b200_ctx();
    c->device = device;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->stream_sym, cudaStreamNonBlocking));
    {
        std::unique_lock<std::shared_mutex> lock(g_live_mu);
        g_live.insert(c);
    }
    CU(cudaEventCreateWithFlags(&c->ev_sym, cudaEventDisableTiming));
    CU(cudaEventCreate(&c->ev0));
    CU(cudaEventCreate(&c->ev1));
    if (const char *e = getenv("MARGIN_B200_MODE")) c->mode_sequential = strcmp(e, "sequential") == 0;
    if (const char *e = getenv("MARGIN_B200_TILE_LEN")) {
        c->tile_len = std::max(10, atoi(e) / 10 * 10);
        c->tile_len_fixed = true;
    }
    if (const char *e = getenv("MARGIN_B200_PIPELINE")) c->pipeline_k = std::max(1, std::min(16, atoi(e)));
    if (const char *e = getenv("MARGIN_B200_KERNEL_TIMING")) c->kernel_timing = atoi(e) != 0;
    if (const char *e = getenv("MARGIN_B200_PIPELINE_FORCE")) c->pipeline_force = atoi(e) != 0;
    if (const char *e = getenv("MARGIN_B200_PIPE_FILL")) c->pipe_fill = std::max(1, std::min(16, atoi(e)));
    if (const char *e = getenv("MARGIN_B200_PIPELINE_MIN_BYTES")) c->pipeline_min_bytes = std::max(0ll, atoll(e));
    if (const char *e = getenv("MARGIN_B200_TILE_MAX")) c->tile_len = std::max(200, atoi(e) / 10 * 10);
    {
        // lane groups of 8 the backward tile kernel keeps resident: occupancy x SMs x groups per CTA
        int sms = 0, per_sm = 0;
        CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mb::backward_tile_kernel<8, kTileThreads, true, false, true>,
                                                         kTileThreads, 0));
        c->bwd_slots = sms * per_sm * (kTileThreads / 8);
    }
    if (const char *e = getenv("MARGIN_B200_WARMUP")) {
        c->warmup = std::max(1, atoi(e));
        c->warmup_env = true;
    }
    if (const char *e = getenv("MARGIN_B200_PRECISION")) margin_b200_set_precision(c, strcmp(e, "f32") == 0 ? MARGIN_B200_F32 : MARGIN_B200_F64);
    if (const char *e = getenv("MARGIN_B200_GUARD_TOL32")) c->guard_tol32 = atof(e);
    if (const char *e = getenv("MARGIN_B200_GAP_MIN")) c->gap_min = std::max(1, atoi(e));
    if (const char *e = getenv("MARGIN_B200_GUARD")) c->guard_on = atoi(e) != 0;
    if (const char *e = getenv("MARGIN_B200_GUARD_TOL")) c->guard_tol = atof(e);
    if (const char *e = getenv("MARGIN_B200_GUARD_STATS")) c->guard_stats = atoi(e) != 0;
    if (const char *e = getenv("MARGIN_B200_REPAIR_FACTOR")) c->repair_factor = std::max(2, atoi(e));
    c->counted = true;
    g_live_contexts[device & (kMaxDevices - 1)].fetch_add(1);
    *out = c;
    return 0;
}

void margin_b200_destroy(margin_b200_ctx *c) {
    if (!c) return;
    for (margin_b200_ctx *sub : c->subs) margin_b200_destroy(sub);
    c->subs.clear();
    if (c->fwd_sub) margin_b200_destroy(c->fwd_sub);
    c->fwd_sub = nullptr;
    if (c->counted) g_live_contexts[c->device & (kMaxDevices - 1)].fetch_sub(1);
    enter_ctx(c);
    cudaStreamSynchronize(c->stream);
    if (c->stream_sym) cudaStreamSynchronize(c->stream_sym);
    DevBuf *bufs[] = {&c->d_models, &c->d_rsm[0], &c->d_rsm[1], &c->d_symx, &c->d_symy, &c->d_anchors, &c->d_jobs,
                      &c->d_kl, &c->d_kr, &c->d_cell_off, &c->d_job_cells, &c->d_job_maxw, &c->d_job_nseg, &c->d_err,
                      &c->d_segs, &c->d_seg_counts, &c->d_seg_off, &c->d_scan_tmp, &c->d_slot_ring_off, &c->d_ring,
                      &c->d_ids, &c->d_aln_first_slot, &c->d_F, &c->d_Fpriv, &c->d_guard_worst, &c->d_class_cells, &c->d_esc_counts, &c->d_esc_off, &c->d_esc, &c->d_esc_off_out, &c->d_F32, &c->d_shiftF, &c->d_bnd32, &c->d_emit_f[0], &c->d_emit_f[1],
                      &c->d_emit_f_ptrs, &c->d_stage, &c->d_out, &c->d_off_out, &c->d_fwd_out,
                      &c->d_geo, &c->d_tiles, &c->d_bnd, &c->d_slot_bnd_off, &c->d_emit5[0], &c->d_emit5[1],
                      &c->d_px, &c->d_py, &c->d_stage2, &c->d_chunks, &c->d_job_first_chunk, &c->d_chunk_sum, &c->d_chunk_maxw,
                      &c->d_totals, &c->d_band_out, &c->d_anchors_delta, &c->d_sym_err, &c->d_tiny_ids, &c->d_fwd_flags, &c->d_fwd_counts, &c->d_chunks_fwd, &c->d_job_first_chunk_fwd, &c->d_arena, &c->d_offs, &c->d_status};
    for (DevBuf *b : bufs) b->release();
    c->h_stage_in.release();
    c->h_small.release();
    c->h_gaps.release();
    c->h_arena.release();
    c->h_out.release();
    c->h_off.release();
    for (PinBuf &b : c->h_pipe_out) b.release();
    c->h_tiny_ids.release();
    c->h_esc.release();
    c->h_esc_off.release();
    c->h_pipe_off.release();
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaStreamSynchronize(c->stream_sym);
    cudaEventDestroy(c->ev_sym);
    {
        std::unique_lock<std::shared_mutex> lock(g_live_mu);  // no thread is waiting on these streams from here on
        g_live.erase(c);
    }
    if (tls_ctx == c) tls_ctx = nullptr;
    cudaStreamDestroy(c->stream_sym);
    cudaStreamDestroy(c->stream);
    delete c;
}

int margin_b200_set_model(margin_b200_ctx *c, int slot, const margin_b200_model *m) {
    if (!c || !m) return fail("margin_b200_set_model: null argument");
    if (slot < 0 || slot > 1) return fail("margin_b200_set_model: slot must be 0 or 1");
    if (m->rle && m->max_rl != MARGIN_B200_MAX_RL) return fail("margin_b200_set_model: max_rl must be %d", MARGIN_B200_MAX_RL);
    CU(enter_ctx(c));
    if (!c->is_sub) {
        if (c->user_models[slot].get() != m) c->user_models[slot].reset(new margin_b200_model(*m));
        for (margin_b200_ctx *sub : c->subs)
            if (margin_b200_set_model(sub, slot, m)) return 1;
    }
    mb::DevModel &d = c->host_models[slot];
    memset(&d, 0, sizeof(d));
    memcpy(d.t, m->t, sizeof(d.t));
    memcpy(d.match, m->match, sizeof(d.match));
    for (int b = 0; b < 4; b++) {
        d.gap_x[b] = m->gap_x[b];
        d.gap_y[b] = m->gap_y[b];
    }
    d.gap_x[4] = -1.386294361;  // ambiguous symbol, impl/stateMachine.c:363-368
    d.gap_y[4] = -1.386294361;
    const int tx[3] = {MB_T_GAP_OPEN_X, MB_T_GAP_EXTEND_X, MB_T_GAP_SWITCH_TO_X};
    const int ty[3] = {MB_T_GAP_OPEN_Y, MB_T_GAP_EXTEND_Y, MB_T_GAP_SWITCH_TO_Y};
    for (int b = 0; b < 5; b++) {
        for (int k = 0; k < 3; k++) {
            d.ex_t[b][k] = d.gap_x[b] + m->t[tx[k]];  // the reference's (eP + tP), pairwiseAligner.c:313
            d.ey_t[b][k] = d.gap_y[b] + m->t[ty[k]];
        }
    }
    d.end_p[0][0] = m->t[MB_T_MATCH_CONTINUE];  // impl/stateMachine.c:532-545
    d.end_p[0][1] = m->t[MB_T_MATCH_FROM_GAP_X];
    d.end_p[0][2] = m->t[MB_T_MATCH_FROM_GAP_Y];
    d.end_p[1][0] = (m->t[MB_T_GAP_OPEN_X] + m->t[MB_T_GAP_OPEN_Y]) / 2.0;  // :547-560
    d.end_p[1][1] = m->t[MB_T_GAP_EXTEND_X];
    d.end_p[1][2] = m->t[MB_T_GAP_EXTEND_Y];
    d.rle = m->rle;
    d.max_rl = MARGIN_B200_MAX_RL;
    const size_t nr = (size_t) 4 * MARGIN_B200_MAX_RL * MARGIN_B200_MAX_RL;
    std::vector<double> scaled(nr);
    for (size_t i = 0; i < nr; i++) scaled[i] = 2.3025 * m->rsm[i];  // impl/stateMachine.c:735-736
    if (c->d_rsm[slot].ensure(nr * sizeof(double))) return 1;
    CU(cudaMemcpyAsync(c->d_rsm[slot].p, scaled.data(), nr * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    d.rsm_scaled = c->d_rsm[slot].as<double>();
    // folded match emission of the tile kernels, indexed by packed symbols (pairhmm_tiles.cuh): sub + 2.3025*rsm,
    // one rounding, as impl/stateMachine.c:733-738 forms it
    {
        const int R = MARGIN_B200_MAX_RL, P = 5 * R;
        std::vector<double> ep((size_t) P * P, 0.0);
        for (int xb = 0; xb < 5; xb++)
            for (int yb = 0; yb < 5; yb++) {
                const double sub = (xb >= 4 || yb >= 4) ? -2.772588722 : m->match[xb * 4 + yb];
                const int tb = xb >= 4 ? 0 : xb;
                for (int rx = 0; rx < R; rx++)
                    for (int ry = 0; ry < R; ry++) {
                        double v = sub;
                        if (m->rle) {
                            const double r = scaled[((size_t) tb * R + rx) * R + ry];
                            v = sub + r;
                        }
                        ep[(size_t) (xb * R + rx) * P + (yb * R + ry)] = v;
                    }
            }
        if (c->d_emit5[slot].ensure(ep.size() * sizeof(double))) return 1;
        CU(cudaMemcpyAsync(c->d_emit5[slot].p, ep.data(), ep.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        // the same table in FP32 for the fast tile kernels, and the two tables' addresses
        std::vector<float> epf(ep.begin(), ep.end());
        if (c->d_emit_f[slot].ensure(epf.size() * sizeof(float)) || c->d_emit_f_ptrs.ensure(2 * sizeof(float *))) return 1;
        CU(cudaMemcpyAsync(c->d_emit_f[slot].p, epf.data(), epf.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
        const float *ptrs[2] = {c->d_emit_f[0].as<float>(), c->d_emit_f[1].as<float>()};
        CU(cudaMemcpyAsync(c->d_emit_f_ptrs.p, ptrs, sizeof(ptrs), cudaMemcpyHostToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        d.emit_packed = c->d_emit5[slot].as<double>();
    }
    if (c->d_models.ensure(2 * sizeof(mb::DevModel))) return 1;
    CU(cudaMemcpyAsync(c->d_models.p, c->host_models, 2 * sizeof(mb::DevModel), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->have_model[slot] = true;
    return 0;
}

int margin_b200_set_params(margin_b200_ctx *c, const margin_b200_params *p) {
    if (!c || !p) return fail("margin_b200_set_params: null argument");
    // the prerequisites asserted at impl/pairwiseAligner.c:714-718
    if (p->traceback_diagonals < 1) return fail("traceBackDiagonals must be >= 1");
    if (p->diagonal_expansion < 0 || p->diagonal_expansion % 2 != 0) return fail("diagonalExpansion must be even and >= 0");
    if (p->min_diags_between_traceback < 2) return fail("minDiagsBetweenTraceBack must be >= 2");
    if (p->traceback_diagonals + 1 >= p->min_diags_between_traceback)
        return fail("traceBackDiagonals + 1 must be < minDiagsBetweenTraceBack");
    if (!(p->threshold >= 0.0 && p->threshold <= 1.0)) return fail("threshold must be in [0,1]");
    c->params = *p;
    c->have_params = true;
    for (margin_b200_ctx *sub : c->subs)
        if (margin_b200_set_params(sub, p)) return 1;
    // log-domain threshold of the tile kernels: smallest v with exp(v) >= threshold on this device
    if (p->threshold <= 0.0) {
        c->v_thr = -INFINITY;
    } else if (p->threshold >= 1.0) {
        c->v_thr = std::log(p->threshold);
    } else {
        if (check_device(c->device)) return 1;
        CU(enter_ctx(c));
        if (c->d_err.ensure(sizeof(double))) return 1;
        mb::find_log_threshold_kernel<<<1, 1, 0, c->stream>>>(p->threshold, std::log(p->threshold), c->d_err.as<double>());
        CU(cudaMemcpyAsync(&c->v_thr, c->d_err.p, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
    }
    // the same threshold for FP32 log posteriors: the smallest float that is not below it
    c->v_thr_f = (float) c->v_thr;
    if ((double) c->v_thr_f < c->v_thr) c->v_thr_f = std::nextafterf(c->v_thr_f, INFINITY);
    return 0;
}

int margin_b200_set_precision(margin_b200_ctx *c, int precision) {
    if (!c) return fail("margin_b200_set_precision: null context");
    if (precision != MARGIN_B200_F64 && precision != MARGIN_B200_F32) return fail("margin_b200_set_precision: unknown precision %d", precision);
    c->f32 = precision == MARGIN_B200_F32;
    if (!c->warmup_env) c->warmup = c->f32 ? 128 : 192;
    int sms = 0, per_sm = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device) == cudaSuccess) {
        if (c->f32) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mb::backward_tile_kernel_f<8, kTileThreads, true, false, true>, kTileThreads, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mb::backward_tile_kernel<8, kTileThreads, true, false, true>, kTileThreads, 0);
        if (per_sm > 0) c->bwd_slots = sms * per_sm * (kTileThreads / 8);
    }
    for (margin_b200_ctx *sub : c->subs) margin_b200_set_precision(sub, precision);
    return 0;
}

int margin_b200_get_precision(margin_b200_ctx *c) { return c && c->f32 ? MARGIN_B200_F32 : MARGIN_B200_F64; }

int margin_b200_set_kernel_timing(margin_b200_ctx *c, int on) {
    if (!c) return fail("margin_b200_set_kernel_timing: null context");
    c->kernel_timing = on != 0;
    for (margin_b200_ctx *sub : c->subs) sub->kernel_timing = c->kernel_timing;
    return 0;
}

int margin_b200_set_pair_format(margin_b200_ctx *c, int format) {
    if (!c) return fail("margin_b200_set_pair_format: null context");
    if (format != MARGIN_B200_PAIRS_TRIPLES && format != MARGIN_B200_PAIRS_PACKED64 && format != MARGIN_B200_PAIRS_DELTA32)
        return fail("margin_b200_set_pair_format: unknown format %d", format);
    c->pair_format = format;
    for (margin_b200_ctx *sub : c->subs) sub->pair_format = format;
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// Batch upload: large-gap splitting (host, pure integer) + job table + H2D copies
// ------------------------------------------------------------------------------------------------------------------
namespace {

struct Block {
    int64_t x1, y1, x2, y2;
};

// getSplitPointsP / getSplitPoints (impl/pairwiseAligner.c:913-965)
bool split_p(int64_t &x1, int64_t &y1, int64_t x2, int64_t y2, int64_t x3, int64_t y3, std::vector<Block> &out,
             int64_t limit, bool skip_block) {
    const int64_t l_x2 = x3 - x2, l_y2 = y3 - y2;
    if (l_x2 * l_y2 > limit) {
        const int64_t max_len = (int64_t) std::sqrt((double) limit);
        const int64_t hx = l_x2 / 2 > max_len ? max_len : l_x2 / 2;
        const int64_t hy = l_y2 / 2 > max_len ? max_len : l_y2 / 2;
        if (!skip_block) out.push_back(Block{x1, y1, x2 + hx, y2 + hy});
        x1 = x3 - hx;
        y1 = y3 - hy;
        return true;
    }
    return false;
}

void split_points(const int32_t *anc, int64_t na, int64_t lx, int64_t ly, int64_t limit, bool ragged_l, bool ragged_r,
                  std::vector<Block> &out) {
    out.clear();
    int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    for (int64_t i = 0; i < na; i++) {
        const int64_t x3 = anc[3 * i], y3 = anc[3 * i + 1];
        split_p(x1, y1, x2, y2, x3, y3, out, limit, ragged_l && i == 0);
        x2 = x3 + 1;
        y2 = y3 + 1;
    }
    if (!split_p(x1, y1, x2, y2, lx, ly, out, limit, ragged_l && na == 0) || !ragged_r) out.push_back(Block{x1, y1, lx, ly});
}

int validate_anchors(const int32_t *anc, int64_t na, int64_t lx, int64_t ly, int64_t aln) {
    // the asserts of band_constructDynamic (impl/pairwiseAligner.c:150-158), which are compiled out in the reference
    int64_t px = -1, py = -1;
    for (int64_t i = 0; i < na; i++) {
        const int64_t x = anc[3 * i], y = anc[3 * i + 1], e = anc[3 * i + 2];
        if (x <= px || y <= py || x < 0 || y < 0 || x >= lx || y >= ly || e < 0 || (e & 1))
            return fail("PAIRWISE_ALIGNMENT_EXCEPTION: alignment %lld: anchor %lld (%lld,%lld,%lld) is not strictly increasing / in range / even",
                        (long long) aln, (long long) i, (long long) x, (long long) y, (long long) e);
        px = x;
        py = y;
    }
    return 0;
}

}  // namespace

extern "C" int margin_b200_upload_batch(margin_b200_ctx *c, const margin_b200_batch *b) {
    if (!c || !b) return fail("margin_b200_upload_batch: null argument");
    if (!c->have_params) return fail("margin_b200_upload_batch: margin_b200_set_params has not been called");
    if (b->n < 0) return fail("margin_b200_upload_batch: negative batch size");
    if (check_device(c->device)) return 1;
    CU(enter_ctx(c));
    Trace tr(c);
    const margin_b200_params &p = c->params;
    const int64_t n = b->n;
    c->n_aln = n;
    c->jobs.clear();
    c->aln_first_slot.assign(n + 1, 0);
    c->aln_first_job.assign(n + 1, 0);
    const int64_t total_anchor_triples = n ? b->off_a[n] : 0;
    c->n_anchors_total = total_anchor_triples;
    const int64_t nx = n ? b->off_x[n] : 0, ny = n ? b->off_y[n] : 0;
    // ---- cheap per-alignment checks, then everything that is O(symbols) or O(anchors) happens on the device
    bool used[2] = {false, false};
    bool fits_packed = true;
    for (int64_t i = 0; i < n; i++) {
        const int64_t lx = b->off_x[i + 1] - b->off_x[i], ly = b->off_y[i + 1] - b->off_y[i];
        const int64_t na = b->off_a[i + 1] - b->off_a[i];
        if (lx < 0 || ly < 0 || na < 0) return fail("alignment %lld: negative length", (long long) i);
        if (lx + ly > (int64_t) 1 << 30) return fail("alignment %lld: sequences too long", (long long) i);
        if (lx >= (1 << 20) - 1 || ly >= (1 << 20) - 1) fits_packed = false;
        const uint8_t slot = b->model_slot ? b->model_slot[i] : 0;
        if (slot > 1 || !c->have_model[slot]) return fail("alignment %lld: model slot %d has not been set", (long long) i, slot);
        used[slot] = true;
    }
    if (used[0] && used[1] && c->host_models[0].rle != c->host_models[1].rle)
        return fail("the two state machines of a batch must both use (or both not use) repeat counts");
    const int rle = c->host_models[used[0] ? 0 : 1].rle;
    c->packed_now = c->pair_format == MARGIN_B200_PAIRS_PACKED64 && fits_packed;
    c->fmt_now = c->pair_format == MARGIN_B200_PAIRS_DELTA32 ? MARGIN_B200_PAIRS_DELTA32
                                                             : (c->packed_now ? MARGIN_B200_PAIRS_PACKED64 : MARGIN_B200_PAIRS_TRIPLES);
    c->have_switch = false;
    for (int k = 0; k < 2; k++)
        if (used[k] && (c->host_models[k].t[mb::T_YX] != -INFINITY || c->host_models[k].t[mb::T_XY] != -INFINITY))
            c->have_switch = true;
    const bool packed_in = b->packed_x != nullptr && b->packed_y != nullptr;
    const bool runs_in = b->anchors_runs != nullptr && b->off_r != nullptr;
    if (!packed_in && ((nx && !b->sym_x) || (ny && !b->sym_y))) return fail("margin_b200_upload_batch: no symbols given");
    if (c->d_symx.ensure((nx + 1) * sizeof(uint16_t)) || c->d_symy.ensure((ny + 1) * sizeof(uint16_t)) ||
        c->d_anchors.ensure(((size_t) 3 * total_anchor_triples + 3) * sizeof(int32_t)) ||
        c->d_offs.ensure((size_t) 4 * (n + 1) * sizeof(int64_t)) || c->d_status.ensure(((size_t) 2 * n + 2) * sizeof(int32_t)) ||
        c->d_px.ensure((size_t) nx + 2 * mb::PACK_SLACK) || c->d_py.ensure((size_t) ny + 2 * mb::PACK_SLACK) ||
        c->d_err.ensure(sizeof(double)) ||
        c->h_small.ensure(((size_t) 2 * n + 4) * sizeof(int32_t) + (size_t) 4 * (n + 1) * sizeof(int64_t)))
        return 1;
    CU(cudaStreamSynchronize(c->stream_sym));  // idle unless an upload was never followed by a sweep
    if (total_anchor_triples) {
        if (runs_in) {
            const int64_t n_words = b->off_r[n];
            if (c->d_anchors_delta.ensure((size_t) (n_words + 1) * sizeof(uint32_t))) return 1;
            CU(cudaMemcpyAsync(c->d_anchors_delta.p, b->anchors_runs, (size_t) n_words * sizeof(uint32_t), cudaMemcpyHostToDevice,
                               c->stream));
        } else if (b->anchors_delta) {
            if (c->d_anchors_delta.ensure((size_t) total_anchor_triples * sizeof(uint32_t))) return 1;
            CU(cudaMemcpyAsync(c->d_anchors_delta.p, b->anchors_delta, (size_t) total_anchor_triples * sizeof(uint32_t),
                               cudaMemcpyHostToDevice, c->stream));
        } else {
            if (!b->anchors) return fail("margin_b200_upload_batch: the batch has anchor offsets but no anchors");
            CU(cudaMemcpyAsync(c->d_anchors.p, b->anchors, (size_t) 3 * total_anchor_triples * sizeof(int32_t),
                               cudaMemcpyHostToDevice, c->stream));
        }
    }
    // pinned staging: [status 2n+2 int32 (read-back)][off_a | off_x | off_y, n+1 int64 each]
    int32_t *h_status = c->h_small.as<int32_t>();
    if (n) {
        int64_t *d_offs = c->d_offs.as<int64_t>();
        int64_t *h_offs = reinterpret_cast<int64_t *>(c->h_small.as<char>() + ((size_t) 2 * n + 4) * sizeof(int32_t));
        memcpy(h_offs, b->off_a, (n + 1) * sizeof(int64_t));
        memcpy(h_offs + (n + 1), b->off_x, (n + 1) * sizeof(int64_t));
        memcpy(h_offs + 2 * (n + 1), b->off_y, (n + 1) * sizeof(int64_t));
        if (runs_in) memcpy(h_offs + 3 * (n + 1), b->off_r, (n + 1) * sizeof(int64_t));
        if (ctl_copy(c->stream, d_offs, h_offs, (size_t) (runs_in ? 4 : 3) * (n + 1) * sizeof(int64_t))) return 1;
        ctl_fill_status_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, c->stream>>>(c->d_status.as<int32_t>(), n);
        if (total_anchor_triples && runs_in) {
            if (c->d_sym_err.ensure(16)) return 1;
            CU(cudaMemsetAsync(c->d_sym_err.as<char>() + 8, 0, 8, c->stream));
            mb::expand_anchor_runs_kernel<256><<<(unsigned) n, 256, 0, c->stream>>>(c->d_anchors_delta.as<uint32_t>(), d_offs + 3 * (n + 1),
                                                                                    d_offs, c->d_anchors.as<int32_t>(),
                                                                                    c->d_sym_err.as<int>() + 2);
        } else if (total_anchor_triples && b->anchors_delta)
            mb::expand_anchors_kernel<256><<<(unsigned) n, 256, 0, c->stream>>>(c->d_anchors_delta.as<uint32_t>(), d_offs,
                                                                                c->d_anchors.as<int32_t>());
        mb::check_anchors_kernel<<<(unsigned) n, 256, 0, c->stream>>>(c->d_anchors.as<int32_t>(), d_offs, d_offs + (n + 1),
                                                                     d_offs + 2 * (n + 1), p.split_matrix_bigger_than_this,
                                                                     c->d_status.as<int32_t>());
    }
    // Symbols are queued behind the anchors on their own stream: the host only waits for the anchor checks below, the
    // band phase of the sweep runs while the symbols still travel, and the DP kernels wait for ev_sym.  Packed
    // one-byte copies for the tile kernels (slack on both sides, see pairhmm_tiles.cuh); a run length the model cannot
    // hold is reported by the sweep (d_sym_err).
    if (c->d_sym_err.ensure(16)) return 1;
    CU(cudaMemsetAsync(c->d_px.p, 0, (size_t) nx + 2 * mb::PACK_SLACK, c->stream_sym));
    CU(cudaMemsetAsync(c->d_py.p, 0, (size_t) ny + 2 * mb::PACK_SLACK, c->stream_sym));
    CU(cudaMemsetAsync(c->d_sym_err.p, 0, 8, c->stream_sym));  // [0]: symbols (this stream); [2]: anchor runs (main stream)
    if (packed_in) {
        // one byte per symbol over the link; the uint16 Symbols the general kernels read are rebuilt on the device
        if (nx) CU(cudaMemcpyAsync(c->d_px.as<uint8_t>() + mb::PACK_SLACK, b->packed_x, nx, cudaMemcpyHostToDevice, c->stream_sym));
        if (ny) CU(cudaMemcpyAsync(c->d_py.as<uint8_t>() + mb::PACK_SLACK, b->packed_y, ny, cudaMemcpyHostToDevice, c->stream_sym));
        if (nx) mb::unpack_symbols_kernel<<<(unsigned) ((nx + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_symx.as<uint16_t>(), nx, rle, c->d_sym_err.as<int>());
        if (ny) mb::unpack_symbols_kernel<<<(unsigned) ((ny + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_py.as<uint8_t>() + mb::PACK_SLACK, c->d_symy.as<uint16_t>(), ny, rle, c->d_sym_err.as<int>());
    } else {
        if (nx) CU(cudaMemcpyAsync(c->d_symx.p, b->sym_x, nx * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream_sym));
        if (ny) CU(cudaMemcpyAsync(c->d_symy.p, b->sym_y, ny * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream_sym));
        if (nx) mb::pack_symbols_kernel<<<(unsigned) ((nx + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_symx.as<uint16_t>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, nx, rle, c->d_sym_err.as<int>());
        if (ny) mb::pack_symbols_kernel<<<(unsigned) ((ny + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_symy.as<uint16_t>(), c->d_py.as<uint8_t>() + mb::PACK_SLACK, ny, rle, c->d_sym_err.as<int>());
    }
    CU(cudaEventRecord(c->ev_sym, c->stream_sym));
    if (n && ctl_copy(c->stream, h_status, c->d_status.p, (size_t) 2 * n * sizeof(int32_t))) return 1;
    h_status[2 * n] = 0;
    if (n && runs_in && total_anchor_triples && ctl_copy(c->stream, h_status + 2 * n, c->d_sym_err.as<char>() + 8, 8)) return 1;
    tr.mark("enqueue");
    if (c->after_h2d_enqueue) c->after_h2d_enqueue();
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    tr.mark("h2d+check sync");
    if (h_status[2 * n]) return fail("margin_b200_upload_batch: anchors_runs does not expand to the anchor counts of off_a");

    // ---- job table: one job per alignment, or one per block where getSplitPoints cuts (rare: host path)
    std::vector<int32_t> extra_anchors;
    std::vector<Block> blocks;
    int64_t diag_off = 0;
    int64_t slot_base = 0;
    const int64_t seg_step = p.min_diags_between_traceback - p.traceback_diagonals - 1;
    for (int64_t i = 0; i < n; i++) {
        const int64_t lx = b->off_x[i + 1] - b->off_x[i], ly = b->off_y[i + 1] - b->off_y[i];
        const int64_t na = b->off_a[i + 1] - b->off_a[i];
        const int32_t *anc = b->anchors ? b->anchors + 3 * b->off_a[i] : nullptr;
        if (h_status[2 * i] != 0x7fffffff) return (b->anchors && validate_anchors(anc, na, lx, ly, i)) ? 1 : fail("PAIRWISE_ALIGNMENT_EXCEPTION: alignment %lld: bad anchors", (long long) i);
        const bool rl = b->ragged_left && b->ragged_left[i], rr = b->ragged_right && b->ragged_right[i];
        const uint8_t slot = b->model_slot ? b->model_slot[i] : 0;
        if (h_status[2 * i + 1]) {
            if (!anc && na) return fail("alignment %lld needs large-gap splitting: the batch must also carry its anchors as triples", (long long) i);
            split_points(anc, na, lx, ly, p.split_matrix_bigger_than_this, rl, rr, blocks);
        } else {
            blocks.assign(1, Block{0, 0, lx, ly});
        }
        c->aln_first_slot[i] = (int32_t) slot_base;
        c->aln_first_job[i] = (int32_t) c->jobs.size();
        int64_t j = 0;
        for (size_t k = 0; k < blocks.size(); k++) {
            const Block &bl = blocks[k];
            const int64_t j0 = j;
            if (blocks.size() == 1) j = na;
            else
                while (j < na && (int64_t) anc[3 * j] + anc[3 * j + 1] < bl.x2 + bl.y2) j++;
            mb::Job job;
            memset(&job, 0, sizeof(job));
            job.symx_off = b->off_x[i] + bl.x1;
            job.symy_off = b->off_y[i] + bl.y1;
            job.lx = (int32_t) (bl.x2 - bl.x1);
            job.ly = (int32_t) (bl.y2 - bl.y1);
            job.n_anc = (int32_t) (j - j0);
            if (bl.x1 == 0 && bl.y1 == 0) {
                job.anc_off = b->off_a[i] + j0;
            } else {  // re-based copy (impl/pairwiseAligner.c:1010-1023)
                job.anc_off = total_anchor_triples + (int64_t) extra_anchors.size() / 3;
                for (int64_t q = j0; q < j; q++) {
                    extra_anchors.push_back(anc[3 * q] - (int32_t) bl.x1);
                    extra_anchors.push_back(anc[3 * q + 1] - (int32_t) bl.y1);
                    extra_anchors.push_back(anc[3 * q + 2]);
                }
            }
            job.x1 = (int32_t) bl.x1;
            job.y1 = (int32_t) bl.y1;
            job.aln = (int32_t) i;
            job.slot = slot;
            job.ragged_l = rl || k > 0;
            job.ragged_r = rr || k + 1 < blocks.size();
            job.diag_off = diag_off;
            const int64_t N = (int64_t) job.lx + job.ly;
            diag_off += N + 2;
            job.max_seg = (int32_t) (N / seg_step + 2);
            job.seg_base = (int32_t) slot_base;
            slot_base += job.max_seg;
            if (slot_base > (int64_t) 1 << 30) return fail("batch too large: too many traceback segments");
            c->jobs.push_back(job);
        }
    }
    c->aln_first_slot[n] = (int32_t) slot_base;
    c->aln_first_job[n] = (int32_t) c->jobs.size();
    // chunks of the band construction (pairhmm_kernels.cuh): BAND_CHUNK diagonals of one job each
    c->chunks.clear();
    c->job_first_chunk.assign(c->jobs.size() + 1, 0);
    for (size_t j = 0; j < c->jobs.size(); j++) {
        c->job_first_chunk[j] = (int32_t) c->chunks.size();
        const int entries = c->jobs[j].lx + c->jobs[j].ly + 2;
        for (int d0 = 0; d0 < entries; d0 += mb::BAND_CHUNK) c->chunks.push_back(mb::BandChunk{(int32_t) j, d0});
    }
    c->job_first_chunk[c->jobs.size()] = (int32_t) c->chunks.size();
    c->total_diag = diag_off;
    c->seg_slots = (int) slot_base;
    c->total_slots = 0;
    c->total_cells = 0;

    if (!extra_anchors.empty()) {  // the re-based copies live behind the batch's own anchors
        // grow without losing the anchors already on the device
        const size_t need = ((size_t) 3 * total_anchor_triples + extra_anchors.size() + 3) * sizeof(int32_t);
        if (need > c->d_anchors.cap) {
            DevBuf bigger;
            if (bigger.ensure(need)) return 1;
            CU(cudaMemcpyAsync(bigger.p, c->d_anchors.p, (size_t) 3 * total_anchor_triples * sizeof(int32_t),
                               cudaMemcpyDeviceToDevice, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            c->d_anchors.release();
            c->d_anchors = bigger;
        }
        CU(cudaMemcpyAsync(c->d_anchors.as<int32_t>() + 3 * total_anchor_triples, extra_anchors.data(),
                           extra_anchors.size() * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    }
    if (c->d_jobs.ensure((c->jobs.size() + 1) * sizeof(mb::Job)) ||
        c->d_chunks.ensure((c->chunks.size() + 1) * sizeof(mb::BandChunk)) ||
        c->d_job_first_chunk.ensure((c->jobs.size() + 1) * sizeof(int32_t)))
        return 1;
    {
        // job table, band chunks and chunk index: staged in pinned memory, moved by the control-copy kernel
        static_assert(sizeof(mb::Job) % 8 == 0 && sizeof(mb::BandChunk) % 4 == 0, "control copies move 4-byte words");
        const size_t b_jobs = c->jobs.size() * sizeof(mb::Job), b_chunks = c->chunks.size() * sizeof(mb::BandChunk),
                     b_first = (c->jobs.size() + 1) * sizeof(int32_t);
        const size_t o_chunks = (b_jobs + 15) / 16 * 16, o_first = o_chunks + (b_chunks + 15) / 16 * 16;
        if (c->h_stage_in.ensure(o_first + b_first)) return 1;
        char *h = c->h_stage_in.as<char>();
        if (b_jobs) memcpy(h, c->jobs.data(), b_jobs);
        if (b_chunks) memcpy(h + o_chunks, c->chunks.data(), b_chunks);
        memcpy(h + o_first, c->job_first_chunk.data(), b_first);
        if (ctl_copy(c->stream, c->d_jobs.p, h, b_jobs) || ctl_copy(c->stream, c->d_chunks.p, h + o_chunks, b_chunks) ||
            ctl_copy(c->stream, c->d_job_first_chunk.p, h + o_first, b_first))
            return 1;
    }
    tr.mark("jobs");
    if (!extra_anchors.empty()) CU(cudaStreamSynchronize(c->stream));  // pageable host memory about to go out of scope
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Kernel pipeline
// ------------------------------------------------------------------------------------------------------------------
namespace {

mb::Params dev_params(const margin_b200_params &p) {
    mb::Params q;
    q.threshold = p.threshold;
    q.min_diags = (int32_t) p.min_diags_between_traceback;
    q.tb_diags = (int32_t) p.traceback_diagonals;
    q.expansion = (int32_t) p.diagonal_expansion;
    q.dynamic = p.dynamic_anchor_expansion;
    return q;
}


inline int group_of(int maxw) { return maxw <= 8 ? 8 : (maxw <= 16 ? 16 : 32); }

// Band geometry (chunk-parallel, 4 launches) + schedule_kernel, then the first host round trip of the pipeline:
// per-job cell counts, widths, segment counts and the segments themselves come back (one pinned buffer, one sync) so
// that F / staging can be sized and the work cut into tiles.  Job::cell_base / ring_w are filled in on the device.
int run_band(margin_b200_ctx *c, bool dynamic, bool with_schedule, bool want_kr) {
    const int nj = (int) c->jobs.size();
    const int nc = (int) c->chunks.size();
    mb::Params q = dev_params(c->params);
    q.dynamic = dynamic ? 1 : 0;
    const size_t S = (size_t) c->seg_slots;
    // one device block mirrored by one pinned landing zone, so that everything the host needs comes back in one copy:
    // [totals 2 x int64][job_cells][job_maxw][job_nseg][segs][err]
    const size_t o_cells = 16, o_maxw = o_cells + (size_t) nj * 8, o_nseg = o_maxw + (size_t) nj * 4,
                 o_segs = o_nseg + (size_t) nj * 4, o_err = o_segs + S * sizeof(mb::Seg), o_gaps = o_err + 8;
    // [err int32][gap count int32][gap records]: the first kGapInline records come back with the control block, the
    // rest (rare) in a second copy; more than gap_cap of them sends the batch down the sequential path
    constexpr int kGapInline = 1024;
    const bool want_gaps = with_schedule && !c->mode_sequential;
    const int gap_cap = want_gaps ? (int) std::min<int64_t>(std::max<int64_t>(c->n_anchors_total + nj, kGapInline), 1 << 20) : 0;
    const size_t o_end = o_gaps + (want_gaps ? kGapInline * sizeof(mb::AnchorGap) : 0);
    const size_t o_dev_end = o_gaps + (size_t) gap_cap * sizeof(mb::AnchorGap);
    if (c->d_kl.ensure((c->total_diag + 1) * sizeof(int32_t)) || c->d_cell_off.ensure((c->total_diag + 1) * sizeof(uint32_t)) ||
        c->d_geo.ensure((c->total_diag + 1) * sizeof(int2)) || c->d_band_out.ensure(o_dev_end) ||
        c->d_chunk_sum.ensure((nc + 1) * sizeof(uint32_t)) || c->d_chunk_maxw.ensure((nc + 1) * sizeof(int32_t)))
        return 1;
    if (want_kr && c->d_kr.ensure((c->total_diag + 1) * sizeof(int32_t))) return 1;
    // per-chunk cell bases and widest diagonals come back too: build_tiles classifies every tile by its own widest diagonal
    const size_t o_cbase = (o_end + 15) / 16 * 16, o_cmaxw = o_cbase + ((size_t) nc * 4 + 15) / 16 * 16,
                 o_all = o_cmaxw + (size_t) nc * 4;
    if (c->h_small.ensure(with_schedule ? o_all : o_end)) return 1;
    char *h = c->h_small.as<char>();
    char *dv = c->d_band_out.as<char>();
    int64_t *d_totals = reinterpret_cast<int64_t *>(dv), *d_job_cells = reinterpret_cast<int64_t *>(dv + o_cells);
    int32_t *d_job_maxw = reinterpret_cast<int32_t *>(dv + o_maxw), *d_job_nseg = reinterpret_cast<int32_t *>(dv + o_nseg);
    mb::Seg *d_segs = reinterpret_cast<mb::Seg *>(dv + o_segs);
    int *d_err = reinterpret_cast<int *>(dv + o_err);
    mb::AnchorGap *d_gaps = reinterpret_cast<mb::AnchorGap *>(dv + o_gaps);
    CU(cudaMemsetAsync(dv, 0, o_gaps, c->stream));
    if (nj > 0) {
        {
            KTimer kt_(c, K_BAND);
            mb::band_chunk_kernel<256><<<nc, 256, 0, c->stream>>>(
                c->d_jobs.as<mb::Job>(), c->d_chunks.as<mb::BandChunk>(), nc, c->d_anchors.as<int32_t>(), q,
                c->d_geo.as<int2>(), want_kr ? c->d_kr.as<int32_t>() : nullptr, c->d_chunk_sum.as<uint32_t>(),
                c->d_chunk_maxw.as<int32_t>(), d_err, c->gap_min, gap_cap, d_err + 1, want_gaps ? d_gaps : nullptr);
            mb::band_jobs_kernel<<<(nj + 3) / 4, 128, 0, c->stream>>>(nj, c->d_job_first_chunk.as<int32_t>(),
                                                                     c->d_chunk_sum.as<uint32_t>(),
                                                                     c->d_chunk_maxw.as<int32_t>(), d_job_cells, d_job_maxw);
            mb::band_bases_kernel<512><<<1, 512, 0, c->stream>>>(c->d_jobs.as<mb::Job>(), nj, d_job_cells, d_job_maxw, d_totals);
            mb::band_apply_kernel<256><<<nc, 256, 0, c->stream>>>(c->d_jobs.as<mb::Job>(), c->d_chunks.as<mb::BandChunk>(), nc,
                                                                 c->d_chunk_sum.as<uint32_t>(), c->d_geo.as<int2>(),
                                                                 c->d_kl.as<int32_t>(), c->d_cell_off.as<uint32_t>());
        }
        c->launches += 4;
        if (with_schedule) {
            const int threads = 128, per_block = threads / 32;
            {
                KTimer kt_(c, K_SCHEDULE);
                mb::schedule_kernel<<<(nj + per_block - 1) / per_block, threads, 0, c->stream>>>(
                    c->d_jobs.as<mb::Job>(), nj, c->d_cell_off.as<uint32_t>(), q, d_segs, d_job_nseg);
            }
            c->launches++;
        }
    }
    if (ctl_copy(c->stream, h, dv, (o_end + 3) / 4 * 4)) return 1;
    if (with_schedule && nc > 0 && nj > 0) {
        if (ctl_copy(c->stream, h + o_cbase, c->d_chunk_sum.p, (size_t) nc * 4) ||
            ctl_copy(c->stream, h + o_cmaxw, c->d_chunk_maxw.p, (size_t) nc * 4))
            return 1;
    }
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    if (with_schedule) {
        c->chunk_base.assign(reinterpret_cast<uint32_t *>(h + o_cbase), reinterpret_cast<uint32_t *>(h + o_cbase) + (nj > 0 ? nc : 0));
        c->chunk_maxw.assign(reinterpret_cast<int32_t *>(h + o_cmaxw), reinterpret_cast<int32_t *>(h + o_cmaxw) + (nj > 0 ? nc : 0));
    }
    if (*reinterpret_cast<int *>(h + o_err))
        return fail("PAIRWISE_ALIGNMENT_EXCEPTION: Attempt to create diagonal with invalid coordinates "
                    "(anchors inconsistent with the sequence lengths)");
    c->gaps.clear();
    c->gaps_overflow = false;
    if (want_gaps) {
        const int n_gaps = reinterpret_cast<int *>(h + o_err)[1];
        if (n_gaps > gap_cap) {
            c->gaps_overflow = true;
        } else if (n_gaps > 0) {
            const mb::AnchorGap *hg = reinterpret_cast<const mb::AnchorGap *>(h + o_gaps);
            c->gaps.assign(hg, hg + std::min(n_gaps, kGapInline));
            if (n_gaps > kGapInline) {
                const size_t rest = (size_t) (n_gaps - kGapInline) * sizeof(mb::AnchorGap);
                if (c->h_gaps.ensure(rest)) return 1;
                if (ctl_copy(c->stream, c->h_gaps.as<char>(), dv + o_end, (rest + 3) / 4 * 4)) return 1;
                CU(cudaStreamSynchronize(c->stream));
                const mb::AnchorGap *more = c->h_gaps.as<mb::AnchorGap>();
                c->gaps.insert(c->gaps.end(), more, more + (n_gaps - kGapInline));
            }
            std::sort(c->gaps.begin(), c->gaps.end(), [](const mb::AnchorGap &a, const mb::AnchorGap &b) {
                return a.job != b.job ? a.job < b.job : a.lo < b.lo;
            });
        }
    }
    c->job_cells.assign(reinterpret_cast<int64_t *>(h + o_cells), reinterpret_cast<int64_t *>(h + o_cells) + nj);
    c->job_maxw.assign(reinterpret_cast<int32_t *>(h + o_maxw), reinterpret_cast<int32_t *>(h + o_maxw) + nj);
    c->job_nseg.assign(nj, 0);
    if (with_schedule && nj) {
        memcpy(c->job_nseg.data(), h + o_nseg, (size_t) nj * 4);
        c->segs.resize(S);
        if (S) memcpy(c->segs.data(), h + o_segs, S * sizeof(mb::Seg));
    }
    int64_t cells = 0;
    for (int j = 0; j < nj; j++) {
        c->jobs[j].cell_base = cells;
        c->jobs[j].ring_w = c->job_maxw[j];
        cells += c->job_cells[j];
        if (with_schedule && c->job_nseg[j] > c->jobs[j].max_seg) return fail("internal error: segment bound exceeded");
    }
    if (nj && cells != *reinterpret_cast<int64_t *>(h)) return fail("internal error: cell totals disagree");
    c->total_cells = cells;
    return 0;
}

// A job is cut into tiles when both sequences are non-empty; degenerate one-row matrices, MARGIN_B200_MODE=sequential,
// an overflowing anchor-gap table and jobs a failed seam guard sent back (job_force_seq) take the sequential
// per-alignment sweep.  How a tile runs (lane-group width, or the general wide path) is decided per TILE, from the widest
// diagonal of the tile's own range: one wide spot in a 70 k-symbol read costs one tile, not the read.
bool job_is_tiled(const margin_b200_ctx *c, int j) {
    if (c->mode_sequential || c->gaps_overflow) return false;
    if (j < (int) c->job_force_seq.size() && c->job_force_seq[j]) return false;
    const mb::Job &job = c->jobs[j];
    return job.lx >= 1 && job.ly >= 1;
}

// Widest diagonal among d_lo .. d_hi of job j, to the granularity of the band chunks (an upper bound).
int range_maxw(const margin_b200_ctx *c, int j, int d_lo, int d_hi) {
    const int c0 = c->job_first_chunk[j], c1 = c->job_first_chunk[j + 1];
    if (c1 <= c0 || c->chunk_maxw.empty()) return c->job_maxw[j];
    const int k0 = std::min(c1 - 1, c0 + std::max(d_lo, 0) / mb::BAND_CHUNK), k1 = std::min(c1 - 1, c0 + std::max(d_hi, 0) / mb::BAND_CHUNK);
    int mw = 0;
    for (int k = k0; k <= k1; k++) mw = std::max(mw, c->chunk_maxw[k]);
    return mw;
}

// Upper bound of the band cells on diagonals d_lo .. d_hi of job j: the cells of the band chunks they touch.
int64_t range_cells_bound(const margin_b200_ctx *c, int j, int d_lo, int d_hi) {
    const int c0 = c->job_first_chunk[j], c1 = c->job_first_chunk[j + 1];
    if (c1 <= c0 || c->chunk_base.empty()) return c->job_cells[j];
    const int k0 = std::min(c1 - 1, c0 + std::max(d_lo, 0) / mb::BAND_CHUNK), k1 = std::min(c1 - 1, c0 + std::max(d_hi, 0) / mb::BAND_CHUNK);
    const int64_t hi = k1 + 1 < c1 ? (int64_t) c->chunk_base[k1 + 1] : c->job_cells[j];
    return hi - (int64_t) c->chunk_base[k0];
}

// Cuts every traceback segment into tiles (see pairhmm_tiles.cuh) and lays the tiles out in the order of the
// reference's final lists: alignments in batch order, blocks of an alignment in order, later tracebacks first, and
// inside a traceback the lowest diagonals first (each tile's tuples are reversed by the gather).
// Number of tiles a traceback segment of `span` diagonals is cut into at tile length L: the fewest tiles no longer than
// L, all of (nearly) the same length, so that the groups of a warp finish together and the tile count falls smoothly
// as L grows (a fixed step with a merged remainder left a fifth of the slots empty on some chunk sizes).
inline int tiles_of_span(int span, int L) { return span <= L ? 1 : (span + L - 1) / L; }
inline int tile_step(int span, int L) {
    const int nt = tiles_of_span(span, L);
    return std::max(10, ((span + nt - 1) / nt + 9) / 10 * 10);
}

// Tile length for this batch.  The backward kernel is the longest launch and runs bwd_slots lane groups of 8 at a time
// (occupancy x SMs); a launch takes about waves x (L + W) steps, so L is chosen among a few candidates (the lengths
// that fill 1, 2, 3 ... waves, capped at the configured length) to minimise ceil(tiles / slots) x (L + W).
// MARGIN_B200_TILE_LEN pins L instead.
int pick_tile_len(margin_b200_ctx *c) {
    if (c->tile_len_fixed || c->bwd_slots <= 0) return c->tile_len;
    const int nj = (int) c->jobs.size();
    std::vector<int32_t> &spans = c->span_tmp;
    spans.clear();
    int64_t T = 0;
    for (int j = 0; j < nj; j++) {
        if (!job_is_tiled(c, j)) continue;
        for (int s = 0; s < c->job_nseg[j]; s++) {
            const mb::Seg &sg = c->segs[c->jobs[j].seg_base + s];
            spans.push_back(sg.tb_from - sg.tb_to);
            T += sg.tb_from - sg.tb_to;
        }
    }
    if (T <= 0) return c->tile_len;
    const int Lmax = c->tile_len, Lmin = 200, W = c->warmup;
    const int64_t slots = c->bwd_slots;
    const int64_t n_min = std::max<int64_t>(1, (T + slots * Lmax - 1) / (slots * Lmax));
    int best_L = Lmax;
    double best_cost = 1e300;
    auto count = [&](int L) {
        int64_t cnt = 0;
        for (int32_t sp : spans) cnt += tiles_of_span(sp, L);
        return cnt;
    };
    for (int64_t n = n_min; n <= n_min + 2; n++) {
        // the shortest length whose tiles fit n waves: start where n waves would be exactly full and grow until the
        // per-segment remainders fit too (a launch takes about waves x (L + W) steps)
        int L = (int) std::min<int64_t>(Lmax, std::max<int64_t>(Lmin, (T / (n * slots) + 9) / 10 * 10));
        int64_t cnt = count(L);
        if (cnt > n * slots && L < Lmax) {  // bisect [L, min(Lmax, L + 640)] in steps of 10 (the count falls as L grows)
            int lo = L / 10, hi = std::min(Lmax, L + 640) / 10;
            while (lo < hi) {
                const int mid = (lo + hi) / 2;
                if (count(mid * 10) > n * slots) lo = mid + 1; else hi = mid;
            }
            L = lo * 10;
            cnt = count(L);
        }
        const double cost = (double) ((cnt + slots - 1) / slots) * (L + W);
        if (cost < best_cost) {
            best_cost = cost;
            best_L = L;
        }
    }
    return best_L;
}

// Warm-up ends for a tile edge at diagonal `at` of a job: the sweep starts W diagonals away, not counting diagonals
// inside anchor-free stretches (g[0 .. ng), sorted, disjoint) — there the band is a corridor along one sequence, the
// state distribution carries over unchanged for its whole length, and a flat start inside it would never converge to
// the true one (a 611-base deletion of the bundled ONT region did exactly that with a fixed 128).
inline int warmup_start_below(const mb::AnchorGap *g, int ng, int at, int W) {
    int k = ng - 1;
    while (k >= 0 && g[k].lo >= at) k--;  // stretches entirely above the edge do not matter
    int need = W, pos = at;
    for (; k >= 0 && need > 0; k--) {
        const int clear = pos - std::min(pos, g[k].hi);  // anchored diagonals between the stretch and pos
        if (clear >= need) return pos - need;
        need -= clear;
        pos = std::min(pos, g[k].lo);
    }
    return std::max(0, pos - need);
}
inline int warmup_end_above(const mb::AnchorGap *g, int ng, int at, int W, int limit) {
    int k = 0;
    while (k < ng && g[k].hi <= at) k++;
    int need = W, pos = at;
    for (; k < ng && need > 0; k++) {
        const int clear = std::max(pos, g[k].lo) - pos;
        if (clear >= need) return std::min(limit, pos + need);
        need -= clear;
        pos = std::max(pos, g[k].hi);
    }
    return std::min(limit, pos + need);
}

// Tile edges of one traceback segment, from tb_from downwards in steps of L (multiples of 10 below tb_from, so that
// every tile's top diagonal is one of the reference's total-probability samples).  An edge never lands inside an
// anchor-free stretch: it moves down to the stretch's start, so a long stretch (whose warm-ups would each have to
// cross all of it) belongs to ONE tile, and an alignment without anchors is one tile per traceback.
void cut_segment(const mb::AnchorGap *g, int ng, int tb_to, int tb_from, int L, std::vector<int32_t> &edges) {
    edges.clear();
    edges.push_back(tb_from);
    const int step = tile_step(tb_from - tb_to, L);
    int cur = tb_from;
    while (cur - tb_to > step) {
        int b = cur - step;
        for (int k = 0; k < ng; k++) {
            if (g[k].lo < b && b < g[k].hi) {
                b = cur - (cur - g[k].lo + 9) / 10 * 10;
                break;
            }
        }
        if (b - tb_to < 10) break;  // nothing worth a tile of its own is left
        edges.push_back(b);
        cur = b;
    }
    edges.push_back(tb_to);
}

void build_tiles(margin_b200_ctx *c, int L) {
    const int nj = (int) c->jobs.size();
    const int W = c->warmup;
    const int gmax = c->force_group ? c->force_group : 32;  // widest diagonal the lane-group tile kernels take
    size_t gap_at = 0;  // c->gaps is sorted by job and jobs are visited in ascending order
    c->tile_len_used = L;
    {
        const size_t guess = (size_t) (c->total_diag / std::max(L, 1)) + 2 * c->segs.size() + 16;
        c->tiles.reserve(guess);
        c->tile_maxw.reserve(guess);
        c->tile_seg.reserve(guess);
        c->seams.reserve(2 * guess);
    }
    c->tiles.clear();
    c->tile_maxw.clear();
    c->tile_seg.clear();
    c->tile_rerun.clear();
    c->job_gap_first.assign(nj + 1, (int32_t) c->gaps.size());
    c->seams.clear();
    c->job_first_tile.assign(nj + 1, 0);
    c->wide_cells = c->seq_cells = c->priv_cells = c->stage_cells = 0;
    std::vector<int32_t> edges;
    std::vector<std::pair<int32_t, int32_t>> blocks;  // [first, last) tile of every segment of the current job
    for (int64_t i = 0; i < c->n_aln; i++) {
        c->aln_first_slot[i] = (int32_t) c->tiles.size();
        for (int j = c->aln_first_job[i]; j < c->aln_first_job[i + 1]; j++) {
            const mb::Job &job = c->jobs[j];
            const bool tiled = job_is_tiled(c, j);
            c->job_first_tile[j] = (int32_t) c->tiles.size();
            while (gap_at < c->gaps.size() && c->gaps[gap_at].job < j) gap_at++;
            size_t gap_end = gap_at;
            while (gap_end < c->gaps.size() && c->gaps[gap_end].job == j) gap_end++;
            const mb::AnchorGap *g = c->gaps.data() + gap_at;
            const int ng = (int) (gap_end - gap_at);
            c->job_gap_first[j] = (int32_t) gap_at;
            if (!tiled) c->seq_cells += c->job_cells[j];
            blocks.clear();
            for (int s = c->job_nseg[j] - 1; s >= 0; s--) {
                const mb::Seg &sg = c->segs[job.seg_base + s];
                const int32_t first = (int32_t) c->tiles.size();
                if (!tiled) {
                    mb::Tile t;
                    t.job = j;
                    t.tb_from = sg.tb_from;
                    t.tb_to = sg.tb_to;
                    t.D = sg.D;
                    t.f0 = 0;
                    t.flags = sg.at_end ? mb::TILE_AT_END : 0;
                    t.priv_off = -1;
                    t.stage_off = c->stage_cells;
                    c->stage_cells += range_cells_bound(c, j, t.tb_to + 1, t.tb_from);
                    c->tiles.push_back(t);
                    c->tile_maxw.push_back(c->job_maxw[j]);
                    c->tile_seg.push_back(job.seg_base + s);
                    continue;
                }
                cut_segment(g, ng, sg.tb_to, sg.tb_from, L, edges);
                for (int k = (int) edges.size() - 2; k >= 0; k--) {  // lowest diagonals first
                    mb::Tile t;
                    t.job = j;
                    t.tb_from = edges[k];
                    t.tb_to = edges[k + 1];
                    t.D = (k == 0) ? sg.D : warmup_end_above(g, ng, t.tb_from, W, sg.D);
                    t.f0 = warmup_start_below(g, ng, t.tb_to, W);
                    const int mw = c->job_maxw[j] <= 8 ? c->job_maxw[j] : range_maxw(c, j, t.f0, t.D + 1);  // 8 = narrowest lane group
                    const bool wide = mw > gmax;
                    t.flags = ((t.D == sg.D && sg.at_end) ? mb::TILE_AT_END : 0) | (wide ? mb::TILE_WIDE : mb::TILE_BND);
                    t.priv_off = -1;
                    t.stage_off = 0;
                    if (wide) {
                        t.priv_off = c->priv_cells;
                        c->priv_cells += range_cells_bound(c, j, t.f0, t.tb_from);
                        t.stage_off = c->stage_cells;
                        const int64_t own = range_cells_bound(c, j, t.tb_to + 1, t.tb_from);
                        c->stage_cells += own;
                        c->wide_cells += own;
                    }
                    c->tiles.push_back(t);
                    c->tile_maxw.push_back(mw);
                    c->tile_seg.push_back(job.seg_base + s);
                }
                blocks.emplace_back(first, (int32_t) c->tiles.size());
            }
        }
    }
    for (int j = nj - 1; j >= 0; j--)  // jobs without gaps point at their successor's first entry
        if (c->job_gap_first[j] > c->job_gap_first[j + 1]) c->job_gap_first[j] = c->job_gap_first[j + 1];
    c->tile_binw = c->tile_maxw;
    if (getenv("MARGIN_B200_TILE_STATS")) {  // widest diagonal per tile, weighted by the diagonals the tile walks
        int64_t hist[34] = {0}, steps = 0;
        for (size_t t = 0; t < c->tiles.size(); t++) {
            const int64_t len = 2 * (int64_t) (c->tiles[t].D - c->tiles[t].f0);  // forward + backward
            hist[std::min(c->tile_maxw[t], 33)] += len;
            steps += len;
        }
        int longest = 0;
        for (size_t t = 0; t < c->tiles.size(); t++) longest = std::max(longest, c->tiles[t].D - c->tiles[t].f0);
        fprintf(stderr, "margin_b200 tile stats: %zu tiles for %d slots, tile length %d, longest sweep %d, %lld diagonal steps; "
                "share of steps by widest diagonal:", c->tiles.size(), c->bwd_slots, L, longest, (long long) steps);
        for (int w = 0; w < 34; w++)
            if (hist[w]) fprintf(stderr, " %d:%.3f", w, (double) hist[w] / (double) std::max<int64_t>(steps, 1));
        fprintf(stderr, "\n");
    }
    c->aln_first_slot[c->n_aln] = (int32_t) c->tiles.size();
    c->job_first_tile[nj] = (int32_t) c->tiles.size();
    c->total_slots = (int) c->tiles.size();
}

// Seams of the batch (guard_check_kernel): for every job, walking its tiles from the lowest diagonals up, the forward
// seam with the tile below and, inside a traceback, the backward seam.  Built after the forward launches (finish_bins), so
// that this piece of host work runs beside them.
void build_seams(margin_b200_ctx *c) {
    c->seams.clear();
    if (!c->guard_on) return;
    const int nj = (int) c->jobs.size();
    std::vector<std::pair<int32_t, int32_t>> blocks;  // [first, last) tile of every segment of the job, later tracebacks first
    for (int j = 0; j < nj; j++) {
        const int32_t t0 = c->job_first_tile[j], t1 = c->job_first_tile[j + 1];
        if (t1 <= t0 || !(c->tiles[t0].flags & (mb::TILE_BND | mb::TILE_WIDE))) continue;  // swept sequentially: exact
        blocks.clear();
        for (int32_t u = t0; u < t1; u++) {
            if (u == t0 || c->tile_seg[u] != c->tile_seg[u - 1]) blocks.emplace_back(u, u + 1);
            else blocks.back().second = u + 1;
        }
        int32_t prev = -1;
        for (int bi = (int) blocks.size() - 1; bi >= 0; bi--) {
            bool prev_same_segment = false;
            for (int32_t u = blocks[bi].first; u < blocks[bi].second; u++) {
                if (prev >= 0) {
                    const mb::Tile &lo = c->tiles[prev], &up = c->tiles[u];
                    if (up.f0 > 0) c->seams.push_back(mb::Seam{prev, u, up.tb_to, 0});
                    // inside a traceback the lower tile's backward sweep is warmed up through the upper tile's range
                    if (prev_same_segment && lo.D > lo.tb_from + 1 && lo.D < c->segs[c->tile_seg[prev]].D)
                        c->seams.push_back(mb::Seam{prev, u, lo.tb_from + 1, 1});
                }
                prev = u;
                prev_same_segment = true;
            }
        }
    }
}

// Repair of failed seams: the tile whose warm-up had not converged (the upper tile of a forward seam, the lower tile of a
// backward seam) gets `factor` times the warm-up on that side and is marked to run again; every other tile, its slot,
// its staging and its boundary sections stay exactly as they are.  Returns false when a repair cannot be made in
// place (the longer range contains a diagonal that changes the tile's kernel class, or nothing is left to extend):
// the caller then sends the jobs concerned to the sequential sweep and runs the whole batch again.
bool repair_tiles(margin_b200_ctx *c, const int32_t *flags, int factor) {
    const int S = c->total_slots, gmax = c->force_group ? c->force_group : 32;
    c->tile_rerun.assign(S, 0);
    bool in_place = true;
    for (size_t k = 0; k < c->seams.size(); k++) {
        if (!flags[k]) continue;
        const mb::Seam &sm = c->seams[k];
        const int s = sm.backward ? sm.lower : sm.upper;
        mb::Tile &t = c->tiles[s];
        const int j = t.job;
        const mb::AnchorGap *g = c->gaps.data() + c->job_gap_first[j];
        const int ng = c->job_gap_first[j + 1] - c->job_gap_first[j];
        const mb::Seg &sg = c->segs[c->tile_seg[s]];
        const int W = c->warmup * factor;
        if (!sm.backward) {
            const int f0 = warmup_start_below(g, ng, t.tb_to, W);
            if (f0 >= t.f0) in_place = false;  // already at the job's first diagonal
            t.f0 = f0;
        } else {
            const int D = warmup_end_above(g, ng, t.tb_from, W, sg.D);
            if (D <= t.D) in_place = false;
            t.D = D;
            if (D == sg.D && sg.at_end) t.flags |= mb::TILE_AT_END;
        }
        const int mw = range_maxw(c, j, t.f0, t.D + 1);
        if ((mw > gmax) != ((t.flags & mb::TILE_WIDE) != 0)) in_place = false;
        if ((t.flags & mb::TILE_WIDE)) in_place = false;  // a wide tile's private F was sized for its old range
        c->tile_binw[s] = mw;
        c->tile_rerun[s] = 1;
        if (!in_place) c->job_force_seq[j] = 1;
    }
    return in_place;
}

// Id lists for the launches, each sorted by work (longest first) so that the groups sharing a warp finish together.
// Layout of d_ids: [seq fwd 8|16|32][tile fwd 8|16|32][generic bwd 8|16|32][tile bwd 8|16|32].
struct Bins {
    int seq_fwd[4];
    int tile_fwd[4];
    int gen_bwd[4];
    int tile_bwd[4] = {0, 0, 0, 0};  // filled by finish_bins()
};

// Ascending sort of keys whose low 32 bits (the ids) are already ascending: stable LSD counting passes of 11 bits
// over the high half, skipping every digit on which all keys agree (tile lengths span ~11 bits and most batches use
// one lane-group width, so this is one or two passes over 2048 buckets).  O(n); a comparison sort of ~10^4 tiles was
// the largest host cost of a call, and 16-bit digits spent more time on their 65536 buckets than on the keys.
void sort_keys(std::vector<uint64_t> &k, std::vector<uint64_t> &tmp, std::vector<uint32_t> &cnt) {
    const size_t n = k.size();
    if (n < 2) return;
    uint64_t all_or = 0, all_and = ~0ull;
    for (size_t i = 0; i < n; i++) {
        all_or |= k[i];
        all_and &= k[i];
    }
    const uint64_t differ = (all_or ^ all_and) >> 32;
    tmp.resize(n);
    constexpr int BITS = 11;
    for (int shift = 32; shift < 64; shift += BITS) {
        const uint64_t mask = (1ull << std::min(BITS, 64 - shift)) - 1;
        if (((differ >> (shift - 32)) & mask) == 0) continue;
        cnt.assign(mask + 2, 0);
        for (size_t i = 0; i < n; i++) cnt[((k[i] >> shift) & mask) + 1]++;
        for (uint64_t b = 0; b < mask; b++) cnt[b + 1] += cnt[b];
        for (size_t i = 0; i < n; i++) tmp[cnt[(k[i] >> shift) & mask]++] = k[i];
        k.swap(tmp);
    }
}

int build_bins(margin_b200_ctx *c, bool with_slots, Bins &bins, int force_group) {
    const int nj = (int) c->jobs.size();
    const int S = with_slots ? c->total_slots : 0;
    auto bin_of = [&](int maxw) {
        const int g = force_group ? force_group : group_of(maxw);
        return g == 8 ? 0 : (g == 16 ? 1 : 2);
    };
    // sort keys: (lane-group bin ascending, work descending, id ascending) packed into one integer
    auto key = [](int bin, int work, int id) {
        return ((uint64_t) bin << 60) | ((uint64_t) (0x0fffffff - std::min(work, 0x0fffffff)) << 32) | (uint32_t) id;
    };
    std::vector<uint64_t> &k0 = c->keys[0], &k1 = c->keys[1], &k2 = c->keys[2], &k3 = c->keys[3];
    k0.clear();
    k1.clear();
    k2.clear();
    k3.clear();
    const bool partial = with_slots && !c->tile_rerun.empty();  // repair pass: only the marked tiles run again
    // general forward sweeps: whole jobs (id = job) ...
    for (int j = 0; j < nj && !partial; j++)
        if ((!with_slots || !job_is_tiled(c, j)) && !(!with_slots && j < (int) c->tiny_mask.size() && c->tiny_mask[j]))
            k0.push_back(key(bin_of(c->job_maxw[j]), c->jobs[j].lx + c->jobs[j].ly, j));
    int64_t ring_cells = 0, bnd_cells = 0;
    const int n_seams = 2 * S;  // capacity: the seams themselves are built and sent by finish_bins()
    // arena layout (16-byte aligned sections): tiles | ring_off | bnd_off | aln_first_slot | seams | guard flags | ids
    const size_t o_tiles = 0, o_ring = o_tiles + (((size_t) S + 1) * sizeof(mb::Tile) + 15) / 16 * 16,
                 o_bnd = o_ring + ((size_t) S + 2) * 8, o_first = o_bnd + ((size_t) S + 2) * 8,
                 o_seams = o_first + (((size_t) c->n_aln + 1) * 4 + 15) / 16 * 16,
                 o_flags = o_seams + ((size_t) n_seams + 1) * sizeof(mb::Seam),
                 o_ids = o_flags + (((size_t) n_seams + 4) * 4 + 15) / 16 * 16,
                 o_end = o_ids + ((size_t) nj + 4 * (size_t) S + 4) * 4;
    static_assert(sizeof(mb::Seam) == 16 && sizeof(mb::Tile) % 8 == 0, "arena sections stay aligned");
    if (c->h_arena.ensure(o_end) || c->d_arena.ensure(o_end)) return 1;
    char *h = c->h_arena.as<char>();
    int64_t *ring_off = reinterpret_cast<int64_t *>(h + o_ring), *bnd_off = reinterpret_cast<int64_t *>(h + o_bnd);
    if (with_slots) {
        for (int s = 0; s < S; s++) {
            const mb::Tile &t = c->tiles[s];
            // every tile owns three boundary sections of its widest diagonal (mb::BND_*)
            bnd_off[s] = bnd_cells;
            bnd_cells += 3 * (int64_t) std::max(c->tile_maxw[s], 1);
            ring_off[s] = 0;
            if (!(t.flags & mb::TILE_BND)) {
                ring_off[s] = ring_cells;
                ring_cells += 3 * (int64_t) c->jobs[t.job].ring_w;
            }
            if (partial && !c->tile_rerun[s]) continue;
            const int mw = c->tile_binw[s];
            if (t.flags & mb::TILE_BND) {
                k1.push_back(key(bin_of(mw), t.tb_from - t.f0, s));
                k3.push_back(key(bin_of(mw), t.D - t.tb_to, s));
            } else {
                if (t.flags & mb::TILE_WIDE) k0.push_back(key(bin_of(mw), t.tb_from - t.f0, ~s));  // ... and wide tiles (id = ~slot)
                k2.push_back(key(bin_of(mw), t.D - t.tb_to, s));
            }
        }
        ring_off[S] = 0;
        bnd_off[S] = bnd_cells;
        if (S) memcpy(h + o_tiles, c->tiles.data(), (size_t) S * sizeof(mb::Tile));
        memcpy(h + o_first, c->aln_first_slot.data(), ((size_t) c->n_aln + 1) * 4);
    }
    int32_t *ids = reinterpret_cast<int32_t *>(h + o_ids);
    int n_ids = 0;
    auto emit = [&](std::vector<uint64_t> &k, int *off4) {
        sort_keys(k, c->key_tmp, c->key_cnt);
        size_t q = 0;
        for (int b = 0; b < 3; b++) {
            off4[b] = n_ids;
            while (q < k.size() && (int) (k[q] >> 60) == b) ids[n_ids++] = (int32_t) (uint32_t) k[q++];
        }
        off4[3] = n_ids;
    };
    emit(k0, bins.seq_fwd);
    emit(k1, bins.tile_fwd);
    emit(k2, bins.gen_bwd);
    // The id list of the tiled backward launches is sorted and sent by finish_bins(), after the forward kernels have
    // been launched, so that this piece of host work runs beside them instead of in front of them.
    c->bins_ids_h = ids;
    c->bins_ids_off = o_ids;
    c->bins_n_ids = n_ids;
    c->arena_o_seams = o_seams;
    c->arena_o_flags = o_flags;
    char *dv = c->d_arena.as<char>();
    if (ctl_copy(c->stream, dv, h, o_seams) || ctl_copy(c->stream, dv + o_ids, h + o_ids, (size_t) n_ids * 4)) return 1;
    c->p_tiles = reinterpret_cast<mb::Tile *>(dv + o_tiles);
    c->p_ring_off = reinterpret_cast<int64_t *>(dv + o_ring);
    c->p_bnd_off = reinterpret_cast<int64_t *>(dv + o_bnd);
    c->p_aln_first_slot = reinterpret_cast<int32_t *>(dv + o_first);
    c->p_seams = reinterpret_cast<mb::Seam *>(dv + o_seams);
    c->p_guard_flags = reinterpret_cast<int32_t *>(dv + o_flags);
    c->p_ids = reinterpret_cast<int32_t *>(dv + o_ids);
    if (with_slots) {
        if (c->d_ring.ensure((size_t) (ring_cells + 1) * 3 * sizeof(double)) ||
            c->d_bnd.ensure((size_t) (bnd_cells + 1) * 3 * sizeof(double)) ||
            (c->f32 && c->d_bnd32.ensure((size_t) (bnd_cells + 1) * 4 * sizeof(float))))
            return 1;
    }
    return 0;
}

int finish_bins(margin_b200_ctx *c, Bins &bins) {
    // seams of this pass (a repair pass keeps the ones it is repairing) and their zeroed verdicts
    if (c->tile_rerun.empty()) build_seams(c);
    {
        char *h = c->h_arena.as<char>();
        const size_t n_seams = c->seams.size();
        if (n_seams) memcpy(h + c->arena_o_seams, c->seams.data(), n_seams * sizeof(mb::Seam));
        memset(h + c->arena_o_flags, 0, (n_seams + 4) * 4);
        if (n_seams && ctl_copy(c->stream, c->d_arena.as<char>() + c->arena_o_seams, h + c->arena_o_seams, n_seams * sizeof(mb::Seam)))
            return 1;
        if (ctl_copy(c->stream, c->d_arena.as<char>() + c->arena_o_flags, h + c->arena_o_flags, (n_seams + 4) * 4)) return 1;
    }
    std::vector<uint64_t> &k3 = c->keys[3];
    sort_keys(k3, c->key_tmp, c->key_cnt);
    int32_t *ids = c->bins_ids_h;
    const int first = c->bins_n_ids;
    int n_ids = first;
    size_t q = 0;
    for (int b = 0; b < 3; b++) {
        bins.tile_bwd[b] = n_ids;
        while (q < k3.size() && (int) (k3[q] >> 60) == b) ids[n_ids++] = (int32_t) (uint32_t) k3[q++];
    }
    bins.tile_bwd[3] = n_ids;
    if (n_ids > first)
        if (ctl_copy(c->stream, c->d_arena.as<char>() + c->bins_ids_off + (size_t) first * 4, ids + first,
                     (size_t) (n_ids - first) * 4))
            return 1;
    c->bins_n_ids = n_ids;
    return 0;
}

template <int G>
void launch_forward(margin_b200_ctx *c, int off, int count) {
    if (count <= 0) return;
    const int per_block = kFwdThreads / G;
    KTimer kt_(c, K_FORWARD);
    mb::forward_kernel<G, kFwdThreads><<<(count + per_block - 1) / per_block, kFwdThreads, 0, c->stream>>>(
        c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count, c->d_models.as<mb::DevModel>(),
        c->d_symx.as<uint16_t>(), c->d_symy.as<uint16_t>(), c->d_kl.as<int32_t>(), c->d_cell_off.as<uint32_t>(),
        c->d_F.as<double>(), c->d_Fpriv.as<double>());
    c->launches++;
}

template <int G, bool SW>
void launch_forward_tiles2(margin_b200_ctx *c, int off, int count) {
    const int per_block = kTileThreads / G;
    KTimer kt_(c, K_FORWARD);
    if (c->f32)
        mb::forward_tile_kernel_f<G, kTileThreads, SW><<<(count + per_block - 1) / per_block, kTileThreads, 0, c->stream>>>(
            c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count, c->d_models.as<mb::DevModel>(),
            c->d_emit_f_ptrs.as<const float *>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK,
            c->d_geo.as<int2>(), c->d_F32.as<float>(), c->d_bnd32.as<float>(), c->p_bnd_off, c->d_shiftF.as<float>());
    else
        mb::forward_tile_kernel<G, kTileThreads, SW><<<(count + per_block - 1) / per_block, kTileThreads, 0, c->stream>>>(
            c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count,
            c->d_models.as<mb::DevModel>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK,
            c->d_geo.as<int2>(), c->d_F.as<double>(), c->d_bnd.as<double>(), c->p_bnd_off);
    c->launches++;
}

template <int G>
void launch_forward_tiles(margin_b200_ctx *c, int off, int count) {
    if (count <= 0) return;
    if (c->have_switch) launch_forward_tiles2<G, true>(c, off, count);
    else launch_forward_tiles2<G, false>(c, off, count);
}

template <int G, bool INDELS>
void launch_backward(margin_b200_ctx *c, int off, int count, int64_t stage_stride) {
    if (count <= 0) return;
    const int per_block = kBwdThreads / G;
    KTimer kt_(c, K_BACKWARD);
    mb::backward_kernel<G, kBwdThreads, INDELS><<<(count + per_block - 1) / per_block, kBwdThreads, 0, c->stream>>>(
        c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count,
        c->d_models.as<mb::DevModel>(), c->d_symx.as<uint16_t>(), c->d_symy.as<uint16_t>(), c->d_kl.as<int32_t>(),
        c->d_cell_off.as<uint32_t>(), c->d_F.as<double>(), c->d_Fpriv.as<double>(), dev_params(c->params),
        c->d_ring.as<double>(), c->p_ring_off, c->d_stage.as<mb::Pair>(), stage_stride, c->d_seg_counts.as<int64_t>(),
        c->total_slots, c->d_bnd.as<double>(), c->p_bnd_off);
    c->launches++;
}

template <int G, bool INDELS, bool SW, bool GUARD>
void launch_backward_tiles3(margin_b200_ctx *c, int off, int count, int64_t stage_stride) {
    const int per_block = kTileThreads / G;
    KTimer kt_(c, K_BACKWARD);
    if (c->f32)
        mb::backward_tile_kernel_f<G, kTileThreads, INDELS, SW, GUARD><<<(count + per_block - 1) / per_block, kTileThreads, 0, c->stream>>>(
            c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count, c->d_models.as<mb::DevModel>(),
            c->d_emit_f_ptrs.as<const float *>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK,
            c->d_geo.as<int2>(), c->d_F32.as<float>(), c->d_bnd32.as<float>(), c->p_bnd_off, c->d_shiftF.as<float>(), c->v_thr_f,
            c->d_stage2.as<mb::StagedPair>(), stage_stride, c->d_seg_counts.as<int64_t>(), c->total_slots);
    else
        mb::backward_tile_kernel<G, kTileThreads, INDELS, SW, GUARD><<<(count + per_block - 1) / per_block, kTileThreads, 0, c->stream>>>(
            c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_ids + off, count,
            c->d_models.as<mb::DevModel>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK,
            c->d_geo.as<int2>(), c->d_F.as<double>(), c->d_bnd.as<double>(), c->p_bnd_off, c->v_thr,
            c->d_stage2.as<mb::StagedPair>(), stage_stride, c->d_seg_counts.as<int64_t>(), c->total_slots);
    c->launches++;
}

template <int G, bool INDELS>
void launch_backward_tiles(margin_b200_ctx *c, int off, int count, int64_t stage_stride) {
    if (count <= 0) return;
    const bool guard = c->guard_on && !c->seams.empty();
    if (c->have_switch) {
        if (guard) launch_backward_tiles3<G, INDELS, true, true>(c, off, count, stage_stride);
        else launch_backward_tiles3<G, INDELS, true, false>(c, off, count, stage_stride);
    } else {
        if (guard) launch_backward_tiles3<G, INDELS, false, true>(c, off, count, stage_stride);
        else launch_backward_tiles3<G, INDELS, false, false>(c, off, count, stage_stride);
    }
}

// MARGIN_B200_GUARD_STATS=1: what the seams of this batch looked like (log10 histogram of the largest difference per
// seam, forward and backward seams apart) and how wide its tiles are.  Diagnostic only.
void guard_report(margin_b200_ctx *c, int n_seams, int attempt) {
    std::vector<float> worst((size_t) n_seams);
    if (n_seams) cudaMemcpy(worst.data(), c->d_guard_worst.p, (size_t) n_seams * sizeof(float), cudaMemcpyDeviceToHost);
    int hist[2][20] = {};
    float mx[2] = {0, 0};
    for (int k = 0; k < n_seams; k++) {
        const int b = c->seams[k].backward ? 1 : 0;
        const float v = worst[k];
        mx[b] = std::max(mx[b], v);
        int bin = v <= 0 ? 0 : (int) std::floor(std::log10((double) v)) + 18;  // 1e-18 .. 1e1
        hist[b][std::max(0, std::min(19, bin))]++;
    }
    for (int b = 0; b < 2; b++) {
        fprintf(stderr, "[margin_b200 guard] attempt %d %s seams: max %.3g;", attempt, b ? "backward" : "forward", mx[b]);
        for (int i = 0; i < 20; i++)
            if (hist[b][i]) fprintf(stderr, " 1e%d:%d", i - 18, hist[b][i]);
        fprintf(stderr, "\n");
    }
    int wh[12] = {};
    for (int32_t w : c->tile_maxw) {
        int bin = 0;
        while (bin < 11 && (8 << bin) < w) bin++;
        wh[bin]++;
    }
    fprintf(stderr, "[margin_b200 guard] tiles by widest diagonal:");
    for (int i = 0; i < 12; i++)
        if (wh[i]) fprintf(stderr, " <=%d:%d", 8 << i, wh[i]);
    fprintf(stderr, "; L=%d W=%d\n", c->tile_len_used, c->warmup);
}

int force_group_env() {
    const char *e = getenv("MARGIN_B200_FORCE_GROUP");  // test hook: 8, 16 or 32 exercises the wide-diagonal fallback
    if (!e) return 0;
    int g = atoi(e);
    return (g == 8 || g == 16 || g == 32) ? g : 0;
}

void fill_stats(margin_b200_ctx *c, margin_b200_result *out, int64_t segments, float ms) {
    out->n = c->n_aln;
    out->band_cells = c->total_cells;
    out->diagonals = c->total_diag - (int64_t) c->jobs.size();  // N+1 per job
    out->jobs = (int64_t) c->jobs.size();
    out->segments = segments;
    out->gpu_ms = ms;
    out->kernel_launches = c->launches;
}

}  // namespace

namespace {
// Swaps the device's shared scratch arrays into the context for the duration of a sweep (token held) and back.
struct ScratchLease {
    margin_b200_ctx *c = nullptr;
    void take(margin_b200_ctx *ctx) {
        if (c) return;
        c = ctx;
        swap();
    }
    void swap() {
        ScratchPool &p = g_scratch[c->device & (kMaxDevices - 1)];
        std::swap(c->d_F, p.F);
        std::swap(c->d_F32, p.F32);
        std::swap(c->d_shiftF, p.shiftF);
        std::swap(c->d_Fpriv, p.Fpriv);
        std::swap(c->d_stage, p.stage);
        std::swap(c->d_stage2, p.stage2);
    }
    ~ScratchLease() {
        if (c) swap();
    }
};
}  // namespace

extern "C" int margin_b200_posteriors_resident(margin_b200_ctx *c, int want_indels, margin_b200_result *out) {
    if (!c || !out) return fail("margin_b200_posteriors_resident: null argument");
    if (check_device(c->device)) return 1;
    CU(enter_ctx(c));
    memset(out, 0, sizeof(*out));
    c->last_pipelined = false;
    c->launches = 0;
    c->ev_used = 0;
    c->total_tuples = 0;
    c->list_tuples[0] = c->list_tuples[1] = c->list_tuples[2] = 0;
    c->force_group = force_group_env();
    c->guard_failures = c->guard_reruns = 0;
    const int nj = (int) c->jobs.size();
    const int n_lists = want_indels ? 3 : 1;
    Trace tr(c);
    CU(cudaEventRecord(c->ev0, c->stream));
    if (run_band(c, c->params.dynamic_anchor_expansion != 0, true, false)) return 1;
    tr.mark("band+sync");
    c->job_force_seq.assign(nj, 0);
    const int L = pick_tile_len(c);  // fixed across guard re-runs: the tiles of unaffected jobs must not move
    tr.mark("tile length");
    const int64_t cells = c->total_cells;
    const int64_t stride = cells + 1;  // tile kernels: staging addressed by band cell
    int64_t *bounds = nullptr;
    int S = 0, n_seams = 0;
    int64_t stage_stride = 1;
    bool any_generic = false, any_tiled = false;
    // Several contexts on one device take turns from the first DP launch to the final synchronisation
    // (margin_b200_posteriors_batch); the band phase above, with its host round trip, runs beside the other context's kernels.
    std::unique_lock<std::mutex> token;
    ScratchLease lease;  // destroyed before the token is released

    // forward + backward launches of one pass (all tiles, or the tiles of a repair pass), then the seam guard
    auto run_dp = [&](Bins &bins) -> int {
        CU(cudaStreamWaitEvent(c->stream, c->ev_sym, 0));  // the symbols (upload's second stream) are needed from here on
        launch_forward<8>(c, bins.seq_fwd[0], bins.seq_fwd[1] - bins.seq_fwd[0]);
        launch_forward<16>(c, bins.seq_fwd[1], bins.seq_fwd[2] - bins.seq_fwd[1]);
        launch_forward<32>(c, bins.seq_fwd[2], bins.seq_fwd[3] - bins.seq_fwd[2]);
        launch_forward_tiles<8>(c, bins.tile_fwd[0], bins.tile_fwd[1] - bins.tile_fwd[0]);
        launch_forward_tiles<16>(c, bins.tile_fwd[1], bins.tile_fwd[2] - bins.tile_fwd[1]);
        launch_forward_tiles<32>(c, bins.tile_fwd[2], bins.tile_fwd[3] - bins.tile_fwd[2]);
        if (finish_bins(c, bins)) return 1;
        n_seams = (int) c->seams.size();
        if (want_indels) {
            launch_backward<8, true>(c, bins.gen_bwd[0], bins.gen_bwd[1] - bins.gen_bwd[0], stage_stride);
            launch_backward<16, true>(c, bins.gen_bwd[1], bins.gen_bwd[2] - bins.gen_bwd[1], stage_stride);
            launch_backward<32, true>(c, bins.gen_bwd[2], bins.gen_bwd[3] - bins.gen_bwd[2], stage_stride);
            launch_backward_tiles<8, true>(c, bins.tile_bwd[0], bins.tile_bwd[1] - bins.tile_bwd[0], stride);
            launch_backward_tiles<16, true>(c, bins.tile_bwd[1], bins.tile_bwd[2] - bins.tile_bwd[1], stride);
            launch_backward_tiles<32, true>(c, bins.tile_bwd[2], bins.tile_bwd[3] - bins.tile_bwd[2], stride);
        } else {
            launch_backward<8, false>(c, bins.gen_bwd[0], bins.gen_bwd[1] - bins.gen_bwd[0], stage_stride);
            launch_backward<16, false>(c, bins.gen_bwd[1], bins.gen_bwd[2] - bins.gen_bwd[1], stage_stride);
            launch_backward<32, false>(c, bins.gen_bwd[2], bins.gen_bwd[3] - bins.gen_bwd[2], stage_stride);
            launch_backward_tiles<8, false>(c, bins.tile_bwd[0], bins.tile_bwd[1] - bins.tile_bwd[0], stride);
            launch_backward_tiles<16, false>(c, bins.tile_bwd[1], bins.tile_bwd[2] - bins.tile_bwd[1], stride);
            launch_backward_tiles<32, false>(c, bins.tile_bwd[2], bins.tile_bwd[3] - bins.tile_bwd[2], stride);
        }
        // seam guard: every tile edge is checked (pairhmm_tiles.cuh: guard_check_kernel); the verdict comes back with the
        // final synchronisation (flags[n_seams] = number of failed seams)
        if (n_seams > 0) {
            if (c->guard_stats && c->d_guard_worst.ensure((size_t) n_seams * sizeof(float))) return 1;
            KTimer kt_(c, K_GUARD);
            mb::guard_check_kernel<<<(n_seams + 3) / 4, 128, 0, c->stream>>>(
                c->d_jobs.as<mb::Job>(), c->p_tiles, c->p_seams, n_seams, c->d_geo.as<int2>(), c->d_F.as<double>(),
                c->d_F32.as<float>(), c->d_Fpriv.as<double>(), c->d_bnd.as<double>(), c->d_bnd32.as<float>(), c->p_bnd_off,
                c->f32 ? 1 : 0, c->guard_tol,
                c->guard_tol32, c->guard_floor, c->p_guard_flags, c->p_guard_flags + n_seams,
                c->guard_stats ? c->d_guard_worst.as<float>() : nullptr);
            c->launches++;
        }
        return 0;
    };
    const bool delta = c->fmt_now == MARGIN_B200_PAIRS_DELTA32;
    auto run_gather = [&]() {
        if (delta) {
            if (S > 0 && nj > 0) {
                KTimer kt_(c, K_GATHER);
                mb::delta32_gather_kernel<<<S, 128, 0, c->stream>>>(
                    c->d_jobs.as<mb::Job>(), c->p_tiles, S, c->d_geo.as<int2>(), c->d_seg_counts.as<int64_t>(),
                    c->d_seg_off.as<int64_t>(), c->d_esc_off.as<int64_t>(), c->d_stage2.as<mb::StagedPair>(), stride,
                    c->d_stage.as<mb::Pair>(), stage_stride, c->d_out.as<uint32_t>(), c->d_esc.as<mb::TupleXY>(), n_lists,
                    (int64_t) (c->d_out.cap / sizeof(uint32_t)), (int64_t) (c->d_esc.cap / sizeof(mb::TupleXY)));
                c->launches++;
            }
            return;
        }
        const int64_t cap = (int64_t) (c->d_out.cap / sizeof(mb::Pair));
        if (S > 0 && nj > 0 && any_generic) {
            KTimer kt_(c, K_GATHER);
            mb::gather_kernel<<<S, 128, 0, c->stream>>>(c->p_tiles, S, c->d_seg_counts.as<int64_t>(),
                                                        c->d_seg_off.as<int64_t>(), c->d_stage.as<mb::Pair>(), stage_stride,
                                                        c->d_out.p, c->packed_now ? 1 : 0, n_lists, cap);
            c->launches++;
        }
        if (S > 0 && nj > 0 && any_tiled) {
            KTimer kt_(c, K_GATHER);
            mb::gather_staged_kernel<<<S, 128, 0, c->stream>>>(c->d_jobs.as<mb::Job>(), c->p_tiles, S, c->d_geo.as<int2>(),
                                                               c->d_seg_counts.as<int64_t>(), c->d_seg_off.as<int64_t>(),
                                                               c->d_stage2.as<mb::StagedPair>(), stride, c->d_out.p,
                                                               c->packed_now ? 1 : 0, n_lists, cap);
            c->launches++;
        }
    };
    // scan of the tuple counts, gather into the final lists, per-alignment offsets, then the call's one synchronisation
    auto finish_pass = [&]() -> int {
        // exclusive scan of the tuple counts over [list][slot] (+ one trailing zero so that the total comes out too)
        size_t tmp_bytes = 0;
        const int n_scan = 3 * S + 1;
        CU(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, c->d_seg_counts.as<int64_t>(), c->d_seg_off.as<int64_t>(), n_scan,
                                         c->stream));
        if (c->d_scan_tmp.ensure(tmp_bytes + 16)) return 1;
        {
            KTimer kt_(c, K_SCAN);
            CU(cub::DeviceScan::ExclusiveSum(c->d_scan_tmp.p, tmp_bytes, c->d_seg_counts.as<int64_t>(),
                                             c->d_seg_off.as<int64_t>(), n_scan, c->stream));
        }
        c->launches++;
        // The list totals are only known on the device at this point.  The output buffer is sized from an estimate
        // (0.22 tuples per band cell were measured on real data) and the gather refuses to write past it; the totals come
        // back with the final synchronisation, and an overflow re-runs the gather into an exactly sized buffer.
        bounds = reinterpret_cast<int64_t *>(c->h_small.as<char>());  // pinned; run_band's landing zone is consumed
        ctl_bounds_kernel<<<1, 32, 0, c->stream>>>(c->d_seg_off.as<int64_t>(), S, bounds);
        bounds[4] = 0;
        bounds[5] = 0;
        if (c->d_sym_err.p && ctl_copy(c->stream, bounds + 4, c->d_sym_err.p, 8)) return 1;  // pack_symbols_kernel's verdict
        if (n_seams > 0 && ctl_copy(c->stream, bounds + 5, c->p_guard_flags + n_seams, 4)) return 1;
        bounds[6] = bounds[7] = 0;
        if (any_generic && S > 0) {  // exact cell counts of the wide-tile and sequential paths (result counters)
            if (c->d_class_cells.ensure(16)) return 1;
            CU(cudaMemsetAsync(c->d_class_cells.p, 0, 16, c->stream));
            mb::class_cells_kernel<<<(S + 127) / 128, 128, 0, c->stream>>>(c->d_jobs.as<mb::Job>(), c->p_tiles, S, c->d_geo.as<int2>(),
                                                                        c->d_class_cells.as<unsigned long long>());
            if (ctl_copy(c->stream, bounds + 6, c->d_class_cells.p, 16)) return 1;
            c->launches++;
        }
        const size_t want_cap = (size_t) cells / (want_indels ? 3 : 6) + 4096;
        if (c->d_out.cap / sizeof(mb::Pair) < want_cap && c->d_out.ensure(want_cap * sizeof(mb::Pair))) return 1;
        bounds[8] = bounds[9] = bounds[10] = bounds[11] = 0;
        if (delta && S > 0) {
            // escapes per [list][slot], their scan (same length as the tuple scan: its temporary storage fits), totals
            if (c->d_esc_counts.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) || c->d_esc_off.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
                c->d_esc_off_out.ensure((size_t) 3 * (c->n_aln + 1) * sizeof(int64_t)) ||
                c->d_esc.ensure(((size_t) cells / 32 + (size_t) 3 * S + 4096) * sizeof(mb::TupleXY)))
                return 1;
            CU(cudaMemsetAsync(c->d_esc_counts.p, 0, ((size_t) 3 * S + 1) * sizeof(int64_t), c->stream));
            KTimer kt_(c, K_GATHER);
            mb::delta32_count_kernel<<<S, 128, 0, c->stream>>>(c->d_jobs.as<mb::Job>(), c->p_tiles, S, c->d_geo.as<int2>(),
                                                               c->d_seg_counts.as<int64_t>(), c->d_stage2.as<mb::StagedPair>(), stride,
                                                               c->d_stage.as<mb::Pair>(), stage_stride, n_lists,
                                                               c->d_esc_counts.as<int64_t>());
            CU(cub::DeviceScan::ExclusiveSum(c->d_scan_tmp.p, tmp_bytes, c->d_esc_counts.as<int64_t>(), c->d_esc_off.as<int64_t>(),
                                             n_scan, c->stream));
            ctl_bounds_kernel<<<1, 32, 0, c->stream>>>(c->d_esc_off.as<int64_t>(), S, bounds + 8);
            c->launches += 3;
        }
        run_gather();
        if (delta && S > 0) {
            const int threads = 128;
            mb::list_offsets_kernel<<<(int) ((c->n_aln + 1 + threads - 1) / threads), threads, 0, c->stream>>>(
                c->p_aln_first_slot, (int) c->n_aln, S, c->d_esc_off.as<int64_t>(), c->d_esc_off_out.as<int64_t>());
            c->launches++;
        }
        {
            const int threads = 128;
            const int blocks = (int) ((c->n_aln + 1 + threads - 1) / threads);
            KTimer kt_(c, K_OFFSETS);
            mb::list_offsets_kernel<<<blocks, threads, 0, c->stream>>>(c->p_aln_first_slot, (int) c->n_aln, S,
                                                                       c->d_seg_off.as<int64_t>(), c->d_off_out.as<int64_t>());
            c->launches++;
        }
        CU(cudaEventRecord(c->ev1, c->stream));
        tr.mark("launches");
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
        tr.mark("final sync");
        if (*reinterpret_cast<int *>(bounds + 4))
            return fail("a symbol carries a run length >= %d; the repeat-count model holds 0..%d", MARGIN_B200_MAX_RL,
                        MARGIN_B200_MAX_RL - 1);
        return 0;
    };

    for (int attempt = 0;; attempt++) {
        // ---- a full pass: cut the tiles (jobs sent back by a failed repair are swept sequentially), run everything
        build_tiles(c, L);
        tr.mark("tiles");
        S = c->total_slots;
        any_generic = any_tiled = false;
        for (const mb::Tile &t : c->tiles) {
            if (t.flags & mb::TILE_BND) any_tiled = true;
            else any_generic = true;
        }
        stage_stride = c->stage_cells + 1;  // general kernels: private regions (Tile::stage_off)
        // Everything this context owns is sized and the work lists are built and on their way BEFORE the token is taken:
        // a context's first batches grow a few dozen buffers, and allocations made while holding the token were paid by
        // every thread queueing behind it (16 threads' first chunks: waits of up to 8 s for 0.1 s of kernels).
        {
            size_t tmp_bytes = 0;
            CU(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, c->d_seg_counts.as<int64_t>(), c->d_seg_off.as<int64_t>(), 3 * S + 1,
                                             c->stream));
            const size_t want_cap = (size_t) cells / (want_indels ? 3 : 6) + 4096;
            if (c->d_seg_counts.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
                c->d_seg_off.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
                c->d_off_out.ensure((size_t) 3 * (c->n_aln + 1) * sizeof(int64_t)) || c->d_scan_tmp.ensure(tmp_bytes + 16) ||
                c->d_class_cells.ensure(16) ||
                (c->d_out.cap / sizeof(mb::Pair) < want_cap && c->d_out.ensure(want_cap * sizeof(mb::Pair))))
                return 1;
            if (delta && S > 0 &&
                (c->d_esc_counts.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) || c->d_esc_off.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
                 c->d_esc_off_out.ensure((size_t) 3 * (c->n_aln + 1) * sizeof(int64_t)) ||
                 c->d_esc.ensure(((size_t) cells / 32 + (size_t) 3 * S + 4096) * sizeof(mb::TupleXY))))
                return 1;
        }
        Bins bins;
        if (build_bins(c, true, bins, c->force_group)) return 1;
        tr.mark("bins+h2d");
        if (c->take_token && !token.owns_lock()) {
            token = std::unique_lock<std::mutex>(g_compute_token[c->device & (kMaxDevices - 1)]);
            lease.take(c);
            tr.mark("token");
        }
        bool any_sequential = false;  // whole jobs swept by the general kernels keep their F in the FP64 array
        for (const mb::Tile &t : c->tiles) any_sequential |= !(t.flags & (mb::TILE_BND | mb::TILE_WIDE));
        if (((!c->f32 || any_sequential) && c->d_F.ensure((size_t) (cells + 1) * 3 * sizeof(double))) ||
            (c->f32 && (c->d_F32.ensure((size_t) (cells + 1) * 4 * sizeof(float)) ||
                        c->d_shiftF.ensure((size_t) (c->total_diag + 1) * sizeof(float)))) ||
            c->d_Fpriv.ensure((size_t) (c->priv_cells + 1) * 3 * sizeof(double)) ||
            (any_generic && c->d_stage.ensure((size_t) stage_stride * n_lists * sizeof(mb::Pair))) ||
            (any_tiled && c->d_stage2.ensure((size_t) (cells + 1) * n_lists * sizeof(mb::StagedPair))) ||
            c->d_seg_counts.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
            c->d_seg_off.ensure(((size_t) 3 * S + 1) * sizeof(int64_t)) ||
            c->d_off_out.ensure((size_t) 3 * (c->n_aln + 1) * sizeof(int64_t)))
            return 1;
        CU(cudaMemsetAsync(c->d_seg_counts.p, 0, ((size_t) 3 * S + 1) * sizeof(int64_t), c->stream));
        if (run_dp(bins) || finish_pass()) return 1;
        int failed = *reinterpret_cast<int32_t *>(bounds + 5);
        if (c->guard_stats) guard_report(c, n_seams, attempt);
        if (failed == 0) break;
        // ---- a tile edge where the warmed-up state had not converged.  First the tiles concerned run again with a much
        // longer warm-up, everything else staying in place (repair pass); what still fails, or cannot be repaired in
        // place, sends its job to the sequential sweep (the reference's own recurrence) and the batch runs again.
        if (attempt >= 2) return fail("internal error: the seam guard keeps failing after the affected jobs went sequential");
        c->guard_failures += failed;
        c->guard_reruns++;
        int32_t *pin = reinterpret_cast<int32_t *>(c->h_arena.as<char>());  // the arena's host copy is consumed by now
        if (c->h_arena.cap < (size_t) n_seams * 4) return fail("internal error: arena smaller than the guard flags");
        auto fetch_flags = [&]() -> int {
            if (ctl_copy(c->stream, pin, c->p_guard_flags, (size_t) n_seams * 4)) return 1;
            CU(cudaStreamSynchronize(c->stream));
            return 0;
        };
        if (fetch_flags()) return 1;
        std::vector<int32_t> flags(pin, pin + n_seams);
        if (repair_tiles(c, flags.data(), c->repair_factor)) {
            tr.mark("repair");
            Bins repair_bins;
            if (build_bins(c, true, repair_bins, c->force_group)) return 1;
            if (run_dp(repair_bins) || finish_pass()) return 1;
            failed = *reinterpret_cast<int32_t *>(bounds + 5);
            if (c->guard_stats) guard_report(c, n_seams, 100 + attempt);
            c->tile_rerun.clear();
            if (failed == 0) break;
            c->guard_failures += failed;
            if (fetch_flags()) return 1;
            flags.assign(pin, pin + n_seams);
        }
        for (int k = 0; k < n_seams; k++)
            if (flags[k]) c->job_force_seq[c->tiles[c->seams[k].lower].job] = 1;
        tr.mark("guard re-run");
    }
    c->total_tuples = bounds[3];
    for (int l = 0; l < 3; l++) c->list_tuples[l] = bounds[l + 1] - bounds[l];
    c->esc_total = delta ? bounds[11] : 0;
    for (int l = 0; l < 3; l++) c->esc_list[l] = delta ? bounds[9 + l] - bounds[8 + l] : 0;
    if ((size_t) c->total_tuples > c->d_out.cap / sizeof(mb::Pair) ||
        (delta && (size_t) c->esc_total > c->d_esc.cap / sizeof(mb::TupleXY))) {  // estimate too small: exact size, gather again
        if (delta && c->d_esc.ensure((size_t) (c->esc_total + 1) * sizeof(mb::TupleXY))) return 1;
        if (c->d_out.ensure((size_t) (c->total_tuples + 1) * sizeof(mb::Pair))) return 1;
        run_gather();
        CU(cudaEventRecord(c->ev1, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaGetLastError());
    }
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    int64_t segments = 0;
    for (int j = 0; j < nj; j++) segments += c->job_nseg[j];
    fill_stats(c, out, segments, ms);
    out->tiles = S;
    out->wide_cells = bounds[6];
    out->sequential_cells = bounds[7];
    out->guard_seams = n_seams;
    out->guard_failures = c->guard_failures;
    return 0;
}

extern "C" int margin_b200_download_result(margin_b200_ctx *c, margin_b200_result *out) {
    if (!c || !out) return fail("margin_b200_download_result: null argument");
    CU(enter_ctx(c));
    const int64_t n = c->n_aln;
    const bool delta = c->fmt_now == MARGIN_B200_PAIRS_DELTA32;
    const size_t tuple_bytes = delta ? sizeof(uint32_t) : (c->packed_now ? sizeof(uint64_t) : sizeof(mb::Pair));
    // pinned memory is the slowest thing to allocate: exactly what the format needs, not the widest record
    if (c->h_out.ensure((size_t) (c->total_tuples + 1) * tuple_bytes) || c->h_off.ensure((size_t) 3 * (n + 1) * sizeof(int64_t)))
        return 1;
    if (delta) {
        if (c->h_esc.ensure((size_t) (c->esc_total + 1) * sizeof(mb::TupleXY)) || c->h_esc_off.ensure((size_t) 3 * (n + 1) * sizeof(int64_t)))
            return 1;
        if (c->esc_total)
            CU(cudaMemcpyAsync(c->h_esc.p, c->d_esc.p, (size_t) c->esc_total * sizeof(mb::TupleXY), cudaMemcpyDeviceToHost, c->stream));
        if (ctl_copy(c->stream, c->h_esc_off.p, c->d_esc_off_out.p, (size_t) 3 * (n + 1) * sizeof(int64_t))) return 1;
    }
    if (c->total_tuples)
        CU(cudaMemcpyAsync(c->h_out.p, c->d_out.p, (size_t) c->total_tuples * tuple_bytes, cudaMemcpyDeviceToHost, c->stream));
    if (ctl_copy(c->stream, c->h_off.p, c->d_off_out.p, (size_t) 3 * (n + 1) * sizeof(int64_t))) return 1;
    CU(cudaStreamSynchronize(c->stream));
    const char *base = c->h_out.as<char>();
    int64_t acc = 0;
    for (int l = 0; l < 3; l++) {
        out->pairs[l] = reinterpret_cast<const margin_b200_pair *>(base + (size_t) acc * tuple_bytes);
        out->off[l] = c->h_off.as<int64_t>() + (size_t) l * (n + 1);
        acc += c->list_tuples[l];
    }
    int64_t esc_acc = 0;
    for (int l = 0; l < 3; l++) {
        out->escapes[l] = delta ? c->h_esc.as<margin_b200_xy>() + esc_acc : nullptr;
        out->escape_off[l] = delta ? c->h_esc_off.as<int64_t>() + (size_t) l * (n + 1) : nullptr;
        esc_acc += c->esc_list[l];
    }
    out->pair_format = c->fmt_now;
    return 0;
}

// ---------------------------------------------------------------------------------------------- pipelined batch
//
// A host-memory batch spends as long on PCIe as in the kernels (C2: 70 MB in, 105 MB out, ~1.4 + ~2.0 ms against
// ~2.5 ms of device work).  posteriors_pipelined() cuts the batch into K runs of consecutive alignments and gives
// each to a sub-context (own stream, own arenas) driven by its own host thread: the copies of one run overlap the
// kernels of its neighbours, and the kernels of neighbouring runs share the SMs.  Host->device copies are queued in
// run order (a ticket), so the first run has the link to itself and starts computing early; the tuple lists land
// directly at their final place in three pinned arrays (one per list), run k behind run k-1 as soon as the sizes
// of run k-1 are known, so the result is the same contiguous layout as the unpipelined call without a host copy.
namespace {

struct PipeShared {
    std::mutex mu;
    std::condition_variable cv;
    int h2d_ticket = 0;   // run whose host->device copies may be queued next
    int sized_upto = 0;   // runs whose list sizes are known
    bool overflow = false;
    std::vector<std::array<int64_t, 3>> base;  // [run][list] first tuple of the run in the final list
};

struct PipeRun {
    int64_t first = 0, n = 0;
    std::vector<int64_t> off_x, off_y, off_a;
    margin_b200_batch view{};
    margin_b200_result res{};
    int rc = 0;
    std::string err;
};

int ensure_subs(margin_b200_ctx *c, int K) {
    while ((int) c->subs.size() < K) {
        margin_b200_ctx *sub = nullptr;
        if (margin_b200_create(&sub, c->device)) return 1;
        sub->is_sub = true;
        sub->counted = false;
        g_live_contexts[c->device & (kMaxDevices - 1)].fetch_sub(1);
        {
            // earlier runs outrank later ones, so that a run's kernels are not held up by the runs queued behind it
            // and its lists can start for the host while the others still compute
            int least = 0, greatest = 0;
            CU(cudaDeviceGetStreamPriorityRange(&least, &greatest));
            const int prio = std::min(least, greatest + (int) c->subs.size());
            std::unique_lock<std::shared_mutex> lock(g_live_mu);
            cudaStreamDestroy(sub->stream);
            sub->stream = nullptr;
            CU(cudaStreamCreateWithPriority(&sub->stream, cudaStreamNonBlocking, prio));
        }
        c->subs.push_back(sub);
        for (int slot = 0; slot < 2; slot++)
            if (c->have_model[slot] && margin_b200_set_model(sub, slot, c->user_models[slot].get())) return 1;
        if (margin_b200_set_params(sub, &c->params)) return 1;
    }
    for (int k = 0; k < K; k++) {
        margin_b200_ctx *sub = c->subs[k];
        sub->mode_sequential = c->mode_sequential;
        sub->tile_len = c->tile_len;
        sub->tile_len_fixed = c->tile_len_fixed;
        sub->warmup = c->warmup;
        sub->kernel_timing = c->kernel_timing;
        sub->guard_on = c->guard_on;
        sub->guard_tol = c->guard_tol;
        sub->guard_tol32 = c->guard_tol32;
        sub->f32 = c->f32;
        sub->repair_factor = c->repair_factor;
        sub->gap_min = c->gap_min;
        // Tile length of a run: with bwd_slots / K every run is cut as if it owned 1/K of the SMs (same tiles as the
        // unpipelined call); a larger share means shorter tiles, i.e. a run finishes sooner (its lists reach the host
        // while later runs still compute) at the price of more warm-up diagonals.
        sub->bwd_slots = std::max(kTileThreads / 8, (int) ((int64_t) c->bwd_slots * c->pipe_fill / K));
    }
    return 0;
}

// device -> pinned copies of one run's lists to their final place; the run's per-alignment offsets to its own h_off
int pipe_download(margin_b200_ctx *c, margin_b200_ctx *sub, const std::array<int64_t, 3> &base) {
    const int64_t n = sub->n_aln;
    Trace tr(sub);
    if (sub->h_off.ensure((size_t) 3 * (n + 1) * sizeof(int64_t))) return 1;
    int64_t acc = 0;
    const size_t tb = sub->packed_now ? sizeof(uint64_t) : sizeof(mb::Pair);
    for (int l = 0; l < 3; l++) {
        if (sub->list_tuples[l])
            CU(cudaMemcpyAsync(c->h_pipe_out[l].as<char>() + (size_t) base[l] * tb, sub->d_out.as<char>() + (size_t) acc * tb,
                               (size_t) sub->list_tuples[l] * tb, cudaMemcpyDeviceToHost, sub->stream));
        acc += sub->list_tuples[l];
    }
    if (ctl_copy(sub->stream, sub->h_off.p, sub->d_off_out.p, (size_t) 3 * (n + 1) * sizeof(int64_t))) return 1;
    CU(cudaStreamSynchronize(sub->stream));
    tr.mark("pipe d2h");
    return 0;
}

int posteriors_pipelined(margin_b200_ctx *c, const margin_b200_batch *b, int want_indels, margin_b200_result *out, int K) {
    const int64_t n = b->n;
    if (ensure_subs(c, K)) return 1;
    // ---- runs of consecutive alignments with about equal host bytes
    std::vector<PipeRun> runs(K);
    {
        const int64_t anchor_bytes = b->anchors_delta ? 4 : 12;
        auto bytes_upto = [&](int64_t i) { return anchor_bytes * b->off_a[i] + 2 * (b->off_x[i] + b->off_y[i]); };
        const int64_t total = bytes_upto(n);
        int64_t i = 0;
        for (int k = 0; k < K; k++) {
            runs[k].first = i;
            if (k == K - 1) {
                i = n;
            } else {
                const int64_t want = total * (k + 1) / K;
                while (i < n - (K - 1 - k) && (i <= runs[k].first || bytes_upto(i) < want)) i++;
            }
            runs[k].n = i - runs[k].first;
        }
    }
    for (PipeRun &r : runs) {
        r.off_x.resize(r.n + 1);
        r.off_y.resize(r.n + 1);
        r.off_a.resize(r.n + 1);
        for (int64_t i = 0; i <= r.n; i++) {
            r.off_x[i] = b->off_x[r.first + i] - b->off_x[r.first];
            r.off_y[i] = b->off_y[r.first + i] - b->off_y[r.first];
            r.off_a[i] = b->off_a[r.first + i] - b->off_a[r.first];
        }
        r.view.n = r.n;
        r.view.sym_x = b->sym_x + b->off_x[r.first];
        r.view.sym_y = b->sym_y + b->off_y[r.first];
        r.view.anchors = b->anchors ? b->anchors + 3 * b->off_a[r.first] : nullptr;
        r.view.anchors_delta = b->anchors_delta ? b->anchors_delta + b->off_a[r.first] : nullptr;
        r.view.off_x = r.off_x.data();
        r.view.off_y = r.off_y.data();
        r.view.off_a = r.off_a.data();
        r.view.model_slot = b->model_slot ? b->model_slot + r.first : nullptr;
        r.view.ragged_left = b->ragged_left ? b->ragged_left + r.first : nullptr;
        r.view.ragged_right = b->ragged_right ? b->ragged_right + r.first : nullptr;
    }
    // one tuple format for the whole batch (a run must not fall back on its own)
    bool fits_packed = true;
    for (int64_t i = 0; i < n; i++)
        if (b->off_x[i + 1] - b->off_x[i] >= (1 << 20) - 1 || b->off_y[i + 1] - b->off_y[i] >= (1 << 20) - 1) fits_packed = false;
    const int fmt = (c->pair_format == MARGIN_B200_PAIRS_PACKED64 && fits_packed) ? MARGIN_B200_PAIRS_PACKED64
                                                                                  : MARGIN_B200_PAIRS_TRIPLES;
    for (int k = 0; k < K; k++) c->subs[k]->pair_format = fmt;
    PipeShared sh;
    sh.base.assign(K + 1, {0, 0, 0});
    CU(enter_ctx(c));
    CU(cudaEventRecord(c->ev0, c->stream));
    auto work = [&](int k) {
        PipeRun &r = runs[k];
        margin_b200_ctx *sub = c->subs[k];
        enter_ctx(sub);
        bool released = false;
        auto release = [&]() {
            if (released) return;
            released = true;
            std::lock_guard<std::mutex> lk(sh.mu);
            sh.h2d_ticket = k + 1;
            sh.cv.notify_all();
        };
        {
            std::unique_lock<std::mutex> lk(sh.mu);
            sh.cv.wait(lk, [&] { return sh.h2d_ticket == k; });
        }
        sub->after_h2d_enqueue = release;
        r.rc = margin_b200_upload_batch(sub, &r.view);
        sub->after_h2d_enqueue = nullptr;
        release();
        if (!r.rc) r.rc = margin_b200_posteriors_resident(sub, want_indels, &r.res);
        std::array<int64_t, 3> base;
        bool fits = true;
        {
            std::unique_lock<std::mutex> lk(sh.mu);
            sh.cv.wait(lk, [&] { return sh.sized_upto == k; });
            base = sh.base[k];
            for (int l = 0; l < 3; l++) {
                sh.base[k + 1][l] = base[l] + (r.rc ? 0 : sub->list_tuples[l]);
                if ((size_t) sh.base[k + 1][l] * sizeof(mb::Pair) > c->h_pipe_out[l].cap) fits = false;
            }
            if (!fits) sh.overflow = true;
            fits = fits && !sh.overflow;
            sh.sized_upto = k + 1;
            sh.cv.notify_all();
        }
        if (!r.rc && fits) r.rc = pipe_download(c, sub, base);
        if (r.rc) r.err = g_err;
    };
    {
        std::vector<std::thread> th;
        for (int k = 1; k < K; k++) th.emplace_back(work, k);
        work(0);
        for (std::thread &t : th) t.join();
    }
    for (PipeRun &r : runs)
        if (r.rc) return fail("%s", r.err.c_str());
    if (sh.overflow) {  // first call (or a larger batch than any before): size the pinned lists, fetch again
        for (int l = 0; l < 3; l++)
            if (c->h_pipe_out[l].ensure((size_t) (sh.base[K][l] + sh.base[K][l] / 4 + 1024) * sizeof(mb::Pair))) return 1;
        for (int k = 0; k < K; k++)
            if (pipe_download(c, c->subs[k], sh.base[k])) return 1;
    }
    // ---- offsets of the whole batch, counters
    if (c->h_pipe_off.ensure((size_t) 3 * (n + 1) * sizeof(int64_t))) return 1;
    int64_t *off = c->h_pipe_off.as<int64_t>();
    memset(out, 0, sizeof(*out));
    for (int l = 0; l < 3; l++) {
        for (int k = 0; k < K; k++) {
            const int64_t *so = c->subs[k]->h_off.as<int64_t>() + (size_t) l * (runs[k].n + 1);
            for (int64_t i = 0; i < runs[k].n; i++) off[(size_t) l * (n + 1) + runs[k].first + i] = sh.base[k][l] + so[i];
        }
        off[(size_t) l * (n + 1) + n] = sh.base[K][l];
        out->pairs[l] = c->h_pipe_out[l].as<margin_b200_pair>();
        out->off[l] = off + (size_t) l * (n + 1);
    }
    out->n = n;
    out->pair_format = fmt;
    for (const PipeRun &r : runs) {
        out->band_cells += r.res.band_cells;
        out->diagonals += r.res.diagonals;
        out->jobs += r.res.jobs;
        out->segments += r.res.segments;
        out->tiles += r.res.tiles;
        out->kernel_launches += r.res.kernel_launches;
        out->wide_cells += r.res.wide_cells;
        out->sequential_cells += r.res.sequential_cells;
        out->guard_seams += r.res.guard_seams;
        out->guard_failures += r.res.guard_failures;
    }
    CU(cudaEventRecord(c->ev1, c->stream));
    CU(cudaEventSynchronize(c->ev1));
    CU(cudaEventElapsedTime(&out->gpu_ms, c->ev0, c->ev1));  // the whole call, copies included
    c->n_aln = n;
    c->last_pipelined = true;
    c->last_k = K;
    return 0;
}

}  // namespace

// Several contexts on one device (the chunk scheduler's contexts_per_gpu): left alone, threads that start together stay
// in lock-step (all copy, then all compute, then all copy back) and nothing overlaps.  A per-device compute token makes
// them take turns on the SMs, so one context's copies run beside another's kernels.  A context that is the only one on
// its device pipelines inside the call instead (posteriors_pipelined).  The choice depends on the number of live
// contexts, not on timing, so a context always takes the same path (and its arenas are sized once).
extern "C" int margin_b200_posteriors_batch(margin_b200_ctx *c, const margin_b200_batch *b, int want_indels,
                                            margin_b200_result *out) {
    if (!c || !b || !out) return fail("margin_b200_posteriors_batch: null argument");
    c->last_pipelined = false;
    if (c->is_sub) {
        if (margin_b200_upload_batch(c, b)) return 1;
        if (margin_b200_posteriors_resident(c, want_indels, out)) return 1;
        return margin_b200_download_result(c, out);
    }
    const int dev_slot = c->device & (kMaxDevices - 1);
    if ((g_live_contexts[dev_slot].load() == 1 || c->pipeline_force) && c->pipeline_k > 1 && b->n >= 2 && c->have_params &&
        c->pair_format != MARGIN_B200_PAIRS_DELTA32 && !b->packed_x && !b->anchors_runs) {  // the in-call pipeline predates the compact formats
        const int64_t bytes = 12 * b->off_a[b->n] + 2 * (b->off_x[b->n] + b->off_y[b->n]);
        const int K = (int) std::min<int64_t>(c->pipeline_k, b->n / 2);
        if (K > 1 && bytes >= c->pipeline_min_bytes) return posteriors_pipelined(c, b, want_indels, out, K);
    }
    if (margin_b200_upload_batch(c, b)) return 1;
    c->take_token = true;
    const int rc = margin_b200_posteriors_resident(c, want_indels, out);
    c->take_token = false;
    if (rc) return 1;
    return margin_b200_download_result(c, out);
}

static_assert(sizeof(mb::Pair) == sizeof(margin_b200_pair), "tuple layouts must agree");

extern "C" int margin_b200_last_kernel_ms(margin_b200_ctx *c, float *ms, int32_t *launches, int n_class) {
    if (!c || !ms || !launches) return fail("margin_b200_last_kernel_ms: null argument");
    CU(enter_ctx(c));
    CU(cudaStreamSynchronize(c->stream));
    for (int k = 0; k < n_class; k++) {
        ms[k] = 0.f;
        launches[k] = 0;
    }
    if (c->last_pipelined) {  // sums over the runs (their kernels overlap, so the sum exceeds the call's device time)
        std::vector<float> m(n_class);
        std::vector<int32_t> ln(n_class);
        for (int r = 0; r < c->last_k; r++) {
            if (margin_b200_last_kernel_ms(c->subs[r], m.data(), ln.data(), n_class)) return 1;
            for (int k = 0; k < n_class; k++) {
                ms[k] += m[k];
                launches[k] += ln[k];
            }
        }
        return 0;
    }
    for (int i = 0; i < c->ev_used; i++) {
        float t = 0.f;
        CU(cudaEventElapsedTime(&t, c->ev_pool[2 * i], c->ev_pool[2 * i + 1]));
        const int k = c->ev_class[i];
        if (k < n_class) {
            ms[k] += t;
            launches[k]++;
        }
    }
    return 0;
}

// computeForwardProbability for every alignment of the uploaded batch (impl/pairwiseAligner.c:849-903): static band
// with diagonalExpansion (callers pass no anchors => the full matrix), forward sweep, end-state dot product.
// Large-gap splitting does not apply to this entry point, so upload must have produced one job per alignment.
extern "C" int margin_b200_forward_resident(margin_b200_ctx *c, margin_b200_result *stats) {
    if (!c) return fail("margin_b200_forward_resident: null argument");
    if (check_device(c->device)) return 1;
    CU(enter_ctx(c));
    Trace tr(c);
    c->launches = 0;
    c->ev_used = 0;
    const int nj = (int) c->jobs.size();
    if ((int64_t) nj != c->n_aln)
        return fail("margin_b200_forward: an alignment was split by splitMatrixBiggerThanThis; the forward-only "
                    "path takes whole matrices");
    CU(cudaEventRecord(c->ev0, c->stream));
    CU(cudaStreamWaitEvent(c->stream, c->ev_sym, 0));
    c->force_group = force_group_env();
    // The bubble-graph case: a matrix that is small and has no anchors is swept in registers (no band
    // tables, no F in HBM); anything else takes the general path (band kernels, sequential forward sweep, final dot).
    c->tiny_mask.assign(nj, 0);
    int n_tiny = 0;
    int64_t tiny_cells = 0;
    if (!c->mode_sequential && !c->force_group) {
        for (int j = 0; j < nj; j++) {
            const mb::Job &job = c->jobs[j];
            if (job.n_anc == 0 && std::min(job.lx, job.ly) + 1 <= 32) {
                c->tiny_mask[j] = 1;
                n_tiny++;
                tiny_cells += (int64_t) (job.lx + 1) * (job.ly + 1);
            }
        }
    }
    if (c->d_fwd_out.ensure((size_t) (c->n_aln + 1) * sizeof(double))) return 1;
    tr.mark("fwd classify");
    int64_t general_cells = 0;
    if (n_tiny < nj) {
        // band tables only for the general jobs: a chunk list without the tiny ones stands in for the batch's own
        // during run_band (a tiny job then has no chunks and no cells)
        const bool filtered = n_tiny > 0;
        if (filtered) {
            c->chunks_fwd.clear();
            c->job_first_chunk_fwd.assign(nj + 1, 0);
            for (int j = 0; j < nj; j++) {
                c->job_first_chunk_fwd[j] = (int32_t) c->chunks_fwd.size();
                if (c->tiny_mask[j]) continue;
                const int entries = c->jobs[j].lx + c->jobs[j].ly + 2;
                for (int d0 = 0; d0 < entries; d0 += mb::BAND_CHUNK) c->chunks_fwd.push_back(mb::BandChunk{(int32_t) j, d0});
            }
            c->job_first_chunk_fwd[nj] = (int32_t) c->chunks_fwd.size();
            const size_t b_chunks = c->chunks_fwd.size() * sizeof(mb::BandChunk), b_first = ((size_t) nj + 1) * sizeof(int32_t);
            const size_t o_first = (b_chunks + 15) / 16 * 16;
            // staged behind the tiny id list in this call's own pinned buffer (h_stage_in may still be feeding the
            // upload's copy kernels)
            const size_t o_stage = ((size_t) n_tiny * 4 + 31) / 16 * 16;
            if (c->h_tiny_ids.ensure(o_stage + o_first + b_first) || c->d_chunks_fwd.ensure(b_chunks + 16) ||
                c->d_job_first_chunk_fwd.ensure(b_first))
                return 1;
            char *h = c->h_tiny_ids.as<char>() + o_stage;
            memcpy(h, c->chunks_fwd.data(), b_chunks);
            memcpy(h + o_first, c->job_first_chunk_fwd.data(), b_first);
            if (ctl_copy(c->stream, c->d_chunks_fwd.p, h, b_chunks) ||
                ctl_copy(c->stream, c->d_job_first_chunk_fwd.p, h + o_first, b_first))
                return 1;
            std::swap(c->d_chunks, c->d_chunks_fwd);
            std::swap(c->d_job_first_chunk, c->d_job_first_chunk_fwd);
            c->chunks.swap(c->chunks_fwd);
        }
        const int band_rc = run_band(c, false, false, false);
        if (filtered) {
            std::swap(c->d_chunks, c->d_chunks_fwd);
            std::swap(c->d_job_first_chunk, c->d_job_first_chunk_fwd);
            c->chunks.swap(c->chunks_fwd);
        }
        if (band_rc) return 1;
        Bins bins;
        if (build_bins(c, false, bins, c->force_group)) return 1;
        if (c->d_F.ensure((size_t) (c->total_cells + 1) * 3 * sizeof(double))) return 1;
        launch_forward<8>(c, bins.seq_fwd[0], bins.seq_fwd[1] - bins.seq_fwd[0]);
        launch_forward<16>(c, bins.seq_fwd[1], bins.seq_fwd[2] - bins.seq_fwd[1]);
        launch_forward<32>(c, bins.seq_fwd[2], bins.seq_fwd[3] - bins.seq_fwd[2]);
        const int n_gen = bins.seq_fwd[3] - bins.seq_fwd[0];
        if (n_gen > 0) {
            { KTimer kt_(c, K_FINAL); mb::final_dot_kernel<<<(n_gen + 127) / 128, 128, 0, c->stream>>>(
                  c->d_jobs.as<mb::Job>(), c->p_ids + bins.seq_fwd[0], n_gen, c->d_models.as<mb::DevModel>(),
                  c->d_cell_off.as<uint32_t>(), c->d_F.as<double>(), c->d_fwd_out.as<double>()); }
            c->launches++;
        }
        for (int j = 0; j < nj; j++)
            if (!c->tiny_mask[j]) general_cells += c->job_cells[j];
    }
    tr.mark("fwd general");
    if (n_tiny > 0) {
        if (c->h_tiny_ids.ensure((size_t) n_tiny * 4 + 16) || c->d_tiny_ids.ensure((size_t) n_tiny * 4 + 16)) return 1;
        int32_t *ids = c->h_tiny_ids.as<int32_t>();
        int n16 = 0;
        for (int j = 0; j < nj; j++)
            if (c->tiny_mask[j] && std::min(c->jobs[j].lx, c->jobs[j].ly) + 1 <= 16) ids[n16++] = j;
        int n_all = n16;
        for (int j = 0; j < nj; j++)
            if (c->tiny_mask[j] && std::min(c->jobs[j].lx, c->jobs[j].ly) + 1 > 16) ids[n_all++] = j;
        if (ctl_copy(c->stream, c->d_tiny_ids.p, ids, (size_t) n_tiny * 4)) return 1;
        const int32_t *d_ids = c->d_tiny_ids.as<int32_t>();
        const int e = (int) c->params.diagonal_expansion;
        auto launch = [&](auto g_tag, int off, int count) {
            constexpr int G = decltype(g_tag)::value;
            if (count <= 0) return;
            const int per_block = kTileThreads / G;
            const unsigned grid = (unsigned) ((count + per_block - 1) / per_block);
            KTimer kt_(c, K_FORWARD);
            if (c->have_switch)
                mb::tiny_forward_kernel<G, kTileThreads, true><<<grid, kTileThreads, 0, c->stream>>>(
                    c->d_jobs.as<mb::Job>(), d_ids + off, count, c->d_models.as<mb::DevModel>(),
                    c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK, e,
                    c->d_fwd_out.as<double>());
            else
                mb::tiny_forward_kernel<G, kTileThreads, false><<<grid, kTileThreads, 0, c->stream>>>(
                    c->d_jobs.as<mb::Job>(), d_ids + off, count, c->d_models.as<mb::DevModel>(),
                    c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_py.as<uint8_t>() + mb::PACK_SLACK, e,
                    c->d_fwd_out.as<double>());
            c->launches++;
        };
        launch(std::integral_constant<int, 16>{}, 0, n16);
        launch(std::integral_constant<int, 32>{}, n16, n_tiny - n16);
    }
    CU(cudaEventRecord(c->ev1, c->stream));
    tr.mark("fwd tiny enqueue");
    if (c->h_small.ensure(64)) return 1;
    int *h_sym_err = reinterpret_cast<int *>(c->h_small.as<char>() + 32);
    *h_sym_err = 0;
    if (c->d_sym_err.p && ctl_copy(c->stream, h_sym_err, c->d_sym_err.p, 8)) return 1;
    CU(cudaStreamSynchronize(c->stream));
    tr.mark("fwd sync");
    CU(cudaGetLastError());
    if (*h_sym_err)
        return fail("a symbol carries a run length >= %d; the repeat-count model holds 0..%d", MARGIN_B200_MAX_RL,
                    MARGIN_B200_MAX_RL - 1);
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (getenv("MARGIN_B200_FWD_STATS"))
        fprintf(stderr, "margin_b200 forward: %d matrices (%d register-resident, %lld cells; general path %lld cells), "
                "%d launches, %.2f ms on the device\n", nj, n_tiny, (long long) tiny_cells, (long long) general_cells,
                c->launches, ms);
    c->total_cells = general_cells + tiny_cells;  // tiny matrices are counted whole (their band leaves out far corners at most)
    c->tiny_mask.clear();
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        fill_stats(c, stats, 0, ms);
    }
    return 0;
}

namespace {

// Forward-only batch WITHOUT anchors (the bubble-graph case): nothing per alignment happens on the host.  The
// job table and the work lists are built by forward_jobs_kernel from the offsets; matrices whose shorter side fits a lane
// group are swept in registers; the few larger ones go through the general path as a small batch of their own.
int forward_batch_fast(margin_b200_ctx *c, const margin_b200_batch *b, double *out_log_prob, margin_b200_result *stats) {
    if (!c->have_params) return fail("margin_b200_forward_batch: margin_b200_set_params has not been called");
    if (check_device(c->device)) return 1;
    CU(enter_ctx(c));
    Trace tr(c);
    const int64_t n = b->n;
    const int64_t nx = b->off_x[n], ny = b->off_y[n];
    if (n >= ((int64_t) 1 << 31) - 1) return fail("margin_b200_forward_batch: too many alignments");
    const bool both = c->have_model[0] && c->have_model[1];
    if (both && c->host_models[0].rle != c->host_models[1].rle)
        return fail("the two state machines of a batch must both use (or both not use) repeat counts");
    const int rle = c->host_models[c->have_model[0] ? 0 : 1].rle;
    c->have_switch = false;
    for (int k = 0; k < 2; k++)
        if (c->have_model[k] && (c->host_models[k].t[mb::T_YX] != -INFINITY || c->host_models[k].t[mb::T_XY] != -INFINITY))
            c->have_switch = true;
    c->launches = 0;
    c->ev_used = 0;
    if (c->d_symx.ensure((nx + 1) * sizeof(uint16_t)) || c->d_symy.ensure((ny + 1) * sizeof(uint16_t)) ||
        c->d_px.ensure((size_t) nx + 2 * mb::PACK_SLACK) || c->d_py.ensure((size_t) ny + 2 * mb::PACK_SLACK) ||
        c->d_offs.ensure((size_t) 2 * (n + 1) * sizeof(int64_t)) || c->d_fwd_flags.ensure((size_t) 3 * n + 16) ||
        c->d_jobs.ensure(((size_t) n + 1) * sizeof(mb::Job)) || c->d_tiny_ids.ensure((size_t) 3 * n * sizeof(int32_t) + 16) ||
        c->d_fwd_counts.ensure(64) || c->d_fwd_out.ensure((size_t) (n + 1) * sizeof(double)) || c->d_sym_err.ensure(16) ||
        c->h_small.ensure(128))
        return 1;
    CU(cudaStreamSynchronize(c->stream_sym));
    CU(cudaEventRecord(c->ev0, c->stream));
    // symbols on their own stream, packed to bytes there (as in margin_b200_upload_batch)
    const bool packed_in = b->packed_x != nullptr && b->packed_y != nullptr;
    CU(cudaMemsetAsync(c->d_px.p, 0, (size_t) nx + 2 * mb::PACK_SLACK, c->stream_sym));
    CU(cudaMemsetAsync(c->d_py.p, 0, (size_t) ny + 2 * mb::PACK_SLACK, c->stream_sym));
    CU(cudaMemsetAsync(c->d_sym_err.p, 0, 16, c->stream_sym));
    if (packed_in) {  // the register-resident kernel reads packed bytes only; the unpack kernel is run for its range check
        if (nx) CU(cudaMemcpyAsync(c->d_px.as<uint8_t>() + mb::PACK_SLACK, b->packed_x, nx, cudaMemcpyHostToDevice, c->stream_sym));
        if (ny) CU(cudaMemcpyAsync(c->d_py.as<uint8_t>() + mb::PACK_SLACK, b->packed_y, ny, cudaMemcpyHostToDevice, c->stream_sym));
        if (nx) mb::unpack_symbols_kernel<<<(unsigned) ((nx + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_px.as<uint8_t>() + mb::PACK_SLACK, c->d_symx.as<uint16_t>(), nx, rle, c->d_sym_err.as<int>());
        if (ny) mb::unpack_symbols_kernel<<<(unsigned) ((ny + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_py.as<uint8_t>() + mb::PACK_SLACK, c->d_symy.as<uint16_t>(), ny, rle, c->d_sym_err.as<int>());
    } else {
        if (nx) CU(cudaMemcpyAsync(c->d_symx.p, b->sym_x, nx * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream_sym));
        if (ny) CU(cudaMemcpyAsync(c->d_symy.p, b->sym_y, ny * sizeof(uint16_t), cudaMemcpyHostToDevice, c->stream_sym));
        if (nx) mb::pack_symbols_kernel<<<(unsigned) ((nx + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_symx.as<uint16_t>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK, nx, rle, c->d_sym_err.as<int>());
        if (ny) mb::pack_symbols_kernel<<<(unsigned) ((ny + 255) / 256), 256, 0, c->stream_sym>>>(
                c->d_symy.as<uint16_t>(), c->d_py.as<uint8_t>() + mb::PACK_SLACK, ny, rle, c->d_sym_err.as<int>());
    }
    CU(cudaEventRecord(c->ev_sym, c->stream_sym));
    // offsets and per-alignment flags, then the job table and the work lists on the device
    int64_t *d_offx = c->d_offs.as<int64_t>(), *d_offy = d_offx + (n + 1);
    uint8_t *d_flags = c->d_fwd_flags.as<uint8_t>();
    CU(cudaMemcpyAsync(d_offx, b->off_x, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_offy, b->off_y, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    if (b->model_slot) CU(cudaMemcpyAsync(d_flags, b->model_slot, n, cudaMemcpyHostToDevice, c->stream));
    if (b->ragged_left) CU(cudaMemcpyAsync(d_flags + n, b->ragged_left, n, cudaMemcpyHostToDevice, c->stream));
    if (b->ragged_right) CU(cudaMemcpyAsync(d_flags + 2 * n, b->ragged_right, n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemsetAsync(c->d_fwd_counts.p, 0, 64, c->stream));
    int32_t *ids16 = c->d_tiny_ids.as<int32_t>(), *ids32 = ids16 + n, *ids_gen = ids32 + n;
    int32_t *d_counts = c->d_fwd_counts.as<int32_t>();
    unsigned long long *d_cells = reinterpret_cast<unsigned long long *>(c->d_fwd_counts.as<char>() + 32);
    mb::forward_jobs_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, c->stream>>>(
        n, d_offx, d_offy, b->model_slot ? d_flags : nullptr, b->ragged_left ? d_flags + n : nullptr,
        b->ragged_right ? d_flags + 2 * n : nullptr, c->have_model[0] ? 1 : 0, c->have_model[1] ? 1 : 0,
        c->d_jobs.as<mb::Job>(), ids16, ids32, ids_gen, d_counts, d_cells);
    c->launches++;
    int32_t *h_counts = reinterpret_cast<int32_t *>(c->h_small.as<char>());
    if (ctl_copy(c->stream, h_counts, c->d_fwd_counts.p, 64)) return 1;
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    tr.mark("fwd jobs on device");
    if (h_counts[3]) return fail("margin_b200_forward_batch: an alignment uses a model slot that has not been set");
    const int n16 = h_counts[0], n32 = h_counts[1], n_gen = h_counts[2];
    const int64_t tiny_cells = (int64_t) *reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(h_counts) + 32);
    CU(cudaStreamWaitEvent(c->stream, c->ev_sym, 0));
    const int e = (int) c->params.diagonal_expansion;
    auto launch = [&](auto g_tag, const int32_t *ids, int count) {
        constexpr int G = decltype(g_tag)::value;
        if (count <= 0) return;
        const int per_block = kTileThreads / G;
        const unsigned grid = (unsigned) ((count + per_block - 1) / per_block);
        KTimer kt_(c, K_FORWARD);
        if (c->have_switch)
            mb::tiny_forward_kernel<G, kTileThreads, true><<<grid, kTileThreads, 0, c->stream>>>(
                c->d_jobs.as<mb::Job>(), ids, count, c->d_models.as<mb::DevModel>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK,
                c->d_py.as<uint8_t>() + mb::PACK_SLACK, e, c->d_fwd_out.as<double>());
        else
            mb::tiny_forward_kernel<G, kTileThreads, false><<<grid, kTileThreads, 0, c->stream>>>(
                c->d_jobs.as<mb::Job>(), ids, count, c->d_models.as<mb::DevModel>(), c->d_px.as<uint8_t>() + mb::PACK_SLACK,
                c->d_py.as<uint8_t>() + mb::PACK_SLACK, e, c->d_fwd_out.as<double>());
        c->launches++;
    };
    launch(std::integral_constant<int, 16>{}, ids16, n16);
    launch(std::integral_constant<int, 32>{}, ids32, n32);
    CU(cudaEventRecord(c->ev1, c->stream));
    int *h_sym_err = reinterpret_cast<int *>(c->h_small.as<char>() + 64);
    *h_sym_err = 0;
    if (ctl_copy(c->stream, h_sym_err, c->d_sym_err.p, 8)) return 1;
    if (n) CU(cudaMemcpyAsync(out_log_prob, c->d_fwd_out.p, (size_t) n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    // ---- the matrices that do not fit a lane group: a small batch of their own on the general path
    int64_t general_cells = 0;
    std::vector<int32_t> gen_ids;
    if (n_gen > 0) {
        gen_ids.resize(n_gen);
        CU(cudaMemcpyAsync(gen_ids.data(), ids_gen, (size_t) n_gen * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    }
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaGetLastError());
    tr.mark("fwd tiny");
    if (*h_sym_err)
        return fail("a symbol carries a run length >= %d; the repeat-count model holds 0..%d", MARGIN_B200_MAX_RL,
                    MARGIN_B200_MAX_RL - 1);
    int sub_launches = 0;
    if (n_gen > 0) {
        std::sort(gen_ids.begin(), gen_ids.end());
        if (!c->fwd_sub) {
            if (margin_b200_create(&c->fwd_sub, c->device)) return 1;
            c->fwd_sub->is_sub = true;
            c->fwd_sub->counted = false;
            g_live_contexts[c->device & (kMaxDevices - 1)].fetch_sub(1);
        }
        margin_b200_ctx *sub = c->fwd_sub;
        sub->mode_sequential = true;  // whole matrices, no tiling: the general forward path
        for (int slot = 0; slot < 2; slot++)
            if (c->have_model[slot] && margin_b200_set_model(sub, slot, c->user_models[slot].get())) return 1;
        if (margin_b200_set_params(sub, &c->params)) return 1;
        std::vector<uint16_t> sx, sy;
        std::vector<int64_t> ox(n_gen + 1, 0), oy(n_gen + 1, 0), oa(n_gen + 1, 0);
        std::vector<uint8_t> sl(n_gen, 0), rl(n_gen, 0), rr(n_gen, 0);
        for (int k = 0; k < n_gen; k++) {
            const int64_t i = gen_ids[k];
            if (packed_in) {  // back to Symbols for the general path's small batch
                auto sym = [&](uint8_t p) { return (uint16_t) (rle ? ((unsigned) (p % MARGIN_B200_MAX_RL) << 8) | (p / MARGIN_B200_MAX_RL) : p / MARGIN_B200_MAX_RL); };
                for (int64_t q = b->off_x[i]; q < b->off_x[i + 1]; q++) sx.push_back(sym(b->packed_x[q]));
                for (int64_t q = b->off_y[i]; q < b->off_y[i + 1]; q++) sy.push_back(sym(b->packed_y[q]));
            } else {
                sx.insert(sx.end(), b->sym_x + b->off_x[i], b->sym_x + b->off_x[i + 1]);
                sy.insert(sy.end(), b->sym_y + b->off_y[i], b->sym_y + b->off_y[i + 1]);
            }
            ox[k + 1] = (int64_t) sx.size();
            oy[k + 1] = (int64_t) sy.size();
            sl[k] = b->model_slot ? b->model_slot[i] : 0;
            rl[k] = b->ragged_left ? b->ragged_left[i] : 0;
            rr[k] = b->ragged_right ? b->ragged_right[i] : 0;
        }
        static const uint16_t no_sym[1] = {0};
        static const int32_t no_anc[3] = {0, 0, 0};
        margin_b200_batch gb{};
        gb.n = n_gen;
        gb.sym_x = sx.empty() ? no_sym : sx.data();
        gb.sym_y = sy.empty() ? no_sym : sy.data();
        gb.off_x = ox.data();
        gb.off_y = oy.data();
        gb.off_a = oa.data();
        gb.anchors = no_anc;
        gb.model_slot = sl.data();
        gb.ragged_left = rl.data();
        gb.ragged_right = rr.data();
        std::vector<double> go(n_gen, 0.0);
        margin_b200_result gs;
        if (margin_b200_forward_batch(sub, &gb, go.data(), &gs)) return 1;
        for (int k = 0; k < n_gen; k++) out_log_prob[gen_ids[k]] = go[k];
        general_cells = gs.band_cells;
        sub_launches = gs.kernel_launches;
        tr.mark("fwd general");
    }
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (getenv("MARGIN_B200_FWD_STATS"))
        fprintf(stderr, "margin_b200 forward: %lld matrices (%d register-resident, %lld cells; general path %d matrices, %lld "
                "cells), %d launches, %.2f ms on the device\n", (long long) n, n16 + n32, (long long) tiny_cells, n_gen,
                (long long) general_cells, c->launches + sub_launches, ms);
    c->n_aln = n;
    c->jobs.clear();
    c->total_cells = tiny_cells + general_cells;
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->n = n;
        stats->band_cells = c->total_cells;
        stats->jobs = n;
        stats->gpu_ms = ms;
        stats->kernel_launches = c->launches + sub_launches;
    }
    return 0;
}

}  // namespace

extern "C" int margin_b200_forward_batch(margin_b200_ctx *c, const margin_b200_batch *b, double *out_log_prob,
                                         margin_b200_result *stats) {
    if (!c || !b || !out_log_prob) return fail("margin_b200_forward_batch: null argument");
    // no anchors anywhere (every caller of computeForwardProbability in margin): nothing per alignment on the host
    if (b->n > 0 && b->off_a[b->n] == 0 && !c->mode_sequential && !c->is_sub && !force_group_env() && !getenv("MARGIN_B200_FWD_SLOW"))
        return forward_batch_fast(c, b, out_log_prob, stats);
    // the forward-only path never splits: lift the limit for the upload, then restore it
    const int64_t saved = c->params.split_matrix_bigger_than_this;
    c->params.split_matrix_bigger_than_this = INT64_MAX;
    const int rc = margin_b200_upload_batch(c, b);
    c->params.split_matrix_bigger_than_this = saved;
    if (rc) return 1;
    if (margin_b200_forward_resident(c, stats)) return 1;
    if (b->n)
        CU(cudaMemcpyAsync(out_log_prob, c->d_fwd_out.p, (size_t) b->n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}

// Band geometry only; the test hook for band_constructDynamic / band_construct (tests/pairwiseAlignerTest.c:64-127).
extern "C" int margin_b200_bands(margin_b200_ctx *c, const margin_b200_batch *b, int32_t *xmy_l, int32_t *xmy_r) {
    if (!c || !b || !xmy_l || !xmy_r) return fail("margin_b200_bands: null argument");
    const int64_t saved = c->params.split_matrix_bigger_than_this;
    c->params.split_matrix_bigger_than_this = INT64_MAX;
    const int rc = margin_b200_upload_batch(c, b);
    c->params.split_matrix_bigger_than_this = saved;
    if (rc) return 1;
    c->launches = 0;
    c->ev_used = 0;
    if (run_band(c, c->params.dynamic_anchor_expansion != 0, false, true)) return 1;
    // device layout has N+2 entries per job; the caller's has N+1
    std::vector<int32_t> kl(c->total_diag + 1), kr(c->total_diag + 1);
    CU(cudaMemcpyAsync(kl.data(), c->d_kl.p, c->total_diag * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(kr.data(), c->d_kr.p, c->total_diag * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    int64_t o = 0;
    for (const mb::Job &job : c->jobs) {
        const int64_t N = (int64_t) job.lx + job.ly;
        memcpy(xmy_l + o, kl.data() + job.diag_off, (N + 1) * sizeof(int32_t));
        memcpy(xmy_r + o, kr.data() + job.diag_off, (N + 1) * sizeof(int32_t));
        o += N + 1;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Forward-only likelihoods of small matrices WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The same kernels as margin_b200_forward_batch's register-resident path
// (pack_symbols_kernel, forward_jobs_kernel, tiny_forward_kernel), but every scratch array — packed symbols, job table,
// work lists, counters — lives in ONE file-scope device array of fixed size instead of the context's grow-only buffers.
// Every GPU that loads the library holds these 16 MiB until exit.  All calls (of all contexts on a device) share the
// array and it keeps the last call's contents: the counters and the slack around the packed symbols are cleared by every
// call, everything else is overwritten before it is read; calls that overlap on the GPU race — serialise them.
namespace {

constexpr size_t kNoAllocBytes = (size_t) 16 << 20;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// layout of g_noalloc_scratch
constexpr size_t kNaCounts = 0;  // int32[4] list lengths + bad-slot flag | +32: uint64 cells | +64: int32[2] symbol error
constexpr size_t kNaJobs = 256;
constexpr size_t kNaIds = kNaJobs + (size_t) MARGIN_B200_NOALLOC_MAX_ALIGNMENTS * sizeof(mb::Job);
constexpr size_t kNaPx = (kNaIds + (size_t) 3 * MARGIN_B200_NOALLOC_MAX_ALIGNMENTS * sizeof(int32_t) + 255) / 256 * 256;
constexpr size_t kNaPy = kNaPx + (size_t) MARGIN_B200_NOALLOC_MAX_SYMBOLS + 2 * mb::PACK_SLACK + 128;
constexpr size_t kNaEnd = kNaPy + (size_t) MARGIN_B200_NOALLOC_MAX_SYMBOLS + 2 * mb::PACK_SLACK;
static_assert(kNaEnd <= kNoAllocBytes, "the no-alloc scratch layout must fit its array");
static_assert(kNaPy % 128 == 0 && kNaJobs % 16 == 0, "sections stay aligned");

std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[kMaxDevices];  // device address of the array on each device (the symbol itself is a host-side stand-in)

// once per device, with that device current
int noalloc_scratch(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    unsigned char *&p = g_noalloc_ptr[device & (kMaxDevices - 1)];
    if (p == nullptr) {
        void *addr = nullptr;
        CU(cudaGetSymbolAddress(&addr, g_noalloc_scratch));
        p = static_cast<unsigned char *>(addr);
    }
    *out = p;
    return 0;
}

}  // namespace

extern "C" int margin_b200_forward_tiny_noalloc(margin_b200_ctx *c, int64_t n, const uint16_t *d_sym_x, const int64_t *d_off_x,
                                                int64_t total_x, const uint16_t *d_sym_y, const int64_t *d_off_y,
                                                int64_t total_y, const uint8_t *d_model_slot, double *d_out_log_prob) {
    if (!c || !d_off_x || !d_off_y || !d_out_log_prob) return fail("margin_b200_forward_tiny_noalloc: null argument");
    if (!c->have_params) return fail("margin_b200_forward_tiny_noalloc: margin_b200_set_params has not been called");
    if (n < 0 || total_x < 0 || total_y < 0) return fail("margin_b200_forward_tiny_noalloc: negative size");
    if (n > MARGIN_B200_NOALLOC_MAX_ALIGNMENTS)
        return fail("margin_b200_forward_tiny_noalloc: %lld alignments exceed the fixed scratch (at most %d per call)",
                    (long long) n, MARGIN_B200_NOALLOC_MAX_ALIGNMENTS);
    if (total_x > MARGIN_B200_NOALLOC_MAX_SYMBOLS || total_y > MARGIN_B200_NOALLOC_MAX_SYMBOLS)
        return fail("margin_b200_forward_tiny_noalloc: %lld / %lld symbols exceed the fixed scratch (at most %d per side and call)",
                    (long long) total_x, (long long) total_y, MARGIN_B200_NOALLOC_MAX_SYMBOLS);
    if ((total_x && !d_sym_x) || (total_y && !d_sym_y)) return fail("margin_b200_forward_tiny_noalloc: null symbol array");
    if (!c->have_model[0] && !c->have_model[1]) return fail("margin_b200_forward_tiny_noalloc: no model has been set");
    if (n == 0) return 0;
    if (check_device(c->device)) return 1;
    CU(enter_ctx(c));
    unsigned char *scr = nullptr;
    if (noalloc_scratch(c->device, &scr)) return 1;
    if (c->have_model[0] && c->have_model[1] && c->host_models[0].rle != c->host_models[1].rle)
        return fail("the two state machines of a batch must both use (or both not use) repeat counts");
    const int rle = c->host_models[c->have_model[0] ? 0 : 1].rle;
    bool have_switch = false;
    for (int k = 0; k < 2; k++)
        if (c->have_model[k] && (c->host_models[k].t[mb::T_YX] != -INFINITY || c->host_models[k].t[mb::T_XY] != -INFINITY))
            have_switch = true;
    cudaStream_t st = c->stream;
    int32_t *counts = reinterpret_cast<int32_t *>(scr + kNaCounts);
    unsigned long long *cells = reinterpret_cast<unsigned long long *>(scr + kNaCounts + 32);
    int *sym_err = reinterpret_cast<int *>(scr + kNaCounts + 64);
    mb::Job *jobs = reinterpret_cast<mb::Job *>(scr + kNaJobs);
    int32_t *ids16 = reinterpret_cast<int32_t *>(scr + kNaIds), *ids32 = ids16 + n, *ids_gen = ids32 + n;
    uint8_t *px = scr + kNaPx, *py = scr + kNaPy;
    // what must start at zero: the counters, and the packed symbol arrays with their slack (the kernels read one or two
    // bytes beyond a sequence and rely on finding a valid packed symbol there)
    CU(cudaMemsetAsync(scr + kNaCounts, 0, 256, st));
    CU(cudaMemsetAsync(px, 0, (size_t) total_x + 2 * mb::PACK_SLACK, st));
    CU(cudaMemsetAsync(py, 0, (size_t) total_y + 2 * mb::PACK_SLACK, st));
    if (total_x) mb::pack_symbols_kernel<<<(unsigned) ((total_x + 255) / 256), 256, 0, st>>>(d_sym_x, px + mb::PACK_SLACK, total_x, rle, sym_err);
    if (total_y) mb::pack_symbols_kernel<<<(unsigned) ((total_y + 255) / 256), 256, 0, st>>>(d_sym_y, py + mb::PACK_SLACK, total_y, rle, sym_err);
    mb::forward_jobs_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(n, d_off_x, d_off_y, d_model_slot, nullptr, nullptr,
                                                                       c->have_model[0] ? 1 : 0, c->have_model[1] ? 1 : 0, jobs,
                                                                       ids16, ids32, ids_gen, counts, cells);
    int32_t h[20];  // the counters come back to the stack: no pinned or pageable buffer is allocated for them
    CU(cudaMemcpyAsync(h, scr + kNaCounts, sizeof(h), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (h[3]) return fail("margin_b200_forward_tiny_noalloc: an alignment uses a model slot that has not been set (or has a negative length)");
    if (h[16]) return fail("a symbol carries a run length >= %d; the repeat-count model holds 0..%d", MARGIN_B200_MAX_RL, MARGIN_B200_MAX_RL - 1);
    if (h[2] > 0)
        return fail("margin_b200_forward_tiny_noalloc: %d matrices have a shorter side of more than 31 symbols; this entry point "
                    "only takes matrices that are swept in registers (use margin_b200_forward_batch)", h[2]);
    const int e = (int) c->params.diagonal_expansion;
    auto launch = [&](auto g_tag, const int32_t *ids, int count) {
        constexpr int G = decltype(g_tag)::value;
        if (count <= 0) return;
        const int per_block = kTileThreads / G;
        const unsigned grid = (unsigned) ((count + per_block - 1) / per_block);
        if (have_switch)
            mb::tiny_forward_kernel<G, kTileThreads, true><<<grid, kTileThreads, 0, st>>>(
                jobs, ids, count, c->d_models.as<mb::DevModel>(), px + mb::PACK_SLACK, py + mb::PACK_SLACK, e, d_out_log_prob);
        else
            mb::tiny_forward_kernel<G, kTileThreads, false><<<grid, kTileThreads, 0, st>>>(
                jobs, ids, count, c->d_models.as<mb::DevModel>(), px + mb::PACK_SLACK, py + mb::PACK_SLACK, e, d_out_log_prob);
    };
    launch(std::integral_constant<int, 16>{}, ids16, h[0]);
    launch(std::integral_constant<int, 32>{}, ids32, h[1]);
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    return 0;
}
