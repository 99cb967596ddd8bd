// pairhmm_kernels.cuh — CUDA kernels of the banded pair-HMM path (sm_100a).
//
// Pipeline for a batch (one launch each, all on the context's stream):
//   band_kernel      band geometry per anti-diagonal + cell offsets   (band_constructDynamic, pairwiseAligner.c:120-226)
//   schedule_kernel  traceback segments per job                       (the trigger at pairwiseAligner.c:746-749, :768)
//   forward_kernel   forward sweep, F stored in HBM                   (diagonalCalculationForward, :566-570)
//   backward_kernel  per segment: backward, total probability, posterior tuples staged per segment (:572-681, :753-818)
//   (cub scan)       exclusive scan of per-segment tuple counts in final list order
//   gather_kernel    staged tuples -> final lists in the reference's stList order (:1104-1123)
// and for computeForwardProbability (:849-903): band_kernel + forward_kernel + final_dot_kernel,
// or tiny_forward_kernel when the whole matrix is small (the bubble-graph case).
//
// Work decomposition: a "job" is one alignment block (an alignment, or one piece of it after
// large-gap splitting); a job is swept by a group of G lanes (G = 8, 16 or 32), lane = x mod G,
// so that the three neighbours of a cell are on the same lane (x, y-1) or on lane-1 (x-1, y) and
// (x-1, y-1): two rotate-shuffles per state and no shared memory.  Diagonals wider than G fall back
// to reading the neighbours from the F array in HBM / the B ring in scratch.
#pragma once
#include "pairhmm_device.cuh"

namespace mb {

struct Job {
    int64_t symx_off, symy_off;  // element offsets into the device symbol pools (block start included)
    int64_t anc_off;             // triple offset into the device anchor pool (already re-based to the block)
    int64_t diag_off;            // offset into kl[] / cell_off[] (lx+ly+2 entries per job)
    int64_t cell_base;           // offset of the job's first cell in F / staging (set after band_kernel)
    int64_t ring_off;            // offset (in cells) of the job's backward ring scratch, per segment slot
    int32_t lx, ly;
    int32_t n_anc;
    int32_t x1, y1;              // block origin, added back to the emitted coordinates
    int32_t aln;                 // alignment index in the batch
    int32_t seg_base, max_seg;   // slot range in the segment table
    int32_t ring_w;              // ring capacity in cells per diagonal (max band width of the job)
    uint8_t slot, ragged_l, ragged_r, pad;
};

struct Seg {
    int32_t job;  // -1 = unused slot
    int32_t D, tb_to, tb_from;
    int32_t at_end;
};

// A run of anti-diagonals (lo, hi) of a job without an anchor (band_chunk_kernel).
struct AnchorGap {
    int32_t job, lo, hi;
};

// A unit of backward (and, in tiled mode, forward) work: a traceback segment or a piece of one (pairhmm_tiles.cuh).
// Three kinds:
//   TILE_BND   narrow tile (every diagonal of [f0, D] fits a lane group): forward_tile_kernel / backward_tile_kernel,
//              F in the batch's shared F array, its own copy of F(tb_to) in the boundary buffer;
//   TILE_WIDE  tile with a diagonal wider than 32 cells: the general range kernels below (neighbours through memory)
//              with a PRIVATE copy of F for [f0, tb_from] at priv_off, so that its warm-up never touches cells other
//              tiles own;
//   neither    a whole traceback segment of a job swept sequentially (MARGIN_B200_MODE=sequential, degenerate
//              matrices, guard re-runs): F comes from the job's own forward sweep in the shared array.
struct Tile {
    int32_t job;
    int32_t D;        // diagonal where the backward sweep starts
    int32_t tb_to;    // posteriors are emitted for diagonals tb_to+1 .. tb_from
    int32_t tb_from;
    int32_t f0;       // diagonal where the tile's forward sweep starts (0 = the true start state)
    int32_t flags;    // TILE_* bits
    int64_t priv_off;   // TILE_WIDE: first cell of the tile's private F (cells of f0 .. tb_from); else -1
    int64_t stage_off;  // general backward kernel: first staged tuple of the tile in each list's staging array
};
enum { TILE_AT_END = 1, TILE_BND = 2, TILE_WIDE = 4 };  // AT_END: D is the job's last diagonal

// Each tile owns 3 * w cells in the boundary buffer (w = widest diagonal of the tile): its own F(tb_to), and for the
// seam guard B(tb_to + 1) (the last diagonal of its backward sweep) and B(tb_from + 1) (a warm-up diagonal the tile
// above owns); see guard_check_kernel.
enum { BND_F = 0, BND_B_LOW = 1, BND_B_HIGH = 2 };

// A seam between two tiles of a job: both computed the same cells of diagonal d (one of them during its warm-up).
struct Seam {
    int32_t lower, upper;  // tile slots: `lower` owns diagonals <= d (forward seam) resp. starts above d (backward seam)
    int32_t d;             // the diagonal compared: F(d) for a forward seam, B(d) for a backward seam
    int32_t backward;      // 0 = forward seam, 1 = backward seam
};

struct Params {
    double threshold;
    int32_t min_diags, tb_diags, expansion, dynamic;
};

// ------------------------------------------------------------------------------------------ band

__device__ __forceinline__ int bound_coord(int z, int l) { return z < 0 ? 0 : (z > l ? l : z); }

// band_setCurrentDiagonal (pairwiseAligner.c:96-114); C division semantics kept (operands are non-negative here
// after bounding, except the pre-bound values which are bounded before use).
__device__ __forceinline__ void clip_diagonal(int xay, int xL, int yL, int xU, int yU, int &kl, int &kr) {
    int l = xL - yL, r = xU - yU;
    if ((xay + l) & 1) l += 1;
    if ((xay + r) & 1) r += 1;
    int i = (xay + l) / 2;
    if (i < xL) l += 2 * (xL - i);
    i = (xay - l) / 2;
    if (yL < i) l += 2 * (i - yL);
    i = (xay + r) / 2;
    if (xU < i) r -= 2 * (i - xU);
    i = (xay - r) / 2;
    if (i < yU) r -= 2 * (yU - i);
    kl = l;
    kr = r;
}

// One CTA per job.  For every anti-diagonal d the governing (previous, next) anchor pair is found by binary
// search on the anchors' own anti-diagonals, which makes the walk of band_constructDynamic data-parallel.
// Writes kl[d] and the exclusive prefix sum of the widths (cell_off[d], d = 0..N+1).
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
band_kernel(const Job *__restrict__ jobs, int n_jobs, const int32_t *__restrict__ anchors, Params p,
            int32_t *__restrict__ kl_out, uint32_t *__restrict__ cell_off, int64_t *__restrict__ job_cells,
            int32_t *__restrict__ job_maxw, int32_t *__restrict__ kr_out /* may be null */, int2 *__restrict__ geo_out,
            int *__restrict__ err) {
    const int j = blockIdx.x;
    if (j >= n_jobs) return;
    const Job job = jobs[j];
    const int lx = job.lx, ly = job.ly, N = lx + ly, na = job.n_anc;
    const int32_t *anc = anchors + 3 * job.anc_off;
    __shared__ uint32_t warp_sums[THREADS / 32];
    __shared__ uint32_t carry_s;
    __shared__ int maxw_s;
    if (threadIdx.x == 0) {
        carry_s = 0;
        maxw_s = 0;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base <= N; base += THREADS) {
        const int d = base + threadIdx.x;
        int w = 0, kl = 0, kr = 0;
        if (d <= N) {
            if (d > 0) {
                // smallest anchor index i with (ax+1)+(ay+1) >= d
                int lo = 0, hi = na;
                while (lo < hi) {
                    int mid = (lo + hi) >> 1;
                    int s = anc[3 * mid] + anc[3 * mid + 1] + 2;
                    if (s >= d) hi = mid; else lo = mid + 1;
                }
                int nx = lx, ny = ly, px = 0, py = 0, e;
                if (lo < na) {
                    nx = anc[3 * lo] + 1;
                    ny = anc[3 * lo + 1] + 1;
                    e = p.dynamic ? anc[3 * lo + 2] : p.expansion;
                } else {
                    e = p.dynamic ? (na > 0 ? anc[3 * (na - 1) + 2] : 0) : p.expansion;
                }
                if (lo > 0) {
                    px = anc[3 * (lo - 1)] + 1;
                    py = anc[3 * (lo - 1) + 1] + 1;
                }
                const int pxay = px + py, pxmy = px - py, nxay = nx + ny, nxmy = nx - ny;
                const int xL = bound_coord((pxay + pxmy - e) / 2, lx);
                const int yL = bound_coord((nxay - (nxmy - e)) / 2, ly);
                const int xU = bound_coord((nxay + nxmy + e) / 2, lx);
                const int yU = bound_coord((pxay - (pxmy + e)) / 2, ly);
                clip_diagonal(d, xL, yL, xU, yU, kl, kr);
            }
            w = (kr - kl) / 2 + 1;
            if (kr < kl) {  // diagonal_construct would throw (pairwiseAligner.c:22-27)
                atomicExch(err, 1);
                w = 1;
            }
        }
        // block-wide exclusive scan of w
        uint32_t v = (uint32_t) w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += n;
        }
        if (lane == 31) warp_sums[wid] = v;
        int mw = w;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mw = max(mw, __shfl_xor_sync(0xffffffffu, mw, o));
        if (lane == 0) atomicMax(&maxw_s, mw);
        __syncthreads();
        uint32_t pre = carry_s;
        for (int k = 0; k < wid; k++) pre += warp_sums[k];
        uint32_t excl = pre + v - (uint32_t) w;
        if (d <= N) {
            kl_out[job.diag_off + d] = kl;
            if (kr_out) kr_out[job.diag_off + d] = kr;
            cell_off[job.diag_off + d] = excl;
            geo_out[job.diag_off + d] = make_int2((d + kl) >> 1, (int) excl);  // {lowest x, cell offset}
        }
        __syncthreads();
        if (threadIdx.x == THREADS - 1) carry_s = pre + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        cell_off[job.diag_off + N + 1] = carry_s;
        geo_out[job.diag_off + N + 1] = make_int2(0, (int) carry_s);
        job_cells[j] = carry_s;
        job_maxw[j] = maxw_s;
    }
}

// ------------------------------------------------------------------------------------- anchor check

// One CTA per alignment: the preconditions band_constructDynamic asserts on the anchors (pairwiseAligner.c:150-158:
// strictly increasing in x and y, inside the sequences, even expansion) and whether getSplitPoints (:913-965) would
// cut the alignment (some anchor-free sub-matrix larger than `limit` cells).  status[2*i] = index of the first bad
// anchor or INT_MAX, status[2*i+1] = 1 when the alignment needs the host's split path.
// Compact anchors (include/margin_b200.h: one word per anchor, dx | dy << 12 | (expansion / 2) << 24) back to
// (x, y, expansion) triples: one CTA per alignment, block-wide inclusive scans of dx and dy.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
expand_anchors_kernel(const uint32_t *__restrict__ packed, const int64_t *__restrict__ off_a, int32_t *__restrict__ anchors) {
    const int64_t a0 = off_a[blockIdx.x], na = off_a[blockIdx.x + 1] - a0;
    __shared__ int2 warp_sums[THREADS / 32];
    __shared__ int2 carry_s;
    if (threadIdx.x == 0) carry_s = make_int2(-1, -1);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int64_t base = 0; base < na; base += THREADS) {
        const int64_t k = base + threadIdx.x;
        const uint32_t v = k < na ? packed[a0 + k] : 0u;
        int sx = (int) (v & 0xfffu), sy = (int) ((v >> 12) & 0xfffu);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int nx = __shfl_up_sync(0xffffffffu, sx, o), ny = __shfl_up_sync(0xffffffffu, sy, o);
            if (lane >= o) {
                sx += nx;
                sy += ny;
            }
        }
        if (lane == 31) warp_sums[wid] = make_int2(sx, sy);
        __syncthreads();
        int2 pre = carry_s;
        for (int w = 0; w < wid; w++) {
            pre.x += warp_sums[w].x;
            pre.y += warp_sums[w].y;
        }
        if (k < na) {
            int32_t *q = anchors + 3 * (a0 + k);
            q[0] = pre.x + sx;
            q[1] = pre.y + sy;
            q[2] = (int32_t) (v >> 24) * 2;
        }
        __syncthreads();
        if (threadIdx.x == THREADS - 1) carry_s = make_int2(pre.x + sx, pre.y + sy);
        __syncthreads();
    }
}

// Run-length form of the anchors (include/margin_b200.h: anchors_runs): one word per run,
//   bits 0-7 run = following anchors at step (+1, +1) with the same expansion, bits 8-17 dx, bits 18-27 dy (steps of the
//   run's first anchor from its predecessor, (-1, -1) before an alignment's first anchor), bits 28-31 expansion / 2.
// BAM-derived anchors are one per match column (htsIntegration.c:722-729), so a run covers a whole gap-free stretch and
// the anchors of a chunk travel in ~2 MB instead of 52 MB of triples.  One CTA per alignment, block-wide scans.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
expand_anchor_runs_kernel(const uint32_t *__restrict__ words, const int64_t *__restrict__ off_r,
                          const int64_t *__restrict__ off_a, int32_t *__restrict__ anchors, int *__restrict__ err) {
    const int64_t w0 = off_r[blockIdx.x], nw = off_r[blockIdx.x + 1] - w0;
    const int64_t a0 = off_a[blockIdx.x], na = off_a[blockIdx.x + 1] - a0;
    __shared__ int3 warp_sums[THREADS / 32];
    __shared__ int3 carry_s;  // anchors so far, x and y of the last anchor
    if (threadIdx.x == 0) carry_s = make_int3(0, -1, -1);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int64_t base = 0; base < nw; base += THREADS) {
        const int64_t k = base + threadIdx.x;
        const uint32_t v = k < nw ? words[w0 + k] : 0u;
        const int run = (int) (v & 0xffu), dx = (int) ((v >> 8) & 0x3ffu), dy = (int) ((v >> 18) & 0x3ffu);
        int sc = k < nw ? 1 + run : 0, sx = k < nw ? dx + run : 0, sy = k < nw ? dy + run : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int nc = __shfl_up_sync(0xffffffffu, sc, o), nx = __shfl_up_sync(0xffffffffu, sx, o),
                      ny = __shfl_up_sync(0xffffffffu, sy, o);
            if (lane >= o) {
                sc += nc;
                sx += nx;
                sy += ny;
            }
        }
        if (lane == 31) warp_sums[wid] = make_int3(sc, sx, sy);
        __syncthreads();
        int3 pre = carry_s;
        for (int w = 0; w < wid; w++) {
            pre.x += warp_sums[w].x;
            pre.y += warp_sums[w].y;
            pre.z += warp_sums[w].z;
        }
        if (k < nw) {
            // inclusive sums end at the run's last anchor; its first anchor is `run` steps before
            const int first = pre.x + sc - (1 + run), x0 = pre.y + sx - run, y0 = pre.z + sy - run, e = (int) (v >> 28) * 2;
            if (first + run >= na) {
                atomicExch(err, 1);  // more anchors than off_a announces
            } else {
                int32_t *q = anchors + 3 * (a0 + first);
                for (int j = 0; j <= run; j++) {
                    q[3 * j] = x0 + j;
                    q[3 * j + 1] = y0 + j;
                    q[3 * j + 2] = e;
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == THREADS - 1) carry_s = make_int3(pre.x + sc, pre.y + sx, pre.z + sy);
        __syncthreads();
    }
    if (threadIdx.x == 0 && carry_s.x != (int) na) atomicExch(err, 1);
}

__global__ void check_anchors_kernel(const int32_t *__restrict__ anchors, const int64_t *__restrict__ off_a,
                                     const int64_t *__restrict__ off_x, const int64_t *__restrict__ off_y, int64_t limit,
                                     int32_t *__restrict__ status) {
    const int64_t i = blockIdx.x;
    const int64_t a0 = off_a[i], na = off_a[i + 1] - a0;
    const int64_t lx = off_x[i + 1] - off_x[i], ly = off_y[i + 1] - off_y[i];
    const int32_t *anc = anchors + 3 * a0;
    int bad = 0x7fffffff, split = 0;
    for (int64_t k = threadIdx.x; k <= na; k += blockDim.x) {
        const int64_t px = k > 0 ? anc[3 * (k - 1)] : -1, py = k > 0 ? anc[3 * (k - 1) + 1] : -1;
        int64_t x = lx, y = ly;
        if (k < na) {
            x = anc[3 * k];
            y = anc[3 * k + 1];
            const int e = anc[3 * k + 2];
            if (x <= px || y <= py || x < 0 || y < 0 || x >= lx || y >= ly || e < 0 || (e & 1)) bad = min(bad, (int) k);
        }
        if ((x - (px + 1)) * (y - (py + 1)) > limit) split = 1;
    }
    if (bad != 0x7fffffff) atomicMin(&status[2 * i], bad);
    if (split) atomicOr(&status[2 * i + 1], 1);
}

// ------------------------------------------------------------------------------- band, chunk-parallel version

// The same geometry as band_kernel, parallel over chunks of BAND_CHUNK consecutive diagonals of a job (a 20 kb read has
// ~28 000 diagonals, so one CTA per job leaves most of the GPU idle).  Four launches:
//   band_chunk_kernel   per chunk: {lowest x, width} of every diagonal, chunk-local exclusive scan of the widths
//   band_jobs_kernel    per job (one warp): exclusive scan of the chunk sums -> chunk bases, job cells, widest diagonal
//   band_bases_kernel   one CTA: exclusive scan of the job cells -> Job::cell_base, Job::ring_w, total cells
//   band_apply_kernel   per chunk: adds the chunk base; writes geo {xlo, cell offset}, kl and cell_off
constexpr int BAND_CHUNK = 1024;

struct BandChunk {
    int32_t job;
    int32_t d0;  // first diagonal of the chunk; the chunk covers d0 .. min(d0 + BAND_CHUNK, N + 2) - 1
};

// smallest anchor index i in [lo, hi) with (ax+1)+(ay+1) >= d, else hi
__device__ __forceinline__ int anchor_lower_bound(const int32_t *__restrict__ anc, int lo, int hi, int d) {
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const int s = __ldg(anc + 3 * mid) + __ldg(anc + 3 * mid + 1) + 2;
        if (s >= d) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// The four corner points of the band between the anchor pair governing diagonal(s) with lower bound `lo`
// (band_constructDynamic, pairwiseAligner.c:161-226): the same arithmetic as band_diagonal below.
__device__ __forceinline__ void band_corners(const int32_t *__restrict__ anc, int na, int lx, int ly, int lo,
                                             const Params &p, int &xL, int &yL, int &xU, int &yU) {
    int nx = lx, ny = ly, px = 0, py = 0, e;
    if (lo < na) {
        nx = __ldg(anc + 3 * lo) + 1;
        ny = __ldg(anc + 3 * lo + 1) + 1;
        e = p.dynamic ? __ldg(anc + 3 * lo + 2) : p.expansion;
    } else {
        e = p.dynamic ? (na > 0 ? __ldg(anc + 3 * (na - 1) + 2) : 0) : p.expansion;
    }
    if (lo > 0) {
        px = __ldg(anc + 3 * (lo - 1)) + 1;
        py = __ldg(anc + 3 * (lo - 1) + 1) + 1;
    }
    const int pxay = px + py, pxmy = px - py, nxay = nx + ny, nxmy = nx - ny;
    xL = bound_coord((pxay + pxmy - e) / 2, lx);
    yL = bound_coord((nxay - (nxmy - e)) / 2, ly);
    xU = bound_coord((nxay + nxmy + e) / 2, lx);
    yU = bound_coord((pxay - (pxmy + e)) / 2, ly);
}

// `lo` = anchor_lower_bound over all anchors of the job for diagonal d
__device__ __forceinline__ void band_diagonal(const int32_t *__restrict__ anc, int na, int lx, int ly, int d, int lo,
                                              const Params &p, int &kl, int &kr) {
    kl = 0;
    kr = 0;
    if (d == 0) return;
    int nx = lx, ny = ly, px = 0, py = 0, e;
    if (lo < na) {
        nx = __ldg(anc + 3 * lo) + 1;
        ny = __ldg(anc + 3 * lo + 1) + 1;
        e = p.dynamic ? __ldg(anc + 3 * lo + 2) : p.expansion;
    } else {
        e = p.dynamic ? (na > 0 ? __ldg(anc + 3 * (na - 1) + 2) : 0) : p.expansion;
    }
    if (lo > 0) {
        px = __ldg(anc + 3 * (lo - 1)) + 1;
        py = __ldg(anc + 3 * (lo - 1) + 1) + 1;
    }
    const int pxay = px + py, pxmy = px - py, nxay = nx + ny, nxmy = nx - ny;
    const int xL = bound_coord((pxay + pxmy - e) / 2, lx);
    const int yL = bound_coord((nxay - (nxmy - e)) / 2, ly);
    const int xU = bound_coord((nxay + nxmy + e) / 2, lx);
    const int yU = bound_coord((pxay - (pxmy + e)) / 2, ly);
    clip_diagonal(d, xL, yL, xU, yU, kl, kr);
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
band_chunk_kernel(const Job *__restrict__ jobs, const BandChunk *__restrict__ chunks, int n_chunks,
                  const int32_t *__restrict__ anchors, Params p, int2 *__restrict__ geo, int32_t *__restrict__ kr_out,
                  uint32_t *__restrict__ chunk_sum, int32_t *__restrict__ chunk_maxw, int *__restrict__ err,
                  int gap_min, int gap_cap, int *__restrict__ gap_count, AnchorGap *__restrict__ gaps) {
    const int ci = blockIdx.x;
    if (ci >= n_chunks) return;
    const BandChunk ch = chunks[ci];
    const Job job = jobs[ch.job];
    const int lx = job.lx, ly = job.ly, N = lx + ly, na = job.n_anc;
    const int32_t *anc = anchors + 3 * job.anc_off;
    const int d_end = min(ch.d0 + BAND_CHUNK, N + 2);
    __shared__ uint32_t warp_sums[THREADS / 32];
    __shared__ uint32_t carry_s;
    __shared__ int maxw_s;
    // The anchors that can govern this chunk are a contiguous range [a_lo, a_hi]; their anti-diagonals are staged in
    // shared memory so that the per-diagonal search runs at shared-memory latency.
    constexpr int CAP = 2048;
    __shared__ int s_diag[CAP];
    __shared__ int a_range[2];
    if (threadIdx.x == 0) {
        carry_s = 0;
        maxw_s = 0;
        a_range[0] = anchor_lower_bound(anc, 0, na, ch.d0);
    }
    if (threadIdx.x == 32) a_range[1] = anchor_lower_bound(anc, 0, na, d_end);
    __syncthreads();
    const int a_lo = a_range[0], a_cnt = a_range[1] - a_range[0];
    const bool staged = a_cnt <= CAP;
    if (staged)
        for (int k = threadIdx.x; k < a_cnt; k += THREADS)
            s_diag[k] = __ldg(anc + 3 * (a_lo + k)) + __ldg(anc + 3 * (a_lo + k) + 1) + 2;
    __syncthreads();
    // Stretches of more than gap_min anti-diagonals without an anchor (long indels: the band is a narrow corridor in
    // which the forward/backward state forgets nothing) are reported to the host, which keeps the warm-up of the
    // tiles out of them (build_tiles).  Every anchor is owned by the chunk its diagonal falls into; the stretch after
    // the last anchor by the job's last chunk.
    if (gaps != nullptr) {
        for (int k = threadIdx.x; k < a_cnt; k += THREADS) {
            const int a = a_lo + k;
            const int dk = staged ? s_diag[k] : __ldg(anc + 3 * a) + __ldg(anc + 3 * a + 1) + 2;
            int dp = 0;
            if (k > 0 && staged) dp = s_diag[k - 1];
            else if (a > 0) dp = __ldg(anc + 3 * (a - 1)) + __ldg(anc + 3 * (a - 1) + 1) + 2;
            if (dk - dp > gap_min) {
                const int at = atomicAdd(gap_count, 1);
                if (at < gap_cap) gaps[at] = AnchorGap{ch.job, dp, dk};
            }
        }
        if (threadIdx.x == 0 && ch.d0 + BAND_CHUNK >= N + 2) {
            const int dp = na > 0 ? __ldg(anc + 3 * (na - 1)) + __ldg(anc + 3 * (na - 1) + 1) + 2 : 0;
            if (N - dp > gap_min) {
                const int at = atomicAdd(gap_count, 1);
                if (at < gap_cap) gaps[at] = AnchorGap{ch.job, dp, N};
            }
        }
    }
    // Each thread owns DPT consecutive diagonals: one binary search for the first, then the governing anchor only
    // moves forward; one block-wide scan of the per-thread width sums covers the whole chunk.
    constexpr int DPT = BAND_CHUNK / THREADS;
    static_assert(DPT * THREADS == BAND_CHUNK, "a chunk is THREADS x DPT diagonals");
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int dfirst = ch.d0 + threadIdx.x * DPT;
    int wv[DPT], xv[DPT], krv[DPT];
    int lo_a = 0;
    if (dfirst <= N) {
        if (staged) {
            int lo = 0, hi = a_cnt;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (s_diag[mid] >= dfirst) hi = mid; else lo = mid + 1;
            }
            lo_a = a_lo + lo;
        } else {
            lo_a = anchor_lower_bound(anc, 0, na, dfirst);
        }
    }
    uint32_t tsum = 0;
    int mw = 0;
    int cached_lo = -1, cxL = 0, cyL = 0, cxU = 0, cyU = 0;
#pragma unroll
    for (int k = 0; k < DPT; k++) {
        const int d = dfirst + k;
        int w = 0, kl = 0, kr = 0;
        if (d <= N) {
            if (k > 0) {  // smallest anchor index with (ax+1)+(ay+1) >= d: advances monotonically with d
                if (staged) {
                    while (lo_a - a_lo < a_cnt && s_diag[lo_a - a_lo] < d) lo_a++;
                } else {
                    while (lo_a < na && __ldg(anc + 3 * lo_a) + __ldg(anc + 3 * lo_a + 1) + 2 < d) lo_a++;
                }
            }
            if (d == 0) {
                kl = kr = 0;
            } else {
                if (lo_a != cached_lo) {  // the corner points only depend on the governing anchor pair
                    cached_lo = lo_a;
                    band_corners(anc, na, lx, ly, lo_a, p, cxL, cyL, cxU, cyU);
                }
                clip_diagonal(d, cxL, cyL, cxU, cyU, kl, kr);
            }
            w = (kr - kl) / 2 + 1;
            if (kr < kl) {  // diagonal_construct would throw (pairwiseAligner.c:22-27)
                atomicExch(err, 1);
                w = 1;
            }
        }
        wv[k] = w;
        xv[k] = d <= N ? (d + kl) >> 1 : lx;  // the sentinel entry N+1 keeps a valid column for prefetches
        krv[k] = kr;
        tsum += (uint32_t) w;
        mw = max(mw, w);
    }
    uint32_t v = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    if (lane == 31) warp_sums[wid] = v;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mw = max(mw, __shfl_xor_sync(0xffffffffu, mw, o));
    if (lane == 0) atomicMax(&maxw_s, mw);
    __syncthreads();
    uint32_t pre = 0;
    for (int k = 0; k < wid; k++) pre += warp_sums[k];
    uint32_t run = pre + v - tsum;  // exclusive prefix of this thread's first diagonal inside the chunk
#pragma unroll
    for (int k = 0; k < DPT; k++) {
        const int d = dfirst + k;
        if (d < d_end) {
            geo[job.diag_off + d] = make_int2(xv[k], (int) run);
            if (kr_out) kr_out[job.diag_off + d] = krv[k];
        }
        run += (uint32_t) wv[k];
    }
    if (threadIdx.x == THREADS - 1) carry_s = pre + v;
    __syncthreads();
    if (threadIdx.x == 0) {
        chunk_sum[ci] = carry_s;
        chunk_maxw[ci] = maxw_s;
    }
}

// One warp per job; job_first_chunk[j] .. job_first_chunk[j+1] are the job's chunks.
__global__ void band_jobs_kernel(int n_jobs, const int32_t *__restrict__ job_first_chunk, uint32_t *__restrict__ chunk_sum,
                                 const int32_t *__restrict__ chunk_maxw, int64_t *__restrict__ job_cells,
                                 int32_t *__restrict__ job_maxw) {
    const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= n_jobs) return;
    const int c0 = job_first_chunk[j], c1 = job_first_chunk[j + 1];
    uint32_t carry = 0;
    int mw = 0;
    for (int base = c0; base < c1; base += 32) {
        const int c = base + lane;
        const uint32_t s = c < c1 ? chunk_sum[c] : 0u;
        if (c < c1) mw = max(mw, chunk_maxw[c]);
        uint32_t v = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += n;
        }
        if (c < c1) chunk_sum[c] = carry + v - s;  // exclusive base of the chunk inside the job
        carry += __shfl_sync(0xffffffffu, v, 31);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mw = max(mw, __shfl_xor_sync(0xffffffffu, mw, o));
    if (lane == 0) {
        job_cells[j] = carry;
        job_maxw[j] = mw;
    }
}

// One CTA: exclusive scan of the job cells; fills the device job table and totals[0] = cells of the batch.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
band_bases_kernel(Job *__restrict__ jobs, int n_jobs, const int64_t *__restrict__ job_cells,
                  const int32_t *__restrict__ job_maxw, int64_t *__restrict__ totals) {
    __shared__ int64_t warp_sums[THREADS / 32];
    __shared__ int64_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < n_jobs; base += THREADS) {
        const int j = base + threadIdx.x;
        const int64_t s = j < n_jobs ? job_cells[j] : 0;
        int64_t v = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t n = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += n;
        }
        if (lane == 31) warp_sums[wid] = v;
        __syncthreads();
        int64_t pre = carry_s;
        for (int k = 0; k < wid; k++) pre += warp_sums[k];
        if (j < n_jobs) {
            jobs[j].cell_base = pre + v - s;
            jobs[j].ring_w = job_maxw[j];
        }
        __syncthreads();
        if (threadIdx.x == THREADS - 1) carry_s = pre + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[0] = carry_s;
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
band_apply_kernel(const Job *__restrict__ jobs, const BandChunk *__restrict__ chunks, int n_chunks,
                  const uint32_t *__restrict__ chunk_base, int2 *__restrict__ geo, int32_t *__restrict__ kl_out,
                  uint32_t *__restrict__ cell_off) {
    const int ci = blockIdx.x;
    if (ci >= n_chunks) return;
    const BandChunk ch = chunks[ci];
    const Job job = jobs[ch.job];
    const int N = job.lx + job.ly;
    const int d_end = min(ch.d0 + BAND_CHUNK, N + 2);
    const uint32_t base = chunk_base[ci];
    for (int d = ch.d0 + threadIdx.x; d < d_end; d += THREADS) {
        int2 g = geo[job.diag_off + d];
        g.y = (int) ((uint32_t) g.y + base);
        geo[job.diag_off + d] = g;
        kl_out[job.diag_off + d] = 2 * g.x - d;
        cell_off[job.diag_off + d] = (uint32_t) g.y;
    }
}

// -------------------------------------------------------------------------------------- schedule

// One warp per job: replays the traceback trigger of getPosteriorProbsWithBanding (pairwiseAligner.c:746-768),
// which depends only on band widths.  Segment s of a job is stored at seg_base + s; the host cuts the segments into
// tiles and orders them the way the reference's final lists are ordered (later tracebacks first, :1104-1123).
__global__ void schedule_kernel(const Job *__restrict__ jobs, int n_jobs, const uint32_t *__restrict__ cell_off, Params p,
                                Seg *__restrict__ segs, int32_t *__restrict__ job_nseg) {
    const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= n_jobs) return;
    const Job job = jobs[j];
    const int N = job.lx + job.ly;
    const uint32_t *co = cell_off + job.diag_off;
    int tb_to = 0, s = 0;
    const int wmax = 2 * p.expansion + 1;
    while (N > 0) {
        int D = N, at_end = 1;
        for (int base = tb_to + p.min_diags; base < N; base += 32) {
            const int d = base + lane;
            bool hit = d < N && (int) (co[d + 1] - co[d]) <= wmax;
            unsigned b = __ballot_sync(0xffffffffu, hit);
            if (b) {
                D = base + __ffs(b) - 1;
                at_end = 0;
                break;
            }
        }
        const int tb_from = at_end ? D : D - (p.tb_diags + 1);
        if (lane == 0) {
            Seg sg;
            sg.job = j;
            sg.D = D;
            sg.tb_to = tb_to;
            sg.tb_from = tb_from;
            sg.at_end = at_end;
            if (s < job.max_seg) segs[job.seg_base + s] = sg;
        }
        s++;
        tb_to = tb_from;
        if (at_end) break;
    }
    if (lane == 0) job_nseg[j] = s;
}


// --------------------------------------------------------------------------------------- forward

template <int G>
__device__ __forceinline__ unsigned group_mask() {
    if constexpr (G == 32) {
        return 0xffffffffu;
    } else {
        return ((1u << G) - 1u) << ((threadIdx.x & 31) / G * G);
    }
}

template <int G>
__device__ __forceinline__ Cell shfl_cell(unsigned mask, const Cell &c, int src) {
    Cell r;
    r.m = __shfl_sync(mask, c.m, src, G);
    r.x = __shfl_sync(mask, c.x, src, G);
    r.y = __shfl_sync(mask, c.y, src, G);
    return r;
}

__device__ __forceinline__ Cell load_cell_cg(const double *base, int64_t cell) {
    Cell c;
    const double *q = base + 3 * cell;
    c.m = __ldcg(q);
    c.x = __ldcg(q + 1);
    c.y = __ldcg(q + 2);
    return c;
}

__device__ __forceinline__ void store_cell(double *base, int64_t cell, const Cell &c) {
    double *q = base + 3 * cell;
    q[0] = c.m;
    q[1] = c.x;
    q[2] = c.y;
}

__device__ __forceinline__ unsigned sym_at(const uint16_t *s, int i) {  // getXCharacter / getYCharacter (:535-545)
    return i > 0 ? (unsigned) __ldg(s + i - 1) : 4u;
}

// One forward cell: stateMachine3_cellCalculate + doTransitionForward (stateMachine.c:562-586, pairwiseAligner.c:311-314).
// Each target state folds its three sources left to right starting from -inf; logAdd(-inf, v) == v.
__device__ __forceinline__ Cell forward_cell(const DevModel &m, bool has_lower, const Cell &lower, bool has_middle,
                                             const Cell &middle, bool has_upper, const Cell &upper, unsigned cx,
                                             unsigned cy) {
    Cell c = cell_neg_inf();
    if (has_lower) {
        const double *e = m.ex_t[gap_index(m, cx)];
        double a = dadd(lower.m, e[0]);
        a = logadd(a, dadd(lower.x, e[1]));
        a = logadd(a, dadd(lower.y, e[2]));
        c.x = a;
    }
    if (has_middle) {
        const double e = emit_match(m, cx, cy);
        double a = dadd(middle.m, dadd(e, m.t[T_MM]));
        a = logadd(a, dadd(middle.x, dadd(e, m.t[T_XM])));
        a = logadd(a, dadd(middle.y, dadd(e, m.t[T_YM])));
        c.m = a;
    }
    if (has_upper) {
        const double *e = m.ey_t[gap_index(m, cy)];
        double a = dadd(upper.m, e[0]);
        a = logadd(a, dadd(upper.y, e[1]));
        a = logadd(a, dadd(upper.x, e[2]));
        c.y = a;
    }
    return c;
}

__device__ __forceinline__ int mod_g(int v, int G) { return v & (G - 1); }

// Forward sweep.  One group of G lanes per work item; lane = x mod G.  F (3 doubles per band cell) goes to HBM for the
// backward kernel.  Neighbour diagonals no wider than G stay in registers and move by one rotate-shuffle.
// Work items: id >= 0 is the whole job `id`, swept from the true start state on diagonal 0 to its last diagonal, F in
// the batch's shared array; id < 0 is the TILE_WIDE tile ~id, swept from a flat state on its diagonal f0 (the true start
// state when f0 == 0) to tb_from, F in the tile's private array (see Tile).
template <int G, int THREADS>
__global__ void __launch_bounds__(THREADS)
forward_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids, int n_ids,
               const DevModel *__restrict__ models, const uint16_t *__restrict__ symx,
               const uint16_t *__restrict__ symy, const int32_t *__restrict__ kl_arr,
               const uint32_t *__restrict__ cell_off, double *__restrict__ F, double *__restrict__ Fpriv) {
    __shared__ DevModel sm[2];
    for (int i = threadIdx.x; i < (int) (2 * sizeof(DevModel) / 8); i += THREADS)
        reinterpret_cast<double *>(sm)[i] = reinterpret_cast<const double *>(models)[i];
    __syncthreads();
    const int g = (blockIdx.x * THREADS + threadIdx.x) / G;
    if (g >= n_ids) return;
    const unsigned gm = group_mask<G>();
    const int lane = threadIdx.x & (G - 1);
    const int id = ids[g];
    int d_first = 0, d_last = -1;
    int64_t priv = -1;
    int ji = id;
    if (id < 0) {
        const Tile t = tiles[~id];
        ji = t.job;
        d_first = t.f0;
        d_last = t.tb_from;
        priv = t.priv_off;
    }
    const Job job = jobs[ji];
    const DevModel &m = sm[job.slot];
    const int N = job.lx + job.ly;
    if (id >= 0) d_last = N;
    const uint16_t *sx = symx + job.symx_off, *sy = symy + job.symy_off;
    const int32_t *kls = kl_arr + job.diag_off;
    const uint32_t *co = cell_off + job.diag_off;
    double *Fj = id >= 0 ? F + 3 * job.cell_base : Fpriv + 3 * (priv - (int64_t) co[d_first]);

    Cell f1 = cell_neg_inf(), f2 = cell_neg_inf();
    int xlo1 = 0, w1 = 1, xlo2 = 0, w2 = 0;
    uint32_t co1 = 0, co2 = 0;
    if (d_first == 0) {
        // diagonal 0: the single cell (0,0) holds the start state (stateMachine.c:521-530)
        if (lane == 0) {
            f1.m = job.ragged_l ? MB_NEG_INF : 0.0;
            f1.x = job.ragged_l ? 0.0 : MB_NEG_INF;
            f1.y = f1.x;
            store_cell(Fj, 0, f1);
        }
    } else {
        // a tile's warm-up starts from a flat state on every cell of diagonal f0
        co1 = co[d_first];
        w1 = (int) (co[d_first + 1] - co1);
        xlo1 = (d_first + kls[d_first]) >> 1;
        for (int jc = 0; jc * G < w1; jc++) {
            const int r = mod_g(lane - xlo1, G) + jc * G;
            if (r < w1) {
                f1.m = f1.x = f1.y = 0.0;
                store_cell(Fj, (int64_t) co1 + r, f1);
            }
        }
    }
    const int prev_lane = mod_g(lane - 1, G);
    for (int d = d_first + 1; d <= d_last; d++) {
        const int kl = kls[d];
        const uint32_t cod = co[d];
        const int w = (int) (co[d + 1] - cod);
        const int xlo = (d + kl) >> 1;
        const int xhi = xlo + w - 1;
        const int nch = (w + G - 1) / G;
        if (w1 > G || w2 > G) __syncwarp(gm);  // make the earlier diagonals' stores visible to the group
        Cell cur = cell_neg_inf();
        for (int jc = 0; jc < nch; jc++) {
            const int x = xlo + mod_g(lane - xlo, G) + jc * G;
            const int y = d - x;
            const bool active = x <= xhi;
            const bool has_lower = active && (x - 1 >= xlo1) && (x - 1 < xlo1 + w1);
            const bool has_upper = active && (x >= xlo1) && (x < xlo1 + w1);
            const bool has_middle = active && (w2 > 0) && (x - 1 >= xlo2) && (x - 1 < xlo2 + w2);
            Cell lower, upper, middle;
            if (w1 <= G) {
                lower = shfl_cell<G>(gm, f1, prev_lane);
                upper = f1;
            } else {
                if (has_lower) lower = load_cell_cg(Fj, (int64_t) co1 + (x - 1 - xlo1));
                if (has_upper) upper = load_cell_cg(Fj, (int64_t) co1 + (x - xlo1));
            }
            if (w2 <= G) {
                middle = shfl_cell<G>(gm, f2, prev_lane);
            } else {
                if (has_middle) middle = load_cell_cg(Fj, (int64_t) co2 + (x - 1 - xlo2));
            }
            if (active) {
                Cell c = forward_cell(m, has_lower, lower, has_middle, middle, has_upper, upper, sym_at(sx, x),
                                      sym_at(sy, y));
                store_cell(Fj, (int64_t) cod + (x - xlo), c);
                cur = c;
            }
        }
        f2 = f1;
        xlo2 = xlo1;
        w2 = w1;
        co2 = co1;
        f1 = cur;
        xlo1 = xlo;
        w1 = w;
        co1 = cod;
    }
}


// -------------------------------------------------------------------------------------- backward

struct Pair {
    int32_t w, x, y;
};

// Final tuple store: (w, x, y) int32 triples, or MARGIN_B200_PAIRS_PACKED64 words w | (x+1) << 24 | (y+1) << 44.
__device__ __forceinline__ void store_pair(void *out, int64_t i, int packed, const Pair &q) {
    if (packed)
        reinterpret_cast<unsigned long long *>(out)[i] = (unsigned long long) (uint32_t) q.w |
                                                         ((unsigned long long) (uint32_t) (q.x + 1) << 24) |
                                                         ((unsigned long long) (uint32_t) (q.y + 1) << 44);
    else
        reinterpret_cast<Pair *>(out)[i] = q;
}

// Gather form of diagonalCalculationBackward (pairwiseAligner.c:572-576 with doTransitionBackward :322-325).
// The reference scatters from a cell into its three neighbours; a cell (x,y) of diagonal d therefore receives,
// in this order: its match term from (x+1,y+1) while d+2 is processed, then while d+1 is processed (cells in
// ascending x-y) the gap-Y term from (x,y+1) and the gap-X term from (x+1,y).  Folding in that order from -inf
// reproduces the scatter bit for bit.
__device__ __forceinline__ Cell backward_cell(const DevModel &m, bool has_mm, double b2m, bool has_y, double b1y_y,
                                              bool has_x, double b1x_x, unsigned cx1, unsigned cy1) {
    // cx1 = sX[x] (the symbol consumed when moving to x+1), cy1 = sY[y]
    Cell c = cell_neg_inf();
    if (has_mm) {
        const double e = emit_match(m, cx1, cy1);
        c.m = dadd(b2m, dadd(e, m.t[T_MM]));
        c.x = dadd(b2m, dadd(e, m.t[T_XM]));
        c.y = dadd(b2m, dadd(e, m.t[T_YM]));
    }
    if (has_y) {
        const double *e = m.ey_t[gap_index(m, cy1)];  // [0]=M->Y, [1]=Y->Y, [2]=X->Y
        c.m = logadd(c.m, dadd(b1y_y, e[0]));
        c.x = logadd(c.x, dadd(b1y_y, e[2]));
        c.y = logadd(c.y, dadd(b1y_y, e[1]));
    }
    if (has_x) {
        const double *e = m.ex_t[gap_index(m, cx1)];  // [0]=M->X, [1]=X->X, [2]=Y->X
        c.m = logadd(c.m, dadd(b1x_x, e[0]));
        c.x = logadd(c.x, dadd(b1x_x, e[1]));
        c.y = logadd(c.y, dadd(b1x_x, e[2]));
    }
    return c;
}

// Sequential (non-associative!) logAdd fold of one value per band cell, cells in ascending x, as
// dpDiagonal_dotProduct does (pairwiseAligner.c:450-461).  `val` is this lane's value for chunk jc.
template <int G>
__device__ __forceinline__ double fold_chunk(unsigned gm, double acc, double val, int xlo, int jc, int w) {
    const int n = min(G, w - jc * G);
    for (int i = 0; i < n; i++) {
        const double v = __shfl_sync(gm, val, mod_g(xlo + i, G), G);
        acc = logadd(acc, v);
    }
    return acc;
}

__device__ __forceinline__ bool posterior_weight(double f, double b, double total, double threshold, int32_t &w) {
    // exp((F + B) - total), threshold, clamp, floor(p * 1e7): pairwiseAligner.c:598-608, :629-631
    double pr = exp(dadd(dadd(f, b), -total));
    if (!(pr >= threshold)) return false;
    if (pr > 1.0) pr = 1.0;
    w = (int32_t) floor(dmul(pr, 10000000.0));
    return true;
}

template <int G>
__device__ __forceinline__ void emit_list(unsigned gm, bool ok, int32_t wt, int x, int y, int rank, int xlo, Pair *dst,
                                          int &cnt) {
    unsigned bits = __ballot_sync(gm, ok);
    const int r = mod_g(xlo, G);
    unsigned rot;
    if constexpr (G == 32) {
        rot = __funnelshift_r(bits, bits, r);
    } else {
        bits = (bits >> ((threadIdx.x & 31) / G * G)) & ((1u << G) - 1u);
        rot = ((bits >> r) | (bits << (G - r))) & ((1u << G) - 1u);
    }
    if (ok) {
        const int pos = cnt + __popc(rot & ((1u << rank) - 1u));
        Pair pr;
        pr.w = wt;
        pr.x = x;
        pr.y = y;
        dst[pos] = pr;
    }
    cnt += __popc(bits);
}

// One group of G lanes per traceback segment (the walk-back loop of getPosteriorProbsWithBanding,
// pairwiseAligner.c:753-818).  Segments are independent once F is in HBM, so all tracebacks of all reads run
// concurrently.  Tuples are staged per segment in emission order; gather_kernel puts them in list order.
template <int G, int THREADS, bool INDELS>
__global__ void __launch_bounds__(THREADS)
backward_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ segs, const int32_t *__restrict__ slot_ids,
                int n_ids, const DevModel *__restrict__ models, const uint16_t *__restrict__ symx,
                const uint16_t *__restrict__ symy, const int32_t *__restrict__ kl_arr,
                const uint32_t *__restrict__ cell_off, const double *__restrict__ F, const double *__restrict__ Fpriv,
                Params p, double *__restrict__ ring, const int64_t *__restrict__ slot_ring_off, Pair *__restrict__ stage,
                int64_t stage_stride, int64_t *__restrict__ seg_counts, int total_slots, double *__restrict__ bnd,
                const int64_t *__restrict__ slot_bnd_off) {
    __shared__ DevModel sm[2];
    for (int i = threadIdx.x; i < (int) (2 * sizeof(DevModel) / 8); i += THREADS)
        reinterpret_cast<double *>(sm)[i] = reinterpret_cast<const double *>(models)[i];
    __syncthreads();
    const int g = (blockIdx.x * THREADS + threadIdx.x) / G;
    if (g >= n_ids) return;
    const int slot = slot_ids[g];
    const Tile seg = segs[slot];
    const unsigned gm = group_mask<G>();
    const int lane = threadIdx.x & (G - 1);
    const Job job = jobs[seg.job];
    const DevModel &m = sm[job.slot];
    const uint16_t *sx = symx + job.symx_off, *sy = symy + job.symy_off;
    const int32_t *kls = kl_arr + job.diag_off;
    const uint32_t *co = cell_off + job.diag_off;
    const bool wide = (seg.flags & TILE_WIDE) != 0;
    const double *Fj = wide ? Fpriv + 3 * (seg.priv_off - (int64_t) co[seg.f0]) : F + 3 * job.cell_base;
    const int ring_w = job.ring_w;
    double *rg = ring + 3 * slot_ring_off[slot];
    const int D = seg.D, tb_to = seg.tb_to, tb_from = seg.tb_from;
    const double *endp = m.end_p[((seg.flags & TILE_AT_END) && job.ragged_r) ? 1 : 0];
    Pair *stM = stage + seg.stage_off, *stX = stage + stage_stride + seg.stage_off,
         *stY = stage + 2 * stage_stride + seg.stage_off;
    int cntM = 0, cntX = 0, cntY = 0;
    // seam guard (TILE_WIDE only): B(tb_to + 1) and B(tb_from + 1) go to the tile's boundary sections
    const int64_t bnd_cells = wide ? (slot_bnd_off[slot + 1] - slot_bnd_off[slot]) / 3 : 0;
    double *bnd_t = bnd + 3 * slot_bnd_off[slot];

    Cell b1 = cell_neg_inf(), b2 = cell_neg_inf();
    int xlo1 = 0, w1 = 0, xlo2 = 0, w2 = 0;
    const int next_lane = mod_g(lane + 1, G);
    double total = MB_NEG_INF;
    int done = 0;
    for (int d = D; d > tb_to; d--) {
        const int kl = kls[d];
        const uint32_t cod = co[d];
        const int w = (int) (co[d + 1] - cod);
        const int xlo = (d + kl) >> 1;
        const int xhi = xlo + w - 1;
        const int nch = (w + G - 1) / G;
        const bool post = d <= tb_from;
        const bool do_total = post && (done % 10 == 0);
        if (w > G || w1 > G || w2 > G) __syncwarp(gm);
        double *rg_d = rg + 3 * (int64_t) ((d % 3) * ring_w);
        const double *rg_d1 = rg + 3 * (int64_t) (((d + 1) % 3) * ring_w);
        const double *rg_d2 = rg + 3 * (int64_t) (((d + 2) % 3) * ring_w);
        Cell cur = cell_neg_inf();
        double first = MB_NEG_INF;
        // ---- pass 1: backward values of diagonal d (+ first term of the total probability)
        for (int jc = 0; jc < nch; jc++) {
            const int x = xlo + mod_g(lane - xlo, G) + jc * G;
            const int y = d - x;
            const bool active = x <= xhi;
            Cell b;
            if (d == D) {
                b.m = endp[0];
                b.x = endp[1];
                b.y = endp[2];
            } else {
                const bool has_mm = active && w2 > 0 && (x + 1 >= xlo2) && (x + 1 < xlo2 + w2);
                const bool has_y = active && (x >= xlo1) && (x < xlo1 + w1);
                const bool has_x = active && (x + 1 >= xlo1) && (x + 1 < xlo1 + w1);
                double b2m = 0, b1yy = 0, b1xx = 0;
                if (w2 > 0) {
                    if (w2 <= G) b2m = __shfl_sync(gm, b2.m, next_lane, G);
                    else if (has_mm) b2m = __ldcg(rg_d2 + 3 * (x + 1 - xlo2));
                }
                if (w1 <= G) {
                    b1yy = b1.y;
                    b1xx = __shfl_sync(gm, b1.x, next_lane, G);
                } else {
                    if (has_y) b1yy = __ldcg(rg_d1 + 3 * (x - xlo1) + 2);
                    if (has_x) b1xx = __ldcg(rg_d1 + 3 * (x + 1 - xlo1) + 1);
                }
                b = cell_neg_inf();
                if (active) b = backward_cell(m, has_mm, b2m, has_y, b1yy, has_x, b1xx, sym_at(sx, x + 1), sym_at(sy, y + 1));
            }
            if (active) {
                cur = b;
                if (w > G) {
                    double *q = rg_d + 3 * (x - xlo);
                    q[0] = b.m;
                    q[1] = b.x;
                    q[2] = b.y;
                }
                if (wide && (d == tb_to + 1 || (d == tb_from + 1 && d < D)) && x - xlo < bnd_cells)
                    store_cell(bnd_t, (int64_t) (d == tb_to + 1 ? BND_B_LOW : BND_B_HIGH) * bnd_cells + (x - xlo), b);
            }
            if (do_total) {
                double dot = MB_NEG_INF;
                if (active) {
                    Cell f = load_cell_cg(Fj, (int64_t) cod + (x - xlo));
                    dot = cell_dot(f, b);
                }
                first = fold_chunk<G>(gm, first, dot, xlo, jc, w);
            }
        }
        if (post) {
            if (do_total) {
                // second term (pairwiseAligner.c:586-594): paths that jump over d with a match, i.e. the match block
                // of the forward recursion from F(d-1) into the cells of d+1, dotted with B(d+1).
                total = first;
                if (d < D) {
                    double second = MB_NEG_INF;
                    const int klm = kls[d - 1];
                    const uint32_t com = co[d - 1];
                    const int wm = (int) (co[d] - com);
                    const int xlom = (d - 1 + klm) >> 1;
                    const int nch1 = (w1 + G - 1) / G;
                    for (int jc = 0; jc < nch1; jc++) {
                        const int x = xlo1 + mod_g(lane - xlo1, G) + jc * G;
                        const int y = d + 1 - x;
                        const bool active = x < xlo1 + w1;
                        double val = MB_NEG_INF;
                        if (active) {
                            double gm_ = MB_NEG_INF;
                            if (x - 1 >= xlom && x - 1 < xlom + wm) {
                                Cell f = load_cell_cg(Fj, (int64_t) com + (x - 1 - xlom));
                                const double e = emit_match(m, sym_at(sx, x), sym_at(sy, y));
                                gm_ = dadd(f.m, dadd(e, m.t[T_MM]));
                                gm_ = logadd(gm_, dadd(f.x, dadd(e, m.t[T_XM])));
                                gm_ = logadd(gm_, dadd(f.y, dadd(e, m.t[T_YM])));
                            }
                            double b1m = (w1 <= G) ? b1.m : __ldcg(rg_d1 + 3 * (x - xlo1));
                            val = dadd(gm_, b1m);
                        }
                        second = fold_chunk<G>(gm, second, val, xlo1, jc, w1);
                    }
                    total = logadd(total, second);
                }
            }
            // ---- pass 2: posteriors of diagonal d (diagonalCalculationPosteriorProbs, :637-681)
            for (int jc = 0; jc < nch; jc++) {
                const int rank = mod_g(lane - xlo, G);
                const int x = xlo + rank + jc * G;
                const int y = d - x;
                const bool active = x <= xhi;
                Cell b = cur, f = cell_neg_inf();
                if (active) {
                    if (w > G) b = load_cell_cg(rg_d, x - xlo);
                    f = load_cell_cg(Fj, (int64_t) cod + (x - xlo));
                }
                int32_t wt = 0;
                bool ok = active && x > 0 && y > 0 && posterior_weight(f.m, b.m, total, p.threshold, wt);
                emit_list<G>(gm, ok, wt, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stM, cntM);
                if (INDELS) {
                    ok = active && x > 0 && posterior_weight(f.x, b.x, total, p.threshold, wt);
                    emit_list<G>(gm, ok, wt, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stX, cntX);
                    ok = active && y > 0 && posterior_weight(f.y, b.y, total, p.threshold, wt);
                    emit_list<G>(gm, ok, wt, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stY, cntY);
                }
            }
            done++;
        }
        b2 = b1;
        xlo2 = xlo1;
        w2 = w1;
        b1 = cur;
        xlo1 = xlo;
        w1 = w;
    }
    if (lane == 0) {
        seg_counts[slot] = cntM;
        seg_counts[total_slots + slot] = cntX;
        seg_counts[2 * total_slots + slot] = cntY;
    }
}


// ---------------------------------------------------------------------------------------- gather

// Moves the staged tuples of every segment into the final lists.  seg_off is the exclusive scan of seg_counts over
// [list][slot]; slots are already in final order (schedule_kernel), and inside a segment the reference's pop loop
// (pairwiseAligner.c:1104-1123) reverses the emission order.
__global__ void gather_kernel(const Tile *__restrict__ segs, int total_slots, const int64_t *__restrict__ seg_counts,
                              const int64_t *__restrict__ seg_off, const Pair *__restrict__ stage, int64_t stage_stride,
                              void *__restrict__ out, int packed, int n_lists, int64_t out_capacity) {
    const int slot = blockIdx.x;
    if (seg_off[(int64_t) 3 * total_slots] > out_capacity) return;  // the host re-runs with a larger buffer
    const Tile seg = segs[slot];
    if (seg.flags & TILE_BND) return;  // tiled jobs are gathered by gather_staged_kernel
    const int64_t region = seg.stage_off;
    for (int l = 0; l < n_lists; l++) {
        const int64_t c = seg_counts[(int64_t) l * total_slots + slot];
        const int64_t o = seg_off[(int64_t) l * total_slots + slot];
        const Pair *src = stage + (int64_t) l * stage_stride + region;
        for (int64_t i = threadIdx.x; i < c; i += blockDim.x) store_pair(out, o + i, packed, src[c - 1 - i]);
    }
}

// Per-alignment list offsets (relative to the start of each list) from the scanned slot offsets.
__global__ void list_offsets_kernel(const int32_t *__restrict__ aln_first_slot, int n_aln, int total_slots,
                                    const int64_t *__restrict__ seg_off, int64_t *__restrict__ off_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_aln) return;
    for (int l = 0; l < 3; l++) {
        const int64_t base = seg_off[(int64_t) l * total_slots];
        const int64_t v = i < n_aln ? seg_off[(int64_t) l * total_slots + aln_first_slot[i]]
                                    : seg_off[(int64_t) (l + 1) * total_slots];
        off_out[(int64_t) l * (n_aln + 1) + i] = v - base;
    }
}

// computeForwardProbability's last step (pairwiseAligner.c:881-891): the only cell of the last diagonal is
// (lX, lY); total = cell_dotProduct(F, end state probs).  An empty problem returns LOG_ONE (:860-862).
__global__ void final_dot_kernel(const Job *__restrict__ jobs, const int32_t *__restrict__ ids, int n_ids,
                                 const DevModel *__restrict__ models, const uint32_t *__restrict__ cell_off,
                                 const double *__restrict__ F, double *__restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_ids) return;
    const Job job = jobs[ids[k]];
    const int N = job.lx + job.ly;
    if (N == 0) {
        out[job.aln] = 0.0;
        return;
    }
    const DevModel &m = models[job.slot];
    Cell f = load_cell_cg(F + 3 * job.cell_base, cell_off[job.diag_off + N]);
    Cell e;
    e.m = m.end_p[job.ragged_r ? 1 : 0][0];
    e.x = m.end_p[job.ragged_r ? 1 : 0][1];
    e.y = m.end_p[job.ragged_r ? 1 : 0][2];
    // dpDiagonal_dotProduct folds from -inf: logAdd(-inf, v) == v
    out[job.aln] = cell_dot(f, e);
}

}  // namespace mb
