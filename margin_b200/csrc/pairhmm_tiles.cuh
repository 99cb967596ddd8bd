// pairhmm_tiles.cuh — the tiled (sequence-parallel) forward / backward kernels of the banded pair-HMM (sm_100a).
//
// Why tiles.  A 20 kb read is ~28 000 dependent anti-diagonals of ~5 cells: one alignment offers almost no
// parallelism and a chunk has only ~300 of them, so a sweep per alignment is latency bound (round-1 v1: 1.6 GCUPS).
// The posteriors exp(F + B - total) are invariant under a constant shift of F or of B, and both recurrences forget
// their start after a few dozen diagonals of a banded alignment (the reference relies on exactly this when it
// restarts the backward pass at every traceback point with the end state on EVERY cell and throws the first
// traceBackDiagonals+1 = 41 diagonals away, impl/pairwiseAligner.c:753-768).  So the diagonals of a traceback
// segment are cut into tiles of ~L diagonals; each tile runs its own forward sweep, started `warm-up` diagonals
// early from a flat state, and its own backward sweep, started `warm-up` diagonals late (or exactly at the
// reference's traceback diagonal D with the reference's initial state when that is closer).  Tiles are independent,
// which gives ~N/L times more parallel work.  Where a tile reaches the true start (diagonal 0) or the reference's D
// the computation is operation-for-operation the reference's.
//
// Inside a tile: a group of G lanes (G >= widest diagonal of the job), lane = x mod G.  Neighbour (x, y-1) is the
// lane's own previous value, (x-1, y) comes from lane-1 by one rotate-shuffle, and (x-1, y-1) is the value the
// lane received from lane-1 one diagonal earlier, so a forward step moves 3 doubles, a backward step 2.
// Band geometry, symbols and the folded match emission are fetched 3 / 2 / 1 diagonals ahead (register software
// pipeline), so no load sits on the recurrence's critical path.  All groups of a warp run the same number of
// iterations (the longest tile's; tiles are sorted by length), so every shuffle is a plain full-warp SHFL.
//
// Symbols are pre-packed to one byte, p = min(base,4)*51 + runLength (pack_symbols_kernel), which makes the match
// emission one table read emitP[pX*255 + pY] and removes every clamp: reads one or two symbols outside a sequence
// land on neighbouring sequences or on the zeroed slack at the ends of the arrays and are never used.
#pragma once
#include "pairhmm_kernels.cuh"

#ifndef MB_TOTAL_ON
#define MB_TOTAL_ON true
#endif
namespace mb {

constexpr int PACK_STRIDE = 5 * MARGIN_B200_MAX_RL;  // 255 packed symbol values
constexpr int PACK_SLACK = 64;                       // zeroed bytes in front of and behind the packed arrays
constexpr int FAR_X = 0x40000000;                    // an x no band window contains

// threadIdx.x read once (ptxas otherwise re-reads the special register inside the loops)
__device__ __forceinline__ int tid_x_once() {
    int v;
    asm volatile("mov.u32 %0, %%tid.x;" : "=r"(v));
    return v;
}

// uint16 Symbol -> packed byte.  rle: base = low 8 bits, run length = high 8 bits (impl/stateMachine.c:116-131);
// otherwise the whole symbol is the base (anything >= 4 is ambiguous) and the run length slot is 0.
__global__ void pack_symbols_kernel(const uint16_t *__restrict__ in, uint8_t *__restrict__ out, int64_t n, int rle,
                                    int *__restrict__ err) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned c = in[i];
    const unsigned b = min(rle ? (c & 255u) : c, 4u);
    const unsigned r = rle ? (c >> 8) : 0u;
    if (r >= (unsigned) MARGIN_B200_MAX_RL) atomicExch(err, 2);
    out[i] = (uint8_t) (b * MARGIN_B200_MAX_RL + min(r, (unsigned) MARGIN_B200_MAX_RL - 1u));
}

__device__ __forceinline__ int packed_base(unsigned p) { return (int) ((p * 1286u) >> 16); }  // p / 51 for p < 255

// packed byte -> uint16 Symbol, for the general kernels (wide tiles, sequential sweeps) when the host sent packed symbols
__global__ void unpack_symbols_kernel(const uint8_t *__restrict__ in, uint16_t *__restrict__ out, int64_t n, int rle,
                                      int *__restrict__ err) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned p = in[i];
    if (p >= (unsigned) PACK_STRIDE) atomicExch(err, 2);
    const unsigned b = (p * 1286u) >> 16, r = p - b * MARGIN_B200_MAX_RL;
    out[i] = (uint16_t) (rle ? ((r << 8) | b) : b);
}

__device__ __forceinline__ Cell cell_sel(bool p, const Cell &c) {
    Cell r;
    r.m = p ? c.m : MB_NEG_INF;
    r.x = p ? c.x : MB_NEG_INF;
    r.y = p ? c.y : MB_NEG_INF;
    return r;
}

template <int G>
__device__ __forceinline__ double shfl_g(double v, int src) { return __shfl_sync(0xffffffffu, v, src, G); }

// lookup() (pairwiseAligner.c:282-293) picks its cubic by d <= 1, <= 2.5, <= 4.5 (< 7.5): all multiples of 1/2, so
// ceil(2 d) in 0..16 names the piece exactly, and fma.rp(d, 2, 1.5 * 2^52) leaves that integer in the low word of the
// sum — one FP64 instruction instead of three compares and three selects.  The coefficient table therefore has one row
// per value of ceil(2 d) (row 16: d > 7.5, never used for a result), and eight copies of every row, [row][copy]: lane j
// of a quarter-warp reads copy j, which lies in its own four banks, so the LDS.128 never conflicts.
constexpr int LA_ROWS = 17;
__device__ __forceinline__ int la_piece_of_row(int r) { return r <= 2 ? 0 : (r <= 5 ? 1 : (r <= 9 ? 2 : (r <= 15 ? 3 : 4))); }

struct TileShared {
    double2 lac32[LA_ROWS][8];  // {c3, c2} of the piece of row ceil(2 d)
    double2 lac10[LA_ROWS][8];  // {c1, c0}
    double gx[2][5][4];   // [model][base] {M->X, X->X, Y->X, 0}: the reference's (eP + tP) of the gap-X transitions
    double gy[2][5][4];   // [model][base] {M->Y, Y->Y, X->Y, 0}
    DevModel model[2];
};

template <int THREADS>
__device__ __forceinline__ void tile_shared_init(TileShared &s, const DevModel *__restrict__ models) {
    for (int i = threadIdx.x; i < (int) (2 * sizeof(DevModel) / 8); i += THREADS)
        reinterpret_cast<double *>(s.model)[i] = reinterpret_cast<const double *>(models)[i];
    for (int i = threadIdx.x; i < LA_ROWS * 8; i += THREADS) {
        const int p = la_piece_of_row(i >> 3);
        s.lac32[i >> 3][i & 7] = p < 4 ? make_double2(LA_COEF[4 * p], LA_COEF[4 * p + 1]) : make_double2(0.0, 0.0);
        s.lac10[i >> 3][i & 7] = p < 4 ? make_double2(LA_COEF[4 * p + 2], LA_COEF[4 * p + 3]) : make_double2(0.0, 0.0);
    }
    if (threadIdx.x < 40) {
        const int k = threadIdx.x & 3, b = (threadIdx.x >> 2) % 5, md = threadIdx.x / 20;
        s.gx[md][b][k] = k < 3 ? models[md].ex_t[b][k] : 0.0;
        s.gy[md][b][k] = k < 3 ? models[md].ey_t[b][k] : 0.0;
    }
    __syncthreads();
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ double2 lds_f64x2(uint32_t a) {
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}

__device__ __forceinline__ double lds_f64(uint32_t a) {
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}

// logAdd (pairwiseAligner.c:279-299) for the tile kernels, operation-for-operation the arithmetic of logadd():
//   |x - y| is the reference's (max - min) bit for bit, so the difference is formed first and the hi/lo selects run
//   beside the polynomial instead of in front of it; the piece comes from the row ceil(2 d) (see TileShared);
//   lo == -inf gives d = +inf (or NaN when both are -inf), "d < 7.5" fails and hi comes out.
// lac_s = shared-space address of this lane's copy of TileShared::lac32 (copy = lane & 7).
__device__ __forceinline__ double logadd_t(double x, double y, uint32_t lac_s) {
    const double ds = dadd(x, -y);
    const bool lt = __double2hiint(ds) < 0;
    const double d = fabs(ds);
    const bool small = d < 7.5;
    // lo when the polynomial counts, hi when it does not: one select instead of hi, lo and the final choice
    const double base = (lt == small) ? x : y;
    double t;
    asm("fma.rp.f64 %0, %1, 0d4000000000000000, 0d4338000000000000;" : "=d"(t) : "d"(d));
    const uint32_t row = min((uint32_t) __double2loint(t), (uint32_t) (LA_ROWS - 1));  // inf, NaN, huge d: any row will do
    const uint32_t a = lac_s + row * 128u;
    const double2 c32 = lds_f64x2(a);
    const double2 c10 = lds_f64x2(a + LA_ROWS * 128u);
    double v = dadd(dmul(c32.x, d), c32.y);
    v = dadd(dmul(v, d), c10.x);
    v = dadd(dmul(v, d), c10.y);
    return dadd(small ? v : 0.0, base);  // v + lo, or hi + 0 = hi
}

// A staged posterior: the log posterior v = F + B - total of a cell state that passes the threshold; the gather
// kernel turns it into the reference's weight floor(min(exp(v), 1) * 1e7) (pairwiseAligner.c:598-608).
struct StagedPair {
    double v;
    int32_t x, y;
};

// ------------------------------------------------------------------------------------------------ forward tile
//
// Register software pipeline, period 4 (the loop body is four diagonals, so no value is ever moved between
// registers): ring g[] holds the geometry of d .. d+3, the symbols of d+2 are fetched while d is computed, the
// match emission and the gap-table rows of d+1 are fetched from the symbols loaded one step earlier.
// A lane that is outside its diagonal keeps -inf, so a neighbour is valid exactly when its x lies in the G-wide
// window of the lane mapping of its diagonal, and the select is applied once per target state (all sources of a
// state come from the same neighbour cell).
template <int G, int THREADS, bool SWITCH>
__global__ void __launch_bounds__(THREADS, 5)
forward_tile_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                    int n_ids, const DevModel *__restrict__ models, const uint8_t *__restrict__ px,
                    const uint8_t *__restrict__ py, const int2 *__restrict__ geo, double *__restrict__ F,
                    double *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off) {
    __shared__ __align__(16) TileShared sh;
    tile_shared_init<THREADS>(sh, models);
    const int tid = tid_x_once();
    const int g = (blockIdx.x * THREADS + tid) / G;
    const bool real = g < n_ids;  // groups past the end shadow the last tile and store nothing
    const int lane = tid & (G - 1);
    const int prev_lane = mod_g(lane - 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const DevModel &m = sh.model[job.slot];
    const uint32_t lac_s = smem_u32(&sh.lac32[0][tid & 7]);
    const uint32_t gx_s = smem_u32(&sh.gx[job.slot][0][0]), gy_s = smem_u32(&sh.gy[job.slot][0][0]);
    const int N = job.lx + job.ly;
    const uint8_t *sx = px + job.symx_off - 1, *sy = py + job.symy_off - 1;  // sx[x] = symbol of column x (1-based)
    const int2 *gj = geo + job.diag_off;
    double *Fj = F + 3 * job.cell_base;
    double *bnd = Fbnd + 3 * slot_bnd_off[slot];
    const double *emP = m.emit_packed;
    const double tMM = m.t[T_MM], tXM = m.t[T_XM], tYM = m.t[T_YM];
    const int d0 = t.f0 + 1;  // first diagonal of the loop
    const int tb_to = t.tb_to;

    // ---- diagonal f0: the true start state (stateMachine.c:521-530) or a flat state for the warm-up
    int xlo1, xlo2 = -2 * G;  // nothing on "diagonal f0 - 1" is ever valid
    Cell f1 = cell_neg_inf();
    double midm = MB_NEG_INF, midx = MB_NEG_INF, midy = MB_NEG_INF;
    {
        const int2 ga = __ldg(gj + t.f0), gb = __ldg(gj + t.f0 + 1);
        xlo1 = ga.x;
        const int r = mod_g(lane - xlo1, G);
        if (r < gb.y - ga.y) {
            if (t.f0 == 0) {
                f1.m = job.ragged_l ? MB_NEG_INF : 0.0;
                f1.x = job.ragged_l ? 0.0 : MB_NEG_INF;
                f1.y = f1.x;
            } else {
                f1.m = f1.x = f1.y = 0.0;
            }
            if (real && t.f0 == tb_to) store_cell(bnd, r, f1);
        }
    }
    auto geo_at = [&](int dd) { return __ldg(gj + min(dd, N + 1)); };
    auto lane_x = [&](int xlo) { return xlo + mod_g(lane - xlo, G); };
    // ---- pipeline prologue
    int2 gr[4];
#pragma unroll
    for (int k = 0; k < 4; k++) gr[k] = geo_at(d0 + k);
    unsigned csx[2], csy[2];
    double eM[2];
    uint32_t exa[2], eya[2];
    {
        const int xa = lane_x(gr[0].x);
        const unsigned cx = __ldg(sx + xa), cy = __ldg(sy + (min(d0, N) - xa));
        eM[0] = __ldg(emP + cx * PACK_STRIDE + cy);
        exa[0] = gx_s + 32u * (unsigned) packed_base(cx);
        eya[0] = gy_s + 32u * (unsigned) packed_base(cy);
        const int xb = lane_x(gr[1].x);
        csx[1] = __ldg(sx + xb);
        csy[1] = __ldg(sy + (min(d0 + 1, N) - xb));
    }
    const int n_it = t.tb_from - t.f0;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
    for (int it0 = 0; it0 < n_max; it0 += 4) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int it = it0 + k;
            const int d = d0 + it;
            const bool live = real && it < n_it;
            // ---- loads for the diagonals ahead (independent of the recurrence)
            {
                const int x2 = lane_x(gr[(k + 2) & 3].x);
                csx[k & 1] = __ldg(sx + x2);
                csy[k & 1] = __ldg(sy + (min(d + 2, N) - x2));
                const unsigned cx = csx[(k + 1) & 1], cy = csy[(k + 1) & 1];
                eM[(k + 1) & 1] = __ldg(emP + cx * PACK_STRIDE + cy);
                exa[(k + 1) & 1] = gx_s + 32u * (unsigned) packed_base(cx);
                eya[(k + 1) & 1] = gy_s + 32u * (unsigned) packed_base(cy);
            }
            // ---- diagonal d
            const int xlo = gr[k].x;
            const uint32_t co = (uint32_t) gr[k].y;
            const int w = gr[(k + 1) & 3].y - gr[k].y;
            const int r = mod_g(lane - xlo, G);
            const int x = xlo + r;
            const bool active = r < w;
            const double lm = shfl_g<G>(f1.m, prev_lane);
            const double lx_ = shfl_g<G>(f1.x, prev_lane);
            const double ly_ = shfl_g<G>(f1.y, prev_lane);
            const int xe = active ? x : FAR_X;  // a lane outside its diagonal fails every window test
            const bool ok_lower = (unsigned) (xe - 1 - xlo1) < (unsigned) G;
            const bool ok_upper = (unsigned) (xe - xlo1) < (unsigned) G;
            const bool ok_middle = (unsigned) (xe - 1 - xlo2) < (unsigned) G;
            const double2 ex = lds_f64x2(exa[k & 1]), ey = lds_f64x2(eya[k & 1]);
            const double e = eM[k & 1];
            Cell c;
            // stateMachine3_cellCalculate (stateMachine.c:562-586): X from lower, M from middle, Y from upper
            c.x = logadd_t(dadd(lm, ex.x), dadd(lx_, ex.y), lac_s);
            if (SWITCH) c.x = logadd_t(c.x, dadd(ly_, lds_f64(exa[k & 1] + 16)), lac_s);
            c.m = logadd_t(dadd(midm, dadd(e, tMM)), dadd(midx, dadd(e, tXM)), lac_s);
            c.m = logadd_t(c.m, dadd(midy, dadd(e, tYM)), lac_s);
            c.y = logadd_t(dadd(f1.m, ey.x), dadd(f1.y, ey.y), lac_s);
            if (SWITCH) c.y = logadd_t(c.y, dadd(f1.x, lds_f64(eya[k & 1] + 16)), lac_s);
            c.x = ok_lower ? c.x : MB_NEG_INF;
            c.m = ok_middle ? c.m : MB_NEG_INF;
            c.y = ok_upper ? c.y : MB_NEG_INF;
            {
                // predicated stores (no divergent branch in the loop body); the one diagonal that goes to the boundary
                // buffer instead of F takes a warp-uniform branch
                double *q = Fj + 3 * ((int64_t) co + r);
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.global.f64 [%1], %2;\n\t"
                             "@p st.global.f64 [%1+8], %3;\n\t@p st.global.f64 [%1+16], %4;\n\t}"
                             :: "r"((unsigned) (live && active && d > tb_to)), "l"(q), "d"(c.m), "d"(c.x), "d"(c.y)
                             : "memory");
                if (__any_sync(0xffffffffu, live && d == tb_to)) {
                    if (live && active && d == tb_to) store_cell(bnd, r, c);
                }
            }
            // ---- carry
            midm = lm;
            midx = lx_;
            midy = ly_;
            f1 = c;
            xlo2 = xlo1;
            xlo1 = xlo;
            gr[k] = geo_at(d + 4);
        }
    }
}

// ----------------------------------------------------------------------------------------------- backward tile

// Sequential logAdd folds over the cells of two diagonals in ascending x (dpDiagonal_dotProduct,
// pairwiseAligner.c:450-461), interleaved: `va` / `vb` are this lane's values, diagonal a has wa <= G cells starting
// at x = xa, diagonal b has wb cells starting at xb.  Uniform trip count across the warp (w_max = widest diagonal
// among the warp's groups); a fold past its diagonal's width folds -inf, which leaves the sum unchanged.
template <int G>
__device__ __forceinline__ void fold_diag2(double va, int xa, int wa, double vb, int xb, int wb, int w_max,
                                           uint32_t lac_s, double &acc_a, double &acc_b) {
    acc_a = MB_NEG_INF;
    acc_b = MB_NEG_INF;
#pragma unroll 1
    for (int i = 0; i < w_max; i++) {
        const double a = shfl_g<G>(va, mod_g(xa + i, G));
        const double b = shfl_g<G>(vb, mod_g(xb + i, G));
        acc_a = logadd_t(acc_a, i < wa ? a : MB_NEG_INF, lac_s);
        acc_b = logadd_t(acc_b, i < wb ? b : MB_NEG_INF, lac_s);
    }
}

__device__ __forceinline__ double cell_dot_t(const Cell &a, const Cell &b, uint32_t lac_s) {
    double tp = dadd(a.m, b.m);
    tp = logadd_t(tp, dadd(a.x, b.x), lac_s);
    tp = logadd_t(tp, dadd(a.y, b.y), lac_s);
    return tp;
}

// ----------------------------------------------------------------------------------------- forward only, tiny
//
// computeForwardProbability (pairwiseAligner.c:849-903) for the matrices poa_polish scores alleles with (SURVEY A9:
// lX, lY median 11, max 43, no anchors, ~0.4 M calls per chunk): one group of G lanes per matrix, the whole sweep in
// registers, band geometry computed inline from the four corner points of band_construct without anchors
// (pairwiseAligner.c:161-226 with one "anchor pair" (0,0) -> (lX,lY)), nothing written but the final log-probability
// cell_dotProduct(F(lX,lY), end state) (:881-891).  Same lane mapping, validity rule and logAdd as the tile kernels.
template <int G, int THREADS, bool SWITCH>
__global__ void __launch_bounds__(THREADS)
tiny_forward_kernel(const Job *__restrict__ jobs, const int32_t *__restrict__ ids, int n_ids,
                    const DevModel *__restrict__ models, const uint8_t *__restrict__ px, const uint8_t *__restrict__ py,
                    int expansion, double *__restrict__ out) {
    __shared__ __align__(16) TileShared sh;
    tile_shared_init<THREADS>(sh, models);
    const int tid = tid_x_once();
    const int g = (blockIdx.x * THREADS + tid) / G;
    const bool real = g < n_ids;
    const int lane = tid & (G - 1);
    const int prev_lane = mod_g(lane - 1, G);
    const Job job = jobs[ids[min(g, n_ids - 1)]];
    const DevModel &m = sh.model[job.slot];
    const uint32_t lac_s = smem_u32(&sh.lac32[0][tid & 7]);
    const uint32_t gx_s = smem_u32(&sh.gx[job.slot][0][0]), gy_s = smem_u32(&sh.gy[job.slot][0][0]);
    const int lx = job.lx, ly = job.ly, N = lx + ly;
    const uint8_t *sx = px + job.symx_off - 1, *sy = py + job.symy_off - 1;  // sx[x] = symbol of column x (1-based)
    const double *emP = m.emit_packed;
    const double tMM = m.t[T_MM], tXM = m.t[T_XM], tYM = m.t[T_YM];
    // corner points of the band (band_diagonal with no anchors: previous point (0,0), next point (lX,lY))
    const int nxay = lx + ly, nxmy = lx - ly, e = expansion;
    const int xL = bound_coord((0 + 0 - e) / 2, lx), yL = bound_coord((nxay - (nxmy - e)) / 2, ly);
    const int xU = bound_coord((nxay + nxmy + e) / 2, lx), yU = bound_coord((0 - (0 + e)) / 2, ly);
    auto geo_of = [&](int d, int &xlo, int &w) {  // diagonal d >= 1 (clamped to N)
        int kl, kr;
        clip_diagonal(min(d, N), xL, yL, xU, yU, kl, kr);
        xlo = (min(d, N) + kl) >> 1;
        w = (kr - kl) / 2 + 1;
    };
    auto lane_x = [&](int xlo) { return xlo + mod_g(lane - xlo, G); };
    // diagonal 0: the start state (stateMachine.c:521-530)
    Cell f1 = cell_neg_inf(), res = cell_neg_inf();
    double midm = MB_NEG_INF, midx = MB_NEG_INF, midy = MB_NEG_INF;
    if (lane == 0) {
        f1.m = job.ragged_l ? MB_NEG_INF : 0.0;
        f1.x = job.ragged_l ? 0.0 : MB_NEG_INF;
        f1.y = f1.x;
    }
    int xlo1 = 0, xlo2 = -2 * G;
    // pipeline: geometry two diagonals ahead, symbols one ahead, emission for the current one fetched a step early
    int xloA, wA, xloB, wB;
    geo_of(1, xloA, wA);
    geo_of(2, xloB, wB);
    unsigned cxA, cyA, cxB, cyB;
    {
        const int xa = lane_x(xloA);
        cxA = __ldg(sx + xa);
        cyA = __ldg(sy + (min(1, N) - xa));
        const int xb = lane_x(xloB);
        cxB = __ldg(sx + xb);
        cyB = __ldg(sy + (min(2, N) - xb));
    }
    double eA = __ldg(emP + cxA * PACK_STRIDE + cyA);
    const int n_max = __reduce_max_sync(0xffffffffu, N);
    for (int d = 1; d <= n_max; d++) {
        // ---- loads for the diagonals ahead
        int xloC, wC;
        geo_of(d + 2, xloC, wC);
        const int xc = lane_x(xloC);
        const unsigned cxC = __ldg(sx + xc), cyC = __ldg(sy + (min(d + 2, N) - xc));
        const double eB = __ldg(emP + cxB * PACK_STRIDE + cyB);
        // ---- diagonal d
        const int r = mod_g(lane - xloA, G);
        const int x = xloA + r;
        const bool active = r < wA && d <= N;
        const double lm = shfl_g<G>(f1.m, prev_lane);
        const double lx_ = shfl_g<G>(f1.x, prev_lane);
        const double ly_ = shfl_g<G>(f1.y, prev_lane);
        const int xe = active ? x : FAR_X;
        const bool ok_lower = (unsigned) (xe - 1 - xlo1) < (unsigned) G;
        const bool ok_upper = (unsigned) (xe - xlo1) < (unsigned) G;
        const bool ok_middle = (unsigned) (xe - 1 - xlo2) < (unsigned) G;
        const uint32_t exa = gx_s + 32u * (unsigned) packed_base(cxA), eya = gy_s + 32u * (unsigned) packed_base(cyA);
        const double2 ex = lds_f64x2(exa), ey = lds_f64x2(eya);
        Cell c;
        c.x = logadd_t(dadd(lm, ex.x), dadd(lx_, ex.y), lac_s);
        if (SWITCH) c.x = logadd_t(c.x, dadd(ly_, lds_f64(exa + 16)), lac_s);
        c.m = logadd_t(dadd(midm, dadd(eA, tMM)), dadd(midx, dadd(eA, tXM)), lac_s);
        c.m = logadd_t(c.m, dadd(midy, dadd(eA, tYM)), lac_s);
        c.y = logadd_t(dadd(f1.m, ey.x), dadd(f1.y, ey.y), lac_s);
        if (SWITCH) c.y = logadd_t(c.y, dadd(f1.x, lds_f64(eya + 16)), lac_s);
        c.x = ok_lower ? c.x : MB_NEG_INF;
        c.m = ok_middle ? c.m : MB_NEG_INF;
        c.y = ok_upper ? c.y : MB_NEG_INF;
        if (d == N && active && x == lx) res = c;
        // ---- carry
        midm = lm;
        midx = lx_;
        midy = ly_;
        f1 = c;
        xlo2 = xlo1;
        xlo1 = xloA;
        xloA = xloB;
        wA = wB;
        xloB = xloC;
        wB = wC;
        cxA = cxB;
        cyA = cyB;
        cxB = cxC;
        cyB = cyC;
        eA = eB;
    }
    if (N == 0) res = f1;  // (0,0) is the last cell; the caller special-cases the empty problem below
    const int owner = N == 0 ? 0 : mod_g(lx, G);
    if (real && lane == owner) {
        const double *endp = m.end_p[job.ragged_r ? 1 : 0];
        Cell ec;
        ec.m = endp[0];
        ec.x = endp[1];
        ec.y = endp[2];
        out[job.aln] = N == 0 ? 0.0 : cell_dot_t(res, ec, lac_s);  // an empty problem returns LOG_ONE (:860-862)
    }
}

// Job table and work lists of an anchor-free forward-only batch, built on the device (margin_b200_forward_batch's fast
// path: poa_polish sends ~0.4-2 M tiny matrices per chunk, and a host loop over them — 80-byte jobs, a band chunk each —
// cost more than the DP).  One thread per alignment: writes its Job, and appends its index to the list of the kernel that
// takes it: lane groups of 16 (min(lX, lY) + 1 <= 16), of 32 (<= 32), or the general path (counts[2], ids_gen).
// counts: [0..2] list lengths, [3] bad model slot seen; cells[0] = cells of the register-resident matrices.
__global__ void forward_jobs_kernel(int64_t n, const int64_t *__restrict__ off_x, const int64_t *__restrict__ off_y,
                                    const uint8_t *__restrict__ slot, const uint8_t *__restrict__ ragged_l,
                                    const uint8_t *__restrict__ ragged_r, int have_slot0, int have_slot1,
                                    Job *__restrict__ jobs, int32_t *__restrict__ ids16, int32_t *__restrict__ ids32,
                                    int32_t *__restrict__ ids_gen, int32_t *__restrict__ counts,
                                    unsigned long long *__restrict__ cells) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    int cls = 3;  // no list
    unsigned long long mine = 0;
    if (i < n) {
        Job job;
        job.symx_off = off_x[i];
        job.symy_off = off_y[i];
        job.anc_off = 0;
        job.diag_off = 0;
        job.cell_base = 0;
        job.ring_off = 0;
        job.lx = (int32_t) (off_x[i + 1] - off_x[i]);
        job.ly = (int32_t) (off_y[i + 1] - off_y[i]);
        job.n_anc = 0;
        job.x1 = job.y1 = 0;
        job.aln = (int32_t) i;
        job.seg_base = job.max_seg = 0;
        job.ring_w = 0;
        job.slot = slot ? slot[i] : 0;
        job.ragged_l = ragged_l ? ragged_l[i] : 0;
        job.ragged_r = ragged_r ? ragged_r[i] : 0;
        job.pad = 0;
        if (job.slot > 1 || !(job.slot ? have_slot1 : have_slot0) || job.lx < 0 || job.ly < 0) atomicExch(counts + 3, 1);
        jobs[i] = job;
        const int m = min(job.lx, job.ly) + 1;
        cls = m <= 16 ? 0 : (m <= 32 ? 1 : 2);
        if (cls < 2) mine = (unsigned long long) (job.lx + 1) * (unsigned long long) (job.ly + 1);
    }
    const int lane = threadIdx.x & 31;
    int32_t *lists[3] = {ids16, ids32, ids_gen};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, cls == k);
        if (m == 0) continue;
        int base = 0;
        if (lane == __ffs(m) - 1) base = atomicAdd(counts + k, __popc(m));
        base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
        if (cls == k) lists[k][base + __popc(m & ((1u << lane) - 1u))] = (int32_t) i;
    }
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0 && mine) atomicAdd(cells, mine);
}

// Per-diagonal lane bookkeeping shared by the three staged lists: `bits` of a ballot are cut down to the group,
// rotated so that bit i is the cell x = xlo + i, and a passing cell is appended behind the passing cells below it.
template <int G>
struct StageLane {
    int shift, rot_r;
    unsigned below;  // mask of the cells with smaller x
    __device__ __forceinline__ StageLane(int xlo, int rank, int gshift)
        : shift(gshift), rot_r(mod_g(xlo, G)), below((1u << rank) - 1u) {}
    __device__ __forceinline__ void put(bool ok, double v, int x, int y, StagedPair *dst, int &cnt) const {
        unsigned bits = __ballot_sync(0xffffffffu, ok);
        unsigned rot;
        if constexpr (G == 32) {
            rot = __funnelshift_r(bits, bits, rot_r);
        } else {
            bits = (bits >> shift) & ((1u << G) - 1u);
            rot = (bits >> rot_r) | (bits << (G - rot_r));  // bits above G are masked by `below`
        }
        if (ok) {  // a branch, not a predicated store: most warps stage nothing on most diagonals and skip the address math
            StagedPair sp;
            sp.v = v;
            sp.x = x;
            sp.y = y;
            dst[cnt + __popc(rot & below)] = sp;
        }
        cnt += __popc(bits);
    }
};

// One group of G lanes per tile: the walk-back loop of getPosteriorProbsWithBanding (pairwiseAligner.c:753-818) from
// t.D down to t.tb_to+1, posteriors for diagonals <= t.tb_from.  The total probability is re-sampled every 10th
// posterior diagonal counted from tb_from; tiles are cut so that their tb_from is one of the reference's samples.
// A cell state is staged when its log posterior v >= v_thr, where v_thr is the smallest double whose exp() reaches
// the posterior threshold (found once per context), i.e. exactly when exp(v) >= threshold.
// Same register pipeline as the forward kernel, walking down, with the geometry of d+1 .. d-2 in a move chain.
// Lanes outside their diagonal keep -inf, so the three sources of a cell are valid exactly when their x lies in the
// G-wide window of the lane mapping of their diagonal.
#ifndef MB_BWD_MIN_BLOCKS
#define MB_BWD_MIN_BLOCKS 4
#endif
template <int G, int THREADS, bool INDELS, bool SWITCH, bool GUARD>
__global__ void __launch_bounds__(THREADS, MB_BWD_MIN_BLOCKS)
backward_tile_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                     int n_ids, const DevModel *__restrict__ models, const uint8_t *__restrict__ px,
                     const uint8_t *__restrict__ py, const int2 *__restrict__ geo, const double *__restrict__ F,
                     double *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off, double v_thr,
                     StagedPair *__restrict__ stage, int64_t stage_stride, int64_t *__restrict__ seg_counts,
                     int total_slots) {
    __shared__ __align__(16) TileShared sh;
    tile_shared_init<THREADS>(sh, models);
    const int tid = tid_x_once();
    const int g = (blockIdx.x * THREADS + tid) / G;
    const bool real = g < n_ids;
    const int lane = tid & (G - 1);
    const int next_lane = mod_g(lane + 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const DevModel &m = sh.model[job.slot];
    const uint32_t lac_s = smem_u32(&sh.lac32[0][tid & 7]);
    const uint32_t gx_s = smem_u32(&sh.gx[job.slot][0][0]), gy_s = smem_u32(&sh.gy[job.slot][0][0]);
    const uint8_t *sx = px + job.symx_off, *sy = py + job.symy_off;
    const int2 *gj = geo + job.diag_off;
    const double *Fj = F + 3 * job.cell_base;
    double *bnd = Fbnd + 3 * slot_bnd_off[slot];  // [F(tb_to) | B(tb_to + 1) | B(tb_from + 1)], see Tile
    const int bnd_cells = GUARD ? (int) ((slot_bnd_off[slot + 1] - slot_bnd_off[slot]) / 3) : 0;
    const double *emP = m.emit_packed;
    const double tMM = m.t[T_MM], tXM = m.t[T_XM], tYM = m.t[T_YM];
    const int D = t.D, tb_to = t.tb_to, tb_from = t.tb_from;
    const double *endp = m.end_p[((t.flags & TILE_AT_END) && job.ragged_r) ? 1 : 0];
    const int64_t region = job.cell_base + (uint32_t) __ldg(gj + tb_to + 1).y;
    StagedPair *stM = stage + region, *stX = stage + stage_stride + region, *stY = stage + 2 * stage_stride + region;
    int cntM = 0, cntX = 0, cntY = 0;
    const int xoff = job.x1 - 1, yoff = job.y1 - 1;

    auto geo_at = [&](int dd) { return __ldg(gj + max(dd, 0)); };
    auto lane_x = [&](int xlo) { return xlo + mod_g(lane - xlo, G); };

    Cell b1 = cell_neg_inf();
    double b2m = MB_NEG_INF;  // b(d+2).m of lane+1, i.e. the m shuffled one diagonal earlier
    int xlo2 = -2 * G, w1 = 0;
    double total = MB_NEG_INF;
    int until_total = 0;  // posterior diagonals until the next total-probability sample
    // ---- pipeline prologue: geometry of D+1 .. D-2 (+ D-3 in flight); symbols, emission and F of D, symbols of D-1
    int2 gUp = geo_at(D + 1), gA = geo_at(D), gB = geo_at(D - 1), gC = geo_at(D - 2);
    int2 gn[2];  // landing registers: a geometry word is loaded one step and moved into the chain the next
    gn[1] = geo_at(D - 3);
    unsigned csx[2], csy[2];
    double eM[2];
    uint32_t exa[2], eya[2];
    Cell fc[2];
    {
        const int xa = lane_x(gA.x);
        const unsigned cx = __ldg(sx + xa), cy = __ldg(sy + (D - xa));
        eM[0] = __ldg(emP + cx * PACK_STRIDE + cy);
        exa[0] = gx_s + 32u * (unsigned) packed_base(cx);
        eya[0] = gy_s + 32u * (unsigned) packed_base(cy);
        const int xb = lane_x(gB.x);
        csx[1] = __ldg(sx + xb);
        csy[1] = __ldg(sy + (max(D - 1, 0) - xb));
        fc[0] = cell_neg_inf();
        if (D <= tb_from && xa - gA.x < gUp.y - gA.y) fc[0] = load_cell_cg(Fj, (int64_t) (uint32_t) gA.y + (xa - gA.x));
    }
    const int n_it = D - tb_to;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
    // The loop body is two diagonals (the period of the symbol / emission / F rings); four would leave the geometry
    // chain move-free too but, with the total-probability block in every copy, overflows the instruction cache.
    for (int it0 = 0; it0 < n_max; it0 += 2) {
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int it = it0 + k;
            const bool live = real && it < n_it;
            const int d = D - min(it, n_it - 1);
            gn[k & 1] = geo_at(d - 4);
            // ---- loads for the diagonals below (independent of the recurrence)
            {
                const int x2 = lane_x(gC.x);
                csx[k & 1] = __ldg(sx + x2);
                csy[k & 1] = __ldg(sy + (max(d - 2, 0) - x2));
                const unsigned cx = csx[(k + 1) & 1], cy = csy[(k + 1) & 1];
                eM[(k + 1) & 1] = __ldg(emP + cx * PACK_STRIDE + cy);
                exa[(k + 1) & 1] = gx_s + 32u * (unsigned) packed_base(cx);
                eya[(k + 1) & 1] = gy_s + 32u * (unsigned) packed_base(cy);
                // F of the next diagonal down (used one iteration later)
                const int rn = mod_g(lane - gB.x, G);
                Cell fn = cell_neg_inf();
                if (d - 1 <= tb_from && d - 1 > tb_to && rn < gA.y - gB.y) fn = load_cell_cg(Fj, (int64_t) (uint32_t) gB.y + rn);
                fc[(k + 1) & 1] = fn;
            }
            const int xlo = gA.x, w = gUp.y - gA.y, xlo1 = gUp.x;
            const int rank = mod_g(lane - xlo, G);
            const int x = xlo + rank;
            const int y = d - x;
            const bool active = rank < w;
            const bool post = live && d <= tb_from;
            const bool do_total = post && until_total == 0;
            const Cell f0c = fc[k & 1];
            Cell b;
            {
                // gather form of diagonalCalculationBackward: match term from (x+1,y+1) first, then the gap-Y term from
                // (x,y+1), then the gap-X term from (x+1,y) (see backward_cell in pairhmm_kernels.cuh)
                const double nx = shfl_g<G>(b1.x, next_lane);
                const double nm = shfl_g<G>(b1.m, next_lane);
                const int xe = active ? x : FAR_X;  // a lane outside its diagonal fails every window test
                const bool has_mm = (unsigned) (xe + 1 - xlo2) < (unsigned) G;
                const bool has_y = (unsigned) (xe - xlo1) < (unsigned) G;
                const bool has_x = (unsigned) (xe + 1 - xlo1) < (unsigned) G;
                const double vm = has_mm ? b2m : MB_NEG_INF;
                const double vy = has_y ? b1.y : MB_NEG_INF;
                const double vx = has_x ? nx : MB_NEG_INF;
                const double2 ex = lds_f64x2(exa[k & 1]);  // {M->X, X->X}, +16: Y->X
                const double2 ey = lds_f64x2(eya[k & 1]);  // {M->Y, Y->Y}, +16: X->Y
                const double e = eM[k & 1];
                b.m = dadd(vm, dadd(e, tMM));
                b.x = dadd(vm, dadd(e, tXM));
                b.y = dadd(vm, dadd(e, tYM));
                b.m = logadd_t(b.m, dadd(vy, ey.x), lac_s);
                if (SWITCH) b.x = logadd_t(b.x, dadd(vy, lds_f64(eya[k & 1] + 16)), lac_s);
                b.y = logadd_t(b.y, dadd(vy, ey.y), lac_s);
                b.m = logadd_t(b.m, dadd(vx, ex.x), lac_s);
                b.x = logadd_t(b.x, dadd(vx, ex.y), lac_s);
                if (SWITCH) b.y = logadd_t(b.y, dadd(vx, lds_f64(exa[k & 1] + 16)), lac_s);
                b2m = nm;
                if (k == 0 && it0 == 0) {  // d == D: the backward sweep starts from the end state on every cell (:754-762)
                    b.m = active ? endp[0] : MB_NEG_INF;
                    b.x = active ? endp[1] : MB_NEG_INF;
                    b.y = active ? endp[2] : MB_NEG_INF;
                }
            }
            if (GUARD) {
                // seam guard: the last diagonal of this sweep and the warm-up diagonal just above the tile's own range
                // are kept, so that guard_check_kernel can compare them with the neighbouring tiles' values.  Two
                // diagonals per tile: a warp-uniform branch keeps the stores out of the other iterations.
                const bool low = d == tb_to + 1;
                const bool keep = live && (low || (d == tb_from + 1 && it > 0));
                if (__any_sync(0xffffffffu, keep)) {
                    if (keep && active) store_cell(bnd, (int64_t) (low ? BND_B_LOW : BND_B_HIGH) * bnd_cells + rank, b);
                }
            }
            if (MB_TOTAL_ON && __any_sync(0xffffffffu, do_total)) {
                // diagonalCalculationTotalProbability (pairwiseAligner.c:578-596); groups that do not sample on this
                // iteration run along (uniform control flow) and discard the result
                const int w_max = __reduce_max_sync(0xffffffffu, max(w, w1));
                const double dot = active ? cell_dot_t(f0c, b, lac_s) : MB_NEG_INF;
                // paths that jump over d with a match: match block of the forward step from F(d-1) into the cells
                // of d+1, dotted with B(d+1) (this lane's b1 is its cell of d+1)
                const int xlom = gB.x, wm = gA.y - gB.y;
                const int x1 = lane_x(xlo1);
                const int y1 = d + 1 - x1;
                double val = MB_NEG_INF;
                if (do_total && it > 0 && x1 < xlo1 + w1) {
                    double gm_ = MB_NEG_INF;
                    if ((unsigned) (x1 - 1 - xlom) < (unsigned) wm) {
                        const Cell f = (d - 1 == tb_to && (t.flags & TILE_BND))
                                           ? load_cell_cg(bnd, x1 - 1 - xlom)
                                           : load_cell_cg(Fj, (int64_t) (uint32_t) gB.y + (x1 - 1 - xlom));
                        const double e = __ldg(emP + (unsigned) __ldg(sx + x1 - 1) * PACK_STRIDE + (unsigned) __ldg(sy + y1 - 1));
                        gm_ = dadd(f.m, dadd(e, tMM));
                        gm_ = logadd_t(gm_, dadd(f.x, dadd(e, tXM)), lac_s);
                        gm_ = logadd_t(gm_, dadd(f.y, dadd(e, tYM)), lac_s);
                    }
                    val = dadd(gm_, b1.m);
                }
                double tot, second;
                fold_diag2<G>(dot, xlo, w, val, xlo1, w1, w_max, lac_s, tot, second);
                if (it > 0) tot = logadd_t(tot, second, lac_s);
                if (do_total) total = tot;
            }
            {
                // posteriors of diagonal d (diagonalCalculationPosteriorProbs, :637-681) in the log domain; the ballots
                // are full-warp, so this block is executed by every group on every iteration
                const StageLane<G> sl(xlo, rank, (tid & 31) & ~(G - 1));
                const bool pa = post && active;
                const double vM = dadd(dadd(f0c.m, b.m), -total);
                sl.put(pa && x > 0 && y > 0 && vM >= v_thr, vM, x + xoff, y + yoff, stM, cntM);
                if (INDELS) {
                    const double vX = dadd(dadd(f0c.x, b.x), -total);
                    sl.put(pa && x > 0 && vX >= v_thr, vX, x + xoff, y + yoff, stX, cntX);
                    const double vY = dadd(dadd(f0c.y, b.y), -total);
                    sl.put(pa && y > 0 && vY >= v_thr, vY, x + xoff, y + yoff, stY, cntY);
                }
                if (post) until_total = until_total == 0 ? 9 : until_total - 1;
            }
            // ---- carry
            b1 = b;
            xlo2 = xlo1;
            w1 = w;
            gUp = gA;
            gA = gB;
            gB = gC;
            gC = gn[(k + 1) & 1];
        }
    }
    if (real && lane == 0) {
        seg_counts[slot] = cntM;
        seg_counts[total_slots + slot] = cntX;
        seg_counts[2 * total_slots + slot] = cntY;
    }
}

// ------------------------------------------------------------------------------------------------------ gather

// Moves the staged log posteriors of every tile into the final lists as (weight, x, y): tiles are already in final
// order, inside a tile the reference's pop loop reverses the emission order (pairwiseAligner.c:1104-1123), and the
// weight is floor(min(exp(v), 1) * 1e7) (:598-608).
__global__ void gather_staged_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, int total_slots,
                                     const int2 *__restrict__ geo, const int64_t *__restrict__ seg_counts,
                                     const int64_t *__restrict__ seg_off, const StagedPair *__restrict__ stage,
                                     int64_t stage_stride, void *__restrict__ out, int packed, int n_lists,
                                     int64_t out_capacity) {
    const int slot = blockIdx.x;
    if (seg_off[(int64_t) 3 * total_slots] > out_capacity) return;  // the host re-runs with a larger buffer
    const Tile t = tiles[slot];
    if (!(t.flags & TILE_BND)) return;  // sequential-mode jobs are gathered by gather_kernel
    const Job job = jobs[t.job];
    const int64_t region = job.cell_base + (uint32_t) geo[job.diag_off + t.tb_to + 1].y;
    for (int l = 0; l < n_lists; l++) {
        const int64_t c = seg_counts[(int64_t) l * total_slots + slot];
        const int64_t o = seg_off[(int64_t) l * total_slots + slot];
        const StagedPair *src = stage + (int64_t) l * stage_stride + region;
        for (int64_t i = threadIdx.x; i < c; i += blockDim.x) {
            const StagedPair sp = src[c - 1 - i];
            double pr = exp(sp.v);
            if (pr > 1.0) pr = 1.0;
            Pair q;
            q.w = (int32_t) floor(dmul(pr, 10000000.0));
            q.x = sp.x;
            q.y = sp.y;
            store_pair(out, o + i, packed, q);
        }
    }
}

// --------------------------------------------------------------------------------- compact tuples (DELTA32)
//
// MARGIN_B200_PAIRS_DELTA32 (include/margin_b200.h): one 32-bit record per tuple, weight in the low 24 bits and in the high
// 8 bits the step from the previous tuple of the same list — (anti-diagonal step 0..15) << 4 | (x step + 8) — or 0xFF,
// which takes the tuple's coordinates from the list's escape array instead.  In list order consecutive tuples sit on the
// same or the next few anti-diagonals, so almost every match tuple and most gap tuples fit a step; the first tuple of
// every tile (lists are assembled tile by tile) is always an escape.  Half the bytes of PACKED64 over the host link,
// which is what bounds the end-to-end rate when eight GPUs share one host (measured 15 GB/s per GPU device->host).
struct TupleXY {
    int32_t x, y;
};

__device__ __forceinline__ bool delta32_code(int px, int py, int x, int y, unsigned &code) {
    const int dd = (x + y) - (px + py), dx = x - px;
    code = ((unsigned) dd << 4) | (unsigned) (dx + 8);
    return (unsigned) dd <= 15u && (unsigned) (dx + 8) <= 15u && code != 0xFFu;
}

// Coordinates of entry i (final order) of a tile's list: tiles of the lane-group kernels stage StagedPair records, the
// general kernels Pair records; both in emission order, i.e. reversed.
struct DeltaSrc {
    const StagedPair *sp;
    const Pair *pr;
    int64_t c;
    __device__ __forceinline__ void xy(int64_t i, int &x, int &y) const {
        if (sp) {
            x = sp[c - 1 - i].x;
            y = sp[c - 1 - i].y;
        } else {
            x = pr[c - 1 - i].x;
            y = pr[c - 1 - i].y;
        }
    }
    __device__ __forceinline__ int32_t weight(int64_t i) const {
        if (!sp) return pr[c - 1 - i].w;
        double p = exp(sp[c - 1 - i].v);
        if (p > 1.0) p = 1.0;
        return (int32_t) floor(dmul(p, 10000000.0));
    }
};

__device__ __forceinline__ DeltaSrc delta_src(const Job *jobs, const Tile &t, int slot, int l, int total_slots,
                                              const int2 *geo, const int64_t *seg_counts, const StagedPair *stage2,
                                              int64_t stride2, const Pair *stage1, int64_t stride1) {
    DeltaSrc s;
    s.c = seg_counts[(int64_t) l * total_slots + slot];
    if (t.flags & TILE_BND) {
        const Job job = jobs[t.job];
        s.sp = stage2 + (int64_t) l * stride2 + job.cell_base + (uint32_t) geo[job.diag_off + t.tb_to + 1].y;
        s.pr = nullptr;
    } else {
        s.sp = nullptr;
        s.pr = stage1 + (int64_t) l * stride1 + t.stage_off;
    }
    return s;
}

// escapes per [list][slot]: one CTA per tile
__global__ void delta32_count_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, int total_slots,
                                     const int2 *__restrict__ geo, const int64_t *__restrict__ seg_counts,
                                     const StagedPair *__restrict__ stage2, int64_t stride2, const Pair *__restrict__ stage1,
                                     int64_t stride1, int n_lists, int64_t *__restrict__ esc_counts) {
    const int slot = blockIdx.x;
    const Tile t = tiles[slot];
    __shared__ int total;
    for (int l = 0; l < n_lists; l++) {
        const DeltaSrc src = delta_src(jobs, t, slot, l, total_slots, geo, seg_counts, stage2, stride2, stage1, stride1);
        if (threadIdx.x == 0) total = 0;
        __syncthreads();
        int mine = 0;
        for (int64_t i = threadIdx.x; i < src.c; i += blockDim.x) {
            bool esc = i == 0;
            if (!esc) {
                int px, py, x, y;
                unsigned code;
                src.xy(i - 1, px, py);
                src.xy(i, x, y);
                esc = !delta32_code(px, py, x, y, code);
            }
            mine += esc ? 1 : 0;
        }
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&total, mine);
        __syncthreads();
        if (threadIdx.x == 0) esc_counts[(int64_t) l * total_slots + slot] = total;
        __syncthreads();
    }
}

// records and escapes of every tile at their final places; blockDim.x must be 128
__global__ void delta32_gather_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, int total_slots,
                                      const int2 *__restrict__ geo, const int64_t *__restrict__ seg_counts,
                                      const int64_t *__restrict__ seg_off, const int64_t *__restrict__ esc_off,
                                      const StagedPair *__restrict__ stage2, int64_t stride2,
                                      const Pair *__restrict__ stage1, int64_t stride1, uint32_t *__restrict__ out,
                                      TupleXY *__restrict__ esc_out, int n_lists, int64_t out_capacity,
                                      int64_t esc_capacity) {
    const int slot = blockIdx.x;
    if (seg_off[(int64_t) 3 * total_slots] > out_capacity || esc_off[(int64_t) 3 * total_slots] > esc_capacity) return;
    const Tile t = tiles[slot];
    __shared__ int warp_tot[4];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int l = 0; l < n_lists; l++) {
        const DeltaSrc src = delta_src(jobs, t, slot, l, total_slots, geo, seg_counts, stage2, stride2, stage1, stride1);
        const int64_t o = seg_off[(int64_t) l * total_slots + slot];
        int64_t eo = esc_off[(int64_t) l * total_slots + slot];
        for (int64_t base = 0; base < src.c; base += blockDim.x) {
            const int64_t i = base + threadIdx.x;
            bool esc = false;
            unsigned code = 0xFFu;
            int x = 0, y = 0;
            if (i < src.c) {
                src.xy(i, x, y);
                esc = i == 0;
                if (!esc) {
                    int px, py;
                    src.xy(i - 1, px, py);
                    esc = !delta32_code(px, py, x, y, code);
                }
                if (esc) code = 0xFFu;
                out[o + i] = ((uint32_t) src.weight(i) & 0xffffffu) | (code << 24);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, esc);
            if (lane == 0) warp_tot[wid] = __popc(bal);
            __syncthreads();
            int before = __popc(bal & ((1u << lane) - 1u));
            for (int w = 0; w < wid; w++) before += warp_tot[w];
            if (esc) {
                TupleXY e;
                e.x = x;
                e.y = y;
                esc_out[eo + before] = e;
            }
            eo += warp_tot[0] + warp_tot[1] + warp_tot[2] + warp_tot[3];
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------- seam guard
//
// The tiling rests on both recurrences forgetting their start within the warm-up.  That is a property of the data
// (anchor density, band width), so it is CHECKED for every tile edge of every call: at a seam two tiles have
// computed the same diagonal independently — the lower tile's own F(d) against the upper tile's warmed-up F(d), and
// the upper tile's own B(d) against the lower tile's warmed-up B(d).  Each tile carries its own constant shift, so
// the comparison is made after aligning the two copies on the largest value of the converged one; cells that are more
// than `floor_gap` below it in both copies cannot move a posterior and are ignored.  A seam whose copies differ by
// more than `tol` is flagged; the host then re-runs the batch with the affected jobs swept sequentially (exactly the
// reference's recurrence), so a slow-forgetting stretch costs time, never correctness.
// One warp per seam.
__global__ void guard_check_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles,
                                   const Seam *__restrict__ seams, int n_seams, const int2 *__restrict__ geo,
                                   const double *__restrict__ F, const float *__restrict__ F32,
                                   const double *__restrict__ Fpriv, const double *__restrict__ bnd,
                                   const float *__restrict__ bnd32, const int64_t *__restrict__ slot_bnd_off, int f32,
                                   double tol, double tol32,
                                   double floor_gap, int32_t *__restrict__ flags, int32_t *__restrict__ n_failed,
                                   float *__restrict__ worst_out /* may be null: largest difference per seam */) {
    const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (k >= n_seams) return;
    const Seam sm = seams[k];
    const Tile lo = tiles[sm.lower], up = tiles[sm.upper];
    const Job job = jobs[lo.job];
    const int2 *gj = geo + job.diag_off;
    const int64_t cod = (uint32_t) gj[sm.d].y;
    const int w = (int) ((uint32_t) gj[sm.d + 1].y - (uint32_t) gj[sm.d].y);
    // A copy of a diagonal: FP64 cells of three doubles (general kernels, FP64 tile kernels) or, with the FP32 tile
    // kernels (f32), float4 cells for everything a lane-group tile stores.
    struct Side {
        const void *p;
        bool is_float;
        __device__ double at(int i) const {
            return is_float ? (double) reinterpret_cast<const float *>(p)[4 * (i / 3) + i % 3] : reinterpret_cast<const double *>(p)[i];
        }
    };
    auto sect = [&](int slot, int which) {
        const int64_t cells = (slot_bnd_off[slot + 1] - slot_bnd_off[slot]) / 3;
        const int64_t cell = slot_bnd_off[slot] + which * cells;
        const bool fl = f32 && (tiles[slot].flags & TILE_BND);
        return Side{fl ? (const void *) (bnd32 + 4 * cell) : (const void *) (bnd + 3 * cell), fl};
    };
    auto priv_at = [&](const Tile &t) {
        return Side{Fpriv + 3 * (t.priv_off - (int64_t) (uint32_t) gj[t.f0].y + cod), false};
    };
    auto shared_at = [&]() {
        return f32 ? Side{F32 + 4 * (job.cell_base + cod), true} : Side{F + 3 * (job.cell_base + cod), false};
    };
    Side conv, warm;  // the copy at the end of a long sweep / the copy at the end of a warm-up
    if (!sm.backward) {
        conv = (lo.flags & TILE_WIDE) ? priv_at(lo) : shared_at();
        warm = (up.flags & TILE_WIDE) ? priv_at(up) : sect(sm.upper, BND_F);
    } else {
        conv = sect(sm.upper, BND_B_LOW);
        warm = sect(sm.lower, BND_B_HIGH);
    }
    if (conv.is_float || warm.is_float) tol = tol32;
    // largest value of the converged copy and where it is
    double best = MB_NEG_INF;
    int at = -1;
    for (int i = lane; i < 3 * w; i += 32) {
        const double v = conv.at(i);
        if (v > best) {
            best = v;
            at = i;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oa = __shfl_xor_sync(0xffffffffu, at, o);
        if (ob > best || (ob == best && oa >= 0 && (at < 0 || oa < at))) {
            best = ob;
            at = oa;
        }
    }
    bool bad = false;
    double worst = 0.0;
    if (at < 0) {  // nothing finite: the other copy must be empty as well
        for (int i = lane; i < 3 * w; i += 32) bad |= warm.at(i) != MB_NEG_INF;
    } else {
        const double shift = warm.at(at) - best;
        if (!(fabs(shift) < 1e300)) bad = true;
        for (int i = lane; i < 3 * w; i += 32) {
            const double a = conv.at(i), wv = warm.at(i), b = wv - shift;
            if (a == MB_NEG_INF && wv == MB_NEG_INF) continue;
            if (a < best - floor_gap && b < best - floor_gap) continue;
            const double diff = fabs(a - b);
            if (!(diff <= tol)) bad = true;
            worst = (diff == diff) ? fmax(worst, diff) : 1e300;
        }
    }
    bad = __any_sync(0xffffffffu, bad);
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if (lane == 0) {
        flags[k] = bad ? 1 : 0;
        if (bad) atomicAdd(n_failed, 1);
        if (worst_out) worst_out[k] = bad && worst <= tol ? 3e38f : (float) fmin(worst, 3e38);
    }
}

// Band cells per execution path (result counters): out[0] = cells owned by TILE_WIDE tiles, out[1] = cells of tiles
// swept sequentially (neither TILE_BND nor TILE_WIDE).
__global__ void class_cells_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, int n_tiles,
                                   const int2 *__restrict__ geo, unsigned long long *__restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_tiles) return;
    const Tile t = tiles[s];
    if (t.flags & TILE_BND) return;
    const int2 *gj = geo + jobs[t.job].diag_off;
    const unsigned long long own = (uint32_t) gj[t.tb_from + 1].y - (uint32_t) gj[t.tb_to + 1].y;
    atomicAdd(out + ((t.flags & TILE_WIDE) ? 0 : 1), own);
}

// exp(v) >= threshold  <=>  v >= v_thr for the device's exp(): scans the doubles around log(threshold) and returns
// the smallest one whose exp() reaches the threshold (exp is monotone there; the window is +-256 ulps).
__global__ void find_log_threshold_kernel(double threshold, double guess, double *out) {
    long long bits = __double_as_longlong(guess);  // guess < 0: larger bit pattern = more negative
    long long best = bits - 257;
    for (long long k = -256; k <= 256; k++) {
        const double v = __longlong_as_double(bits - k);  // ascending v as k grows
        if (exp(v) >= threshold) {
            best = bits - k;
            break;
        }
    }
    *out = __longlong_as_double(best);
}

}  // namespace mb
