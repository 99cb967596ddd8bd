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

namespace mb {

constexpr int PACK_STRIDE = 5 * MARGIN_B200_MAX_RL;  // 255 packed symbol values
constexpr int PACK_SLACK = 64;                       // zeroed bytes in front of and behind the packed arrays

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

__device__ __forceinline__ Cell cell_sel(bool p, const Cell &c) {
    Cell r;
    r.m = p ? c.m : MB_NEG_INF;
    r.x = p ? c.x : MB_NEG_INF;
    r.y = p ? c.y : MB_NEG_INF;
    return r;
}

template <int G>
__device__ __forceinline__ double shfl_g(double v, int src) { return __shfl_sync(0xffffffffu, v, src, G); }

struct TileShared {
    DevModel model[2];
    double lac[16];
};

template <int THREADS>
__device__ __forceinline__ void tile_shared_init(TileShared &s, const DevModel *__restrict__ models) {
    for (int i = threadIdx.x; i < (int) (2 * sizeof(DevModel) / 8); i += THREADS)
        reinterpret_cast<double *>(s.model)[i] = reinterpret_cast<const double *>(models)[i];
    if (threadIdx.x < 16) s.lac[threadIdx.x] = LA_COEF[threadIdx.x];
    __syncthreads();
}

// A staged posterior: the log posterior v = F + B - total of a cell state that passes the threshold; the gather
// kernel turns it into the reference's weight floor(min(exp(v), 1) * 1e7) (pairwiseAligner.c:598-608).
struct StagedPair {
    double v;
    int32_t x, y;
};

// ------------------------------------------------------------------------------------------------ forward tile
template <int G, int THREADS, bool SWITCH>
__global__ void __launch_bounds__(THREADS)
forward_tile_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                    int n_ids, const DevModel *__restrict__ models, const uint8_t *__restrict__ px,
                    const uint8_t *__restrict__ py, const int2 *__restrict__ geo, double *__restrict__ F,
                    double *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off) {
    __shared__ __align__(16) TileShared sh;
    tile_shared_init<THREADS>(sh, models);
    const int g = (blockIdx.x * THREADS + threadIdx.x) / G;
    const bool real = g < n_ids;  // groups past the end shadow the last tile and store nothing
    const int lane = threadIdx.x & (G - 1);
    const int prev_lane = mod_g(lane - 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const DevModel &m = sh.model[job.slot];
    const double *lac = sh.lac;
    const int N = job.lx + job.ly;
    const uint8_t *sx = px + job.symx_off, *sy = py + job.symy_off;
    const int2 *gj = geo + job.diag_off;
    double *Fj = F + 3 * job.cell_base;
    double *bnd = Fbnd + 3 * slot_bnd_off[slot];
    const double *emP = m.emit_packed;
    const double tMM = m.t[T_MM], tXM = m.t[T_XM], tYM = m.t[T_YM];

    // ---- diagonal f0: the true start state (stateMachine.c:521-530) or a flat state for the warm-up
    int d = t.f0;
    const int2 g0 = __ldg(gj + d), g1 = __ldg(gj + d + 1);
    int xlo1 = g0.x, w1 = g1.y - g0.y, xlo2 = 0, w2 = 0;
    Cell f1 = cell_neg_inf(), mid = cell_neg_inf();
    {
        const int x = xlo1 + mod_g(lane - xlo1, G);
        if (x < xlo1 + w1) {
            if (d == 0) {
                f1.m = job.ragged_l ? MB_NEG_INF : 0.0;
                f1.x = job.ragged_l ? 0.0 : MB_NEG_INF;
                f1.y = f1.x;
            } else {
                f1.m = f1.x = f1.y = 0.0;
            }
            if (real && d == t.tb_to) store_cell(bnd, x - xlo1, f1);
        }
    }
    // ---- software pipeline prologue: geometry of d+1..d+3, symbols of d+1, d+2, emission of d+1
    auto geo_at = [&](int dd) { return __ldg(gj + min(dd, N + 1)); };
    auto sym_x_at = [&](int2 gg) { return (unsigned) __ldg(sx + (gg.x + mod_g(lane - gg.x, G) - 1)); };
    auto sym_y_at = [&](int2 gg, int dd) { return (unsigned) __ldg(sy + (dd - (gg.x + mod_g(lane - gg.x, G)) - 1)); };
    int2 gA = g1, gB = geo_at(d + 2), gC = geo_at(d + 3);  // geometry of d+1, d+2, d+3
    unsigned cx0 = sym_x_at(gA), cy0 = sym_y_at(gA, d + 1);
    unsigned cx1 = sym_x_at(gB), cy1 = sym_y_at(gB, min(d + 2, N));
    double eM0 = __ldg(emP + cx0 * PACK_STRIDE + cy0);

    const int n_it = t.tb_from - t.f0;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
#pragma unroll 2
    for (int it = 0; it < n_max; it++) {
        const bool live = real && it < n_it;
        d = t.f0 + 1 + min(it, n_it - 1);
        // stage loads for the diagonals ahead (independent of the recurrence)
        const int2 gD = geo_at(d + 3);
        const unsigned cx2 = sym_x_at(gC), cy2 = sym_y_at(gC, min(d + 2, N));
        const double eM1 = __ldg(emP + cx1 * PACK_STRIDE + cy1);
        // ---- diagonal d
        const int xlo = gA.x, w = gB.y - gA.y;
        const uint32_t co = (uint32_t) gA.y;
        const int x = xlo + mod_g(lane - xlo, G);
        const bool active = x < xlo + w;
        Cell lower_raw;
        lower_raw.m = shfl_g<G>(f1.m, prev_lane);
        lower_raw.x = shfl_g<G>(f1.x, prev_lane);
        lower_raw.y = shfl_g<G>(f1.y, prev_lane);
        const Cell lower = cell_sel((unsigned) (x - 1 - xlo1) < (unsigned) w1, lower_raw);
        const Cell upper = cell_sel((unsigned) (x - xlo1) < (unsigned) w1, f1);
        const Cell middle = cell_sel((unsigned) (x - 1 - xlo2) < (unsigned) w2, mid);
        const double *ex = m.ex_t[packed_base(cx0)];
        const double *ey = m.ey_t[packed_base(cy0)];
        Cell c;
        // stateMachine3_cellCalculate (stateMachine.c:562-586): X from lower, M from middle, Y from upper
        c.x = dadd(lower.m, ex[0]);
        c.x = logadd_tab(c.x, dadd(lower.x, ex[1]), lac);
        if (SWITCH) c.x = logadd_tab(c.x, dadd(lower.y, ex[2]), lac);
        c.m = dadd(middle.m, dadd(eM0, tMM));
        c.m = logadd_tab(c.m, dadd(middle.x, dadd(eM0, tXM)), lac);
        c.m = logadd_tab(c.m, dadd(middle.y, dadd(eM0, tYM)), lac);
        c.y = dadd(upper.m, ey[0]);
        c.y = logadd_tab(c.y, dadd(upper.y, ey[1]), lac);
        if (SWITCH) c.y = logadd_tab(c.y, dadd(upper.x, ey[2]), lac);
        c = cell_sel(active, c);
        if (live && active) {
            if (d > t.tb_to) store_cell(Fj, (int64_t) co + (x - xlo), c);
            else if (d == t.tb_to) store_cell(bnd, x - xlo, c);
        }
        // rotate
        mid = lower_raw;
        xlo2 = xlo1;
        w2 = w1;
        f1 = c;
        xlo1 = xlo;
        w1 = w;
        gA = gB;
        gB = gC;
        gC = gD;
        cx0 = cx1;
        cy0 = cy1;
        cx1 = cx2;
        cy1 = cy2;
        eM0 = eM1;
    }
}

// ----------------------------------------------------------------------------------------------- backward tile

// Sequential logAdd fold over the cells of one diagonal in ascending x (dpDiagonal_dotProduct, pairwiseAligner.c:450-461);
// `val` is this lane's value, the diagonal has w <= G cells starting at x = xlo.  Uniform trip count across the warp
// (w_max = widest diagonal among the warp's groups); groups with fewer cells fold -inf, which leaves acc unchanged.
template <int G>
__device__ __forceinline__ double fold_diag(double val, int xlo, int w, int w_max, const double *lac) {
    double acc = MB_NEG_INF;
    for (int i = 0; i < w_max; i++) {
        const double v = shfl_g<G>(val, mod_g(xlo + i, G));
        acc = logadd_tab(acc, i < w ? v : MB_NEG_INF, lac);
    }
    return acc;
}

__device__ __forceinline__ double cell_dot_tab(const Cell &a, const Cell &b, const double *lac) {
    double tp = dadd(a.m, b.m);
    tp = logadd_tab(tp, dadd(a.x, b.x), lac);
    tp = logadd_tab(tp, dadd(a.y, b.y), lac);
    return tp;
}

// Appends the passing cells of one diagonal to a staged list in ascending x (the order the reference emits them).
template <int G>
__device__ __forceinline__ void stage_list(bool ok, double v, int x, int y, int rank, int xlo, StagedPair *dst, int &cnt) {
    unsigned bits = __ballot_sync(0xffffffffu, ok);
    const int r = mod_g(xlo, G);
    unsigned rot;
    if constexpr (G == 32) {
        rot = __funnelshift_r(bits, bits, r);
    } else {
        bits = (bits >> ((threadIdx.x & 31) / G * G)) & ((1u << G) - 1u);
        rot = ((bits >> r) | (bits << (G - r))) & ((1u << G) - 1u);
    }
    if (ok) {
        StagedPair sp;
        sp.v = v;
        sp.x = x;
        sp.y = y;
        dst[cnt + __popc(rot & ((1u << rank) - 1u))] = sp;
    }
    cnt += __popc(bits);
}

// One group of G lanes per tile: the walk-back loop of getPosteriorProbsWithBanding (pairwiseAligner.c:753-818) from
// t.D down to t.tb_to+1, posteriors for diagonals <= t.tb_from.  The total probability is re-sampled every 10th
// posterior diagonal counted from tb_from; tiles are cut so that their tb_from is one of the reference's samples.
// A cell state is staged when its log posterior v >= v_thr, where v_thr is the smallest double whose exp() reaches
// the posterior threshold (found once per context), i.e. exactly when exp(v) >= threshold.
template <int G, int THREADS, bool INDELS, bool SWITCH>
__global__ void __launch_bounds__(THREADS)
backward_tile_kernel(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                     int n_ids, const DevModel *__restrict__ models, const uint8_t *__restrict__ px,
                     const uint8_t *__restrict__ py, const int2 *__restrict__ geo, const double *__restrict__ F,
                     const double *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off, double v_thr,
                     StagedPair *__restrict__ stage, int64_t stage_stride, int64_t *__restrict__ seg_counts,
                     int total_slots) {
    __shared__ __align__(16) TileShared sh;
    tile_shared_init<THREADS>(sh, models);
    const int g = (blockIdx.x * THREADS + threadIdx.x) / G;
    const bool real = g < n_ids;
    const int lane = threadIdx.x & (G - 1);
    const int next_lane = mod_g(lane + 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const DevModel &m = sh.model[job.slot];
    const double *lac = sh.lac;
    const uint8_t *sx = px + job.symx_off, *sy = py + job.symy_off;
    const int2 *gj = geo + job.diag_off;
    const double *Fj = F + 3 * job.cell_base;
    const double *bnd = Fbnd + 3 * slot_bnd_off[slot];
    const double *emP = m.emit_packed;
    const double tMM = m.t[T_MM], tXM = m.t[T_XM], tYM = m.t[T_YM];
    const int D = t.D, tb_to = t.tb_to, tb_from = t.tb_from;
    const double *endp = m.end_p[((t.flags & TILE_AT_END) && job.ragged_r) ? 1 : 0];
    const int64_t region = job.cell_base + (uint32_t) __ldg(gj + tb_to + 1).y;
    StagedPair *stM = stage + region, *stX = stage + stage_stride + region, *stY = stage + 2 * stage_stride + region;
    int cntM = 0, cntX = 0, cntY = 0;

    // symbols of cell (x, y) for the step INTO (x+1, y+1): sX[x], sY[y] (0-based)
    auto geo_at = [&](int dd) { return __ldg(gj + max(dd, 0)); };
    auto sym_x_at = [&](int2 gg) { return (unsigned) __ldg(sx + (gg.x + mod_g(lane - gg.x, G))); };
    auto sym_y_at = [&](int2 gg, int dd) { return (unsigned) __ldg(sy + (dd - (gg.x + mod_g(lane - gg.x, G)))); };
    auto f_at = [&](int2 gg) {
        const int x = gg.x + mod_g(lane - gg.x, G);
        return load_cell_cg(Fj, (int64_t) (uint32_t) gg.y + (x - gg.x));
    };

    Cell b1 = cell_neg_inf();
    double b2m = MB_NEG_INF;  // b(d+2).m of lane+1, i.e. the m shuffled one diagonal earlier
    int xlo1 = 0, w1 = 0, xlo2 = 0, w2 = 0;
    double total = MB_NEG_INF;
    int done = 0;
    // pipeline prologue: geometry of D+1 (for the width), D, D-1, D-2; symbols of D, D-1; emission of D; F of D
    int2 gUp = __ldg(gj + D + 1), gA = geo_at(D), gB = geo_at(D - 1), gC = geo_at(D - 2);
    unsigned cx0 = sym_x_at(gA), cy0 = sym_y_at(gA, D), cx1 = sym_x_at(gB), cy1 = sym_y_at(gB, max(D - 1, 0));
    double eM0 = __ldg(emP + cx0 * PACK_STRIDE + cy0);
    Cell fc0 = cell_neg_inf();
    {
        const int x = gA.x + mod_g(lane - gA.x, G);
        if (D <= tb_from && x < gA.x + (gUp.y - gA.y)) fc0 = f_at(gA);
    }
    const int n_it = D - tb_to;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
    for (int it = 0; it < n_max; it++) {
        const bool live = real && it < n_it;
        const int d = D - min(it, n_it - 1);
        const int2 gD = geo_at(d - 3);
        const unsigned cx2 = sym_x_at(gC), cy2 = sym_y_at(gC, max(d - 2, 0));
        const double eM1 = __ldg(emP + cx1 * PACK_STRIDE + cy1);
        const int xlo = gA.x, w = gUp.y - gA.y;
        const int x = xlo + mod_g(lane - xlo, G);
        const int y = d - x;
        const bool active = x < xlo + w;
        // F of the next diagonal down (used one iteration later)
        Cell fc1 = cell_neg_inf();
        {
            const int wn = gA.y - gB.y;
            const int xn = gB.x + mod_g(lane - gB.x, G);
            if (d - 1 <= tb_from && d - 1 > tb_to && xn < gB.x + wn) fc1 = f_at(gB);
        }
        const bool post = live && d <= tb_from;
        const bool do_total = post && (done % 10 == 0);
        Cell b;
        {
            // gather form of diagonalCalculationBackward: match term from (x+1,y+1) first, then the gap-Y term from
            // (x,y+1), then the gap-X term from (x+1,y) (see backward_cell in pairhmm_kernels.cuh)
            const double nx = shfl_g<G>(b1.x, next_lane);
            const double nm = shfl_g<G>(b1.m, next_lane);
            const bool has_mm = (unsigned) (x + 1 - xlo2) < (unsigned) w2;
            const bool has_y = (unsigned) (x - xlo1) < (unsigned) w1;
            const bool has_x = (unsigned) (x + 1 - xlo1) < (unsigned) w1;
            const double vm = has_mm ? b2m : MB_NEG_INF;
            const double vy = has_y ? b1.y : MB_NEG_INF;
            const double vx = has_x ? nx : MB_NEG_INF;
            const double *ex = m.ex_t[packed_base(cx0)];  // [0]=M->X, [1]=X->X, [2]=Y->X
            const double *ey = m.ey_t[packed_base(cy0)];  // [0]=M->Y, [1]=Y->Y, [2]=X->Y
            b.m = dadd(vm, dadd(eM0, tMM));
            b.x = dadd(vm, dadd(eM0, tXM));
            b.y = dadd(vm, dadd(eM0, tYM));
            b.m = logadd_tab(b.m, dadd(vy, ey[0]), lac);
            if (SWITCH) b.x = logadd_tab(b.x, dadd(vy, ey[2]), lac);
            b.y = logadd_tab(b.y, dadd(vy, ey[1]), lac);
            b.m = logadd_tab(b.m, dadd(vx, ex[0]), lac);
            b.x = logadd_tab(b.x, dadd(vx, ex[1]), lac);
            if (SWITCH) b.y = logadd_tab(b.y, dadd(vx, ex[2]), lac);
            b2m = nm;
            if (it == 0) {  // d == D: the backward sweep starts from the end state on every cell (:754-762)
                b.m = endp[0];
                b.x = endp[1];
                b.y = endp[2];
            }
        }
        b = cell_sel(active, b);
        if (__any_sync(0xffffffffu, do_total)) {
            // diagonalCalculationTotalProbability (pairwiseAligner.c:578-596); groups that do not sample on this
            // iteration run along (uniform control flow) and discard the result
            const int w_max = __reduce_max_sync(0xffffffffu, max(w, w1));
            const double dot = active ? cell_dot_tab(fc0, b, lac) : MB_NEG_INF;
            double tot = fold_diag<G>(dot, xlo, w, w_max, lac);
            // paths that jump over d with a match: match block of the forward step from F(d-1) into the cells
            // of d+1, dotted with B(d+1) (this lane's b1 is its cell of d+1)
            const int xlom = gB.x, wm = gA.y - gB.y;
            const int x1 = xlo1 + mod_g(lane - xlo1, G);
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
                    gm_ = logadd_tab(gm_, dadd(f.x, dadd(e, tXM)), lac);
                    gm_ = logadd_tab(gm_, dadd(f.y, dadd(e, tYM)), lac);
                }
                val = dadd(gm_, b1.m);
            }
            const double second = fold_diag<G>(val, xlo1, w1, w_max, lac);
            if (it > 0) tot = logadd_tab(tot, second, lac);
            if (do_total) total = tot;
        }
        {
            // posteriors of diagonal d (diagonalCalculationPosteriorProbs, :637-681) in the log domain; the ballots
            // inside stage_list are full-warp, so this block is executed by every group on every iteration
            const int rank = mod_g(lane - xlo, G);
            const double vM = dadd(dadd(fc0.m, b.m), -total);
            const bool okM = post && active && x > 0 && y > 0 && vM >= v_thr;
            stage_list<G>(okM, vM, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stM, cntM);
            if (INDELS) {
                const double vX = dadd(dadd(fc0.x, b.x), -total);
                const bool okX = post && active && x > 0 && vX >= v_thr;
                stage_list<G>(okX, vX, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stX, cntX);
                const double vY = dadd(dadd(fc0.y, b.y), -total);
                const bool okY = post && active && y > 0 && vY >= v_thr;
                stage_list<G>(okY, vY, x - 1 + job.x1, y - 1 + job.y1, rank, xlo, stY, cntY);
            }
            if (post) done++;
        }
        // rotate
        xlo2 = xlo1;
        w2 = w1;
        b1 = b;
        xlo1 = xlo;
        w1 = w;
        gUp = gA;
        gA = gB;
        gB = gC;
        gC = gD;
        cx0 = cx1;
        cy0 = cy1;
        cx1 = cx2;
        cy1 = cy2;
        eM0 = eM1;
        fc0 = fc1;
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
                                     int64_t stage_stride, Pair *__restrict__ out, int n_lists, int64_t out_capacity) {
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
            out[o + i] = q;
        }
    }
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
