// pairhmm_tiles_f32.cuh — the FP32 "fast" variant of the tile kernels of pairhmm_tiles.cuh (sm_100a).
//
// Same decomposition (independent tiles with warm-up, lane = x mod G, register software pipeline), different
// arithmetic: cell values are FP32 and are kept small by renormalisation, so that the posteriors stay within the
// tolerance BASELINE.json's north_star states (|dp| <= 1e-4) of the reference's FP64 ones (plain
// FP32 loses 3e-2 after 5 850 diagonals, FP32 with a running offset 1.7e-5).  What changes against the FP64 kernels:
//
//   * logAdd (impl/pairwiseAligner.c:279-299) keeps the reference's four cubic pieces — its systematic error of up to
//     5.5e-4 per call is part of the result — but is evaluated as  max(x,y) + g(|x-y|),  g(d) = lookup(d) - d, with the
//     piece picked by ONE rounding-mode add: the breakpoints 1, 2.5, 4.5, 7.5 are multiples of 1/2, so ceil(2d) is
//     the row of a 17-row coefficient table, and fma.rp(d, 2, 1.5 * 2^23) leaves exactly that integer in the low
//     mantissa bits.  10 full-rate instructions instead of 24 FP64 ones; -inf needs no special case (d is clamped to 8,
//     whose row is all zeros).
//   * Every RENORM iterations a tile subtracts the largest value of its current diagonal from all live registers.
//     The forward kernel records the shift of a diagonal it owns in shiftF[diagonal]; the backward kernel folds the
//     shifts of both sweeps into the per-diagonal constant  c = -total - sum(shiftF) + sum(shiftB)  it adds to
//     F + B, so a log posterior is formed from three numbers of magnitude ~10.
//   * F is stored as float4 {m, x, y, -} (16 B per cell instead of 24), one STG.128 / LDG.128 per cell.
//
// The staged records, the gather, the tile tables, the seam guard and the repair logic are shared with the FP64 path.
// Tiles with a diagonal wider than 32 cells and jobs swept sequentially keep the FP64 general kernels (exact).
#pragma once
#include "pairhmm_tiles.cuh"

namespace mb {

constexpr int RENORM = 32;            // iterations between renormalisations
constexpr float LA_MAGIC = 12582912.0f;  // 1.5 * 2^23: fma.rp(d, 2, LA_MAGIC) has ceil(2 d) in its low mantissa bits
constexpr uint32_t LA_MAGIC_BITS = 0x4B400000u;
#define MB_NEG_INF_F (__int_as_float(0xff800000))

struct CellF {
    float m, x, y;
};

__device__ __forceinline__ CellF cellf_neg_inf() {
    CellF c;
    c.m = c.x = c.y = MB_NEG_INF_F;
    return c;
}

struct TileSharedF {
    // g(d) = lookup(d) - d, row = ceil(2 d): {c3, c2, c1 - 1, c0}; row 16 is zero (d >= 7.5).  Eight copies, [row][copy]:
    // lane j of a quarter-warp reads copy j, which lies in its own four banks, so an LDS.128 of eight different rows
    // never conflicts (one copy cost 0.26 extra wavefronts per load on the C2 chunk: rows r and r + 8 share banks).
    float4 lag[17][8];
    float4 gx[2][5];     // [model][base] {M->X, X->X, Y->X, 0}: the reference's (eP + tP) of the gap-X transitions
    float4 gy[2][5];     // [model][base] {M->Y, Y->Y, X->Y, 0}
    float tm[2][4];      // [model] {M->M, X->M, Y->M, 0}
    float endp[2][2][4]; // [model][ragged] end state
    const float *emit[2];
};


template <int THREADS>
__device__ __forceinline__ void tile_shared_init_f(TileSharedF &s, const DevModel *__restrict__ models,
                                                   const float *const *__restrict__ emit_f) {
    for (int i = threadIdx.x; i < 17 * 8; i += THREADS) {
        const int p = la_piece_of_row(i >> 3);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p < 4) {
            v.x = (float) LA_COEF[4 * p];
            v.y = (float) LA_COEF[4 * p + 1];
            v.z = (float) (LA_COEF[4 * p + 2] - 1.0);
            v.w = (float) LA_COEF[4 * p + 3];
        }
        s.lag[i >> 3][i & 7] = v;
    }
    if (threadIdx.x >= 32 && threadIdx.x < 42) {
        const int k = threadIdx.x - 32, md = k / 5, b = k % 5;
        s.gx[md][b] = make_float4((float) models[md].ex_t[b][0], (float) models[md].ex_t[b][1], (float) models[md].ex_t[b][2], 0.f);
        s.gy[md][b] = make_float4((float) models[md].ey_t[b][0], (float) models[md].ey_t[b][1], (float) models[md].ey_t[b][2], 0.f);
    }
    if (threadIdx.x >= 64 && threadIdx.x < 66) {
        const int md = threadIdx.x - 64;
        s.tm[md][0] = (float) models[md].t[T_MM];
        s.tm[md][1] = (float) models[md].t[T_XM];
        s.tm[md][2] = (float) models[md].t[T_YM];
        s.tm[md][3] = 0.f;
        for (int r = 0; r < 2; r++) {
            for (int k = 0; k < 3; k++) s.endp[md][r][k] = (float) models[md].end_p[r][k];
            s.endp[md][r][3] = 0.f;
        }
        s.emit[md] = emit_f[md];
    }
    __syncthreads();
}

__device__ __forceinline__ float4 lds_f32x4(uint32_t a) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}

// logAdd in FP32.  lag_c = shared-space address of this lane's copy of TileSharedF::lag (copy = lane & 7) minus
// 128 * LA_MAGIC_BITS (mod 2^32), so that the row address is one multiply-add on the raw bits of the magic sum.
__device__ __forceinline__ float logadd_f(float x, float y, uint32_t lag_c) {
    const float hi = fmaxf(x, y);
    const float dc = fminf(fabsf(x - y), 8.0f);  // NaN (-inf - -inf) and +inf clamp to 8: the all-zero row
    float t;
    asm("fma.rp.f32 %0, %1, 0f40000000, 0f4B400000;" : "=f"(t) : "f"(dc));
    const uint32_t a = __float_as_uint(t) * 128u + lag_c;
    const float4 c = lds_f32x4(a);
    float v = fmaf(c.x, dc, c.y);
    v = fmaf(v, dc, c.z);
    v = fmaf(v, dc, c.w);
    return hi + v;
}

template <int G>
__device__ __forceinline__ float shfl_gf(float v, int src) { return __shfl_sync(0xffffffffu, v, src, G); }

// largest finite value among the lanes of a group (0 when there is none: nothing to renormalise by)
template <int G>
__device__ __forceinline__ float group_max_f(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o, G));
    return v > -1e30f ? v : 0.f;
}

__device__ __forceinline__ void stg_cell_f(float *q, bool pred, const CellF &c) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.global.v4.f32 [%1], {%2, %3, %4, %5};\n\t}"
                 :: "r"((unsigned) pred), "l"(q), "f"(c.m), "f"(c.x), "f"(c.y), "f"(0.f)
                 : "memory");
}

__device__ __forceinline__ CellF ldg_cell_f(const float *base, int64_t cell) {
    const float4 v = __ldcg(reinterpret_cast<const float4 *>(base) + cell);
    CellF c;
    c.m = v.x;
    c.x = v.y;
    c.y = v.z;
    return c;
}

// ------------------------------------------------------------------------------------------------ forward tile
//
// forward_tile_kernel (pairhmm_tiles.cuh) in FP32: same period-4 register pipeline, validity by lane window, one
// select per target state.  `pend` is the shift a renormalisation applied in front of the next diagonal; every
// diagonal the tile owns gets its entry of shiftF written (0 or the shift), so a re-run of the tile leaves no stale one.
template <int G, int THREADS, bool SWITCH>
__global__ void __launch_bounds__(THREADS, 8)
forward_tile_kernel_f(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                      int n_ids, const DevModel *__restrict__ models, const float *const *__restrict__ emit_f,
                      const uint8_t *__restrict__ px, const uint8_t *__restrict__ py, const int2 *__restrict__ geo,
                      float *__restrict__ F, float *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off,
                      float *__restrict__ shiftF) {
    __shared__ __align__(16) TileSharedF sh;
    tile_shared_init_f<THREADS>(sh, models, emit_f);
    const int tid = tid_x_once();
    const int g = (blockIdx.x * THREADS + tid) / G;
    const bool real = g < n_ids;  // groups past the end shadow the last tile and store nothing
    const int lane = tid & (G - 1);
    const int prev_lane = mod_g(lane - 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const uint32_t lag_c = smem_u32(&sh.lag[0][tid & 7]) - 128u * LA_MAGIC_BITS;
    const uint32_t gx_s = smem_u32(&sh.gx[job.slot][0]), gy_s = smem_u32(&sh.gy[job.slot][0]);
    const int N = job.lx + job.ly;
    const uint8_t *sx = px + job.symx_off - 1, *sy = py + job.symy_off - 1;  // sx[x] = symbol of column x (1-based)
    const int2 *gj = geo + job.diag_off;
    float *Fj = F + 4 * job.cell_base;
    float *bnd = Fbnd + 4 * slot_bnd_off[slot];
    float *sFj = shiftF + job.diag_off;
    const float *emP = sh.emit[job.slot];
    const float tMM = sh.tm[job.slot][0], tXM = sh.tm[job.slot][1], tYM = sh.tm[job.slot][2];
    const int d0 = t.f0 + 1;  // first diagonal of the loop
    const int tb_to = t.tb_to;

    // ---- diagonal f0: the true start state (stateMachine.c:521-530) or a flat state for the warm-up
    int xlo1, xlo2 = -2 * G;  // nothing on "diagonal f0 - 1" is ever valid
    CellF f1 = cellf_neg_inf();
    float midm = MB_NEG_INF_F, midx = MB_NEG_INF_F, midy = MB_NEG_INF_F;
    {
        const int2 ga = __ldg(gj + t.f0), gb = __ldg(gj + t.f0 + 1);
        xlo1 = ga.x;
        const int r = mod_g(lane - xlo1, G);
        if (r < gb.y - ga.y) {
            if (t.f0 == 0) {
                f1.m = job.ragged_l ? MB_NEG_INF_F : 0.f;
                f1.x = job.ragged_l ? 0.f : MB_NEG_INF_F;
                f1.y = f1.x;
            } else {
                f1.m = f1.x = f1.y = 0.f;
            }
            stg_cell_f(bnd + 4 * r, real && t.f0 == tb_to, f1);
        }
    }
    auto geo_at = [&](int dd) { return __ldg(gj + min(dd, N + 1)); };
    auto lane_x = [&](int xlo) { return xlo + mod_g(lane - xlo, G); };
    // ---- pipeline prologue
    int2 gr[4];
#pragma unroll
    for (int k = 0; k < 4; k++) gr[k] = geo_at(d0 + k);
    unsigned csx[2], csy[2];
    float eM[2];
    uint32_t exa[2], eya[2];
    {
        const int xa = lane_x(gr[0].x);
        const unsigned cx = __ldg(sx + xa), cy = __ldg(sy + (min(d0, N) - xa));
        eM[0] = __ldg(emP + cx * PACK_STRIDE + cy);
        exa[0] = gx_s + 16u * (unsigned) packed_base(cx);
        eya[0] = gy_s + 16u * (unsigned) packed_base(cy);
        const int xb = lane_x(gr[1].x);
        csx[1] = __ldg(sx + xb);
        csy[1] = __ldg(sy + (min(d0 + 1, N) - xb));
    }
    const int n_it = t.tb_from - t.f0;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
    float pend = 0.f;
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
                exa[(k + 1) & 1] = gx_s + 16u * (unsigned) packed_base(cx);
                eya[(k + 1) & 1] = gy_s + 16u * (unsigned) packed_base(cy);
            }
            // ---- diagonal d
            const int xlo = gr[k].x;
            const uint32_t co = (uint32_t) gr[k].y;
            const int w = gr[(k + 1) & 3].y - gr[k].y;
            const int r = mod_g(lane - xlo, G);
            const int x = xlo + r;
            const bool active = r < w;
            const float lm = shfl_gf<G>(f1.m, prev_lane);
            const float lx_ = shfl_gf<G>(f1.x, prev_lane);
            const float ly_ = shfl_gf<G>(f1.y, prev_lane);
            const int xe = active ? x : FAR_X;  // a lane outside its diagonal fails every window test
            const bool ok_lower = (unsigned) (xe - 1 - xlo1) < (unsigned) G;
            const bool ok_upper = (unsigned) (xe - xlo1) < (unsigned) G;
            const bool ok_middle = (unsigned) (xe - 1 - xlo2) < (unsigned) G;
            const float4 ex = lds_f32x4(exa[k & 1]), ey = lds_f32x4(eya[k & 1]);
            const float e = eM[k & 1];
            CellF c;
            // stateMachine3_cellCalculate (stateMachine.c:562-586): X from lower, M from middle, Y from upper
            c.x = logadd_f(lm + ex.x, lx_ + ex.y, lag_c);
            if (SWITCH) c.x = logadd_f(c.x, ly_ + ex.z, lag_c);
            c.m = logadd_f(midm + (e + tMM), midx + (e + tXM), lag_c);
            c.m = logadd_f(c.m, midy + (e + tYM), lag_c);
            c.y = logadd_f(f1.m + ey.x, f1.y + ey.y, lag_c);
            if (SWITCH) c.y = logadd_f(c.y, f1.x + ey.z, lag_c);
            c.x = ok_lower ? c.x : MB_NEG_INF_F;
            c.m = ok_middle ? c.m : MB_NEG_INF_F;
            c.y = ok_upper ? c.y : MB_NEG_INF_F;
            {
                stg_cell_f(Fj + 4 * ((int64_t) co + r), live && active && d > tb_to, c);
                if (__any_sync(0xffffffffu, live && d == tb_to)) {  // the boundary diagonal: once per tile
                    stg_cell_f(bnd + 4 * r, live && active && d == tb_to, c);
                }
                if (live && lane == 0 && d > tb_to) sFj[d] = pend;  // frame change in front of this diagonal
                pend = 0.f;
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
        if ((it0 & (RENORM - 1)) == RENORM - 4) {
            // renormalise: every live register moves by the largest value of the last diagonal; the next diagonal's
            // stored values are relative to the new frame
            const float s = group_max_f<G>(fmaxf(f1.m, fmaxf(f1.x, f1.y)));
            f1.m -= s;
            f1.x -= s;
            f1.y -= s;
            midm -= s;
            midx -= s;
            midy -= s;
            pend = s;
        }
    }
}

// ----------------------------------------------------------------------------------------------- backward tile

// fold_diag2 / cell_dot_t of pairhmm_tiles.cuh in FP32 (diagonalCalculationTotalProbability, pairwiseAligner.c:578-596)
template <int G>
__device__ __forceinline__ void fold_diag2_f(float va, int xa, int wa, float vb, int xb, int wb, int w_max,
                                             uint32_t lag_c, float &acc_a, float &acc_b) {
    acc_a = MB_NEG_INF_F;
    acc_b = MB_NEG_INF_F;
#pragma unroll 1
    for (int i = 0; i < w_max; i++) {
        const float a = shfl_gf<G>(va, mod_g(xa + i, G));
        const float b = shfl_gf<G>(vb, mod_g(xb + i, G));
        acc_a = logadd_f(acc_a, i < wa ? a : MB_NEG_INF_F, lag_c);
        acc_b = logadd_f(acc_b, i < wb ? b : MB_NEG_INF_F, lag_c);
    }
}

__device__ __forceinline__ float cell_dot_f(const CellF &a, const CellF &b, uint32_t lag_c) {
    float tp = a.m + b.m;
    tp = logadd_f(tp, a.x + b.x, lag_c);
    tp = logadd_f(tp, a.y + b.y, lag_c);
    return tp;
}

template <int G>
struct StageLaneF {
    int shift, rot_r;
    unsigned below;  // mask of the cells with smaller x
    __device__ __forceinline__ StageLaneF(int xlo, int rank, int gshift)
        : shift(gshift), rot_r(mod_g(xlo, G)), below((1u << rank) - 1u) {}
    __device__ __forceinline__ void put(bool ok, float v, int x, int y, StagedPair *dst, int &cnt) const {
        unsigned bits = __ballot_sync(0xffffffffu, ok);
        unsigned rot;
        if constexpr (G == 32) {
            rot = __funnelshift_r(bits, bits, rot_r);
        } else {
            bits = (bits >> shift) & ((1u << G) - 1u);
            rot = (bits >> rot_r) | (bits << (G - rot_r));  // bits above G are masked by `below`
        }
        if (ok) {
            StagedPair sp;
            sp.v = (double) v;
            sp.x = x;
            sp.y = y;
            dst[cnt + __popc(rot & below)] = sp;
        }
        cnt += __popc(bits);
    }
};

// backward_tile_kernel (pairhmm_tiles.cuh) in FP32.  `c` is the constant added to F + B on the current diagonal:
// minus the total probability sampled every 10th posterior diagonal, corrected by every frame change since the
// sample — the forward sweep's (shiftF, read per diagonal) and this sweep's own renormalisations.
template <int G, int THREADS, bool INDELS, bool SWITCH, bool GUARD>
__global__ void __launch_bounds__(THREADS, 6)
backward_tile_kernel_f(const Job *__restrict__ jobs, const Tile *__restrict__ tiles, const int32_t *__restrict__ ids,
                       int n_ids, const DevModel *__restrict__ models, const float *const *__restrict__ emit_f,
                       const uint8_t *__restrict__ px, const uint8_t *__restrict__ py, const int2 *__restrict__ geo,
                       const float *__restrict__ F, float *__restrict__ Fbnd, const int64_t *__restrict__ slot_bnd_off,
                       const float *__restrict__ shiftF, float v_thr, StagedPair *__restrict__ stage, int64_t stage_stride,
                       int64_t *__restrict__ seg_counts, int total_slots) {
    __shared__ __align__(16) TileSharedF sh;
    tile_shared_init_f<THREADS>(sh, models, emit_f);
    const int tid = tid_x_once();
    const int g = (blockIdx.x * THREADS + tid) / G;
    const bool real = g < n_ids;
    const int lane = tid & (G - 1);
    const int next_lane = mod_g(lane + 1, G);
    const int slot = ids[min(g, n_ids - 1)];
    const Tile t = tiles[slot];
    const Job job = jobs[t.job];
    const uint32_t lag_c = smem_u32(&sh.lag[0][tid & 7]) - 128u * LA_MAGIC_BITS;
    const uint32_t gx_s = smem_u32(&sh.gx[job.slot][0]), gy_s = smem_u32(&sh.gy[job.slot][0]);
    const uint8_t *sx = px + job.symx_off, *sy = py + job.symy_off;
    const int2 *gj = geo + job.diag_off;
    const float *Fj = F + 4 * job.cell_base;
    float *bnd = Fbnd + 4 * slot_bnd_off[slot];  // [F(tb_to) | B(tb_to + 1) | B(tb_from + 1)], see Tile
    const int bnd_cells = (int) ((slot_bnd_off[slot + 1] - slot_bnd_off[slot]) / 3);
    const float *sFj = shiftF + job.diag_off;
    const float *emP = sh.emit[job.slot];
    const float tMM = sh.tm[job.slot][0], tXM = sh.tm[job.slot][1], tYM = sh.tm[job.slot][2];
    const int D = t.D, tb_to = t.tb_to, tb_from = t.tb_from;
    const float *endp = sh.endp[job.slot][((t.flags & TILE_AT_END) && job.ragged_r) ? 1 : 0];
    const int64_t region = job.cell_base + (uint32_t) __ldg(gj + tb_to + 1).y;
    StagedPair *stM = stage + region, *stX = stage + stage_stride + region, *stY = stage + 2 * stage_stride + region;
    int cntM = 0, cntX = 0, cntY = 0;
    const int xoff = job.x1 - 1, yoff = job.y1 - 1;

    auto geo_at = [&](int dd) { return __ldg(gj + max(dd, 0)); };
    auto lane_x = [&](int xlo) { return xlo + mod_g(lane - xlo, G); };

    CellF b1 = cellf_neg_inf();
    float b2m = MB_NEG_INF_F;  // b(d+2).m of lane+1, i.e. the m shuffled one diagonal earlier
    int xlo2 = -2 * G, w1 = 0;
    float c = 0.f;        // see above; meaningful from the first total-probability sample on
    int until_total = 0;  // posterior diagonals until the next total-probability sample
    // ---- pipeline prologue: geometry of D+1 .. D-2 (+ D-3 in flight); symbols, emission and F of D, symbols of D-1
    int2 gUp = geo_at(D + 1), gA = geo_at(D), gB = geo_at(D - 1), gC = geo_at(D - 2);
    int2 gn[2];  // landing registers: a geometry word is loaded one step and moved into the chain the next
    gn[1] = geo_at(D - 3);
    unsigned csx[2], csy[2];
    float eM[2], sf[2];
    uint32_t exa[2], eya[2];
    CellF fc[2];
    {
        const int xa = lane_x(gA.x);
        const unsigned cx = __ldg(sx + xa), cy = __ldg(sy + (D - xa));
        eM[0] = __ldg(emP + cx * PACK_STRIDE + cy);
        exa[0] = gx_s + 16u * (unsigned) packed_base(cx);
        eya[0] = gy_s + 16u * (unsigned) packed_base(cy);
        const int xb = lane_x(gB.x);
        csx[1] = __ldg(sx + xb);
        csy[1] = __ldg(sy + (max(D - 1, 0) - xb));
        fc[0] = cellf_neg_inf();
        if (D <= tb_from && xa - gA.x < gUp.y - gA.y) fc[0] = ldg_cell_f(Fj, (int64_t) (uint32_t) gA.y + (xa - gA.x));
        sf[0] = D <= tb_from ? __ldg(sFj + D) : 0.f;
    }
    const int n_it = D - tb_to;
    const int n_max = __reduce_max_sync(0xffffffffu, n_it);
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
                exa[(k + 1) & 1] = gx_s + 16u * (unsigned) packed_base(cx);
                eya[(k + 1) & 1] = gy_s + 16u * (unsigned) packed_base(cy);
                // F and the forward frame change of the next diagonal down (used one iteration later)
                const int rn = mod_g(lane - gB.x, G);
                CellF fn = cellf_neg_inf();
                const bool own = d - 1 <= tb_from && d - 1 > tb_to;
                if (own && rn < gA.y - gB.y) fn = ldg_cell_f(Fj, (int64_t) (uint32_t) gB.y + rn);
                fc[(k + 1) & 1] = fn;
                sf[(k + 1) & 1] = own ? __ldg(sFj + d - 1) : 0.f;
            }
            const int xlo = gA.x, w = gUp.y - gA.y, xlo1 = gUp.x;
            const int rank = mod_g(lane - xlo, G);
            const int x = xlo + rank;
            const int y = d - x;
            const bool active = rank < w;
            const bool post = live && d <= tb_from;
            const bool do_total = post && until_total == 0;
            const CellF f0c = fc[k & 1];
            const float sfd = sf[k & 1];  // frame change of the forward sweep in front of diagonal d
            CellF b;
            {
                // gather form of diagonalCalculationBackward: match term from (x+1,y+1) first, then the gap-Y term from
                // (x,y+1), then the gap-X term from (x+1,y) (see backward_cell in pairhmm_kernels.cuh)
                const float nx = shfl_gf<G>(b1.x, next_lane);
                const float nm = shfl_gf<G>(b1.m, next_lane);
                const int xe = active ? x : FAR_X;  // a lane outside its diagonal fails every window test
                const bool has_mm = (unsigned) (xe + 1 - xlo2) < (unsigned) G;
                const bool has_y = (unsigned) (xe - xlo1) < (unsigned) G;
                const bool has_x = (unsigned) (xe + 1 - xlo1) < (unsigned) G;
                const float vm = has_mm ? b2m : MB_NEG_INF_F;
                const float vy = has_y ? b1.y : MB_NEG_INF_F;
                const float vx = has_x ? nx : MB_NEG_INF_F;
                const float4 ex = lds_f32x4(exa[k & 1]);  // {M->X, X->X, Y->X}
                const float4 ey = lds_f32x4(eya[k & 1]);  // {M->Y, Y->Y, X->Y}
                const float e = eM[k & 1];
                b.m = vm + (e + tMM);
                b.x = vm + (e + tXM);
                b.y = vm + (e + tYM);
                b.m = logadd_f(b.m, vy + ey.x, lag_c);
                if (SWITCH) b.x = logadd_f(b.x, vy + ey.z, lag_c);
                b.y = logadd_f(b.y, vy + ey.y, lag_c);
                b.m = logadd_f(b.m, vx + ex.x, lag_c);
                b.x = logadd_f(b.x, vx + ex.y, lag_c);
                if (SWITCH) b.y = logadd_f(b.y, vx + ex.z, lag_c);
                b2m = nm;
                if (k == 0 && it0 == 0) {  // d == D: the backward sweep starts from the end state on every cell (:754-762)
                    b.m = active ? endp[0] : MB_NEG_INF_F;
                    b.x = active ? endp[1] : MB_NEG_INF_F;
                    b.y = active ? endp[2] : MB_NEG_INF_F;
                }
            }
            if (GUARD) {
                const bool low = d == tb_to + 1;
                const bool keep = live && (low || (d == tb_from + 1 && it > 0));
                if (__any_sync(0xffffffffu, keep)) {  // two diagonals per tile
                    stg_cell_f(bnd + 4 * ((low ? BND_B_LOW : BND_B_HIGH) * bnd_cells + rank), keep && active, b);
                }
            }
            if (__any_sync(0xffffffffu, do_total)) {
                // diagonalCalculationTotalProbability (pairwiseAligner.c:578-596) in the frame of diagonal d
                const int w_max = __reduce_max_sync(0xffffffffu, max(w, w1));
                const float dot = active ? cell_dot_f(f0c, b, lag_c) : MB_NEG_INF_F;
                const int xlom = gB.x, wm = gA.y - gB.y;
                const int x1 = lane_x(xlo1);
                const int y1 = d + 1 - x1;
                float val = MB_NEG_INF_F;
                if (do_total && it > 0 && x1 < xlo1 + w1) {
                    float gm_ = MB_NEG_INF_F;
                    if ((unsigned) (x1 - 1 - xlom) < (unsigned) wm) {
                        const CellF f = (d - 1 == tb_to) ? ldg_cell_f(bnd, x1 - 1 - xlom)
                                                         : ldg_cell_f(Fj, (int64_t) (uint32_t) gB.y + (x1 - 1 - xlom));
                        const float e = __ldg(emP + (unsigned) __ldg(sx + x1 - 1) * PACK_STRIDE + (unsigned) __ldg(sy + y1 - 1));
                        gm_ = f.m + (e + tMM);
                        gm_ = logadd_f(gm_, f.x + (e + tXM), lag_c);
                        gm_ = logadd_f(gm_, f.y + (e + tYM), lag_c);
                    }
                    val = (gm_ - sfd) + b1.m;  // F(d-1) was stored in the frame before diagonal d's change
                }
                float tot, second;
                fold_diag2_f<G>(dot, xlo, w, val, xlo1, w1, w_max, lag_c, tot, second);
                if (it > 0) tot = logadd_f(tot, second, lag_c);
                if (do_total) c = -tot;
            }
            {
                const StageLaneF<G> sl(xlo, rank, (tid & 31) & ~(G - 1));
                const bool pa = post && active;
                const float vM = (f0c.m + b.m) + c;
                sl.put(pa && x > 0 && y > 0 && vM >= v_thr, vM, x + xoff, y + yoff, stM, cntM);
                if (INDELS) {
                    const float vX = (f0c.x + b.x) + c;
                    sl.put(pa && x > 0 && vX >= v_thr, vX, x + xoff, y + yoff, stX, cntX);
                    const float vY = (f0c.y + b.y) + c;
                    sl.put(pa && y > 0 && vY >= v_thr, vY, x + xoff, y + yoff, stY, cntY);
                }
                if (post) until_total = until_total == 0 ? 9 : until_total - 1;
            }
            c -= sfd;  // the next diagonal down was stored before this frame change
            // ---- carry
            b1 = b;
            xlo2 = xlo1;
            w1 = w;
            gUp = gA;
            gA = gB;
            gB = gC;
            gC = gn[(k + 1) & 1];
        }
        if ((it0 & (RENORM - 1)) == RENORM - 2) {
            const float s = group_max_f<G>(fmaxf(b1.m, fmaxf(b1.x, b1.y)));
            b1.m -= s;
            b1.x -= s;
            b1.y -= s;
            b2m -= s;
            c += s;
        }
    }
    if (real && lane == 0) {
        seg_counts[slot] = cntM;
        seg_counts[total_slots + slot] = cntX;
        seg_counts[2 * total_slots + slot] = cntY;
    }
}

}  // namespace mb
