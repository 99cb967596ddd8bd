// pairhmm_device.cuh — device-side arithmetic of the 3-state pair-HMM (sm_100a).
//
// Everything here is FP64 and rounds every multiply and add separately
// (__dmul_rn / __dadd_rn are never contracted into FMA), because the reference
// is plain SSE2 double arithmetic and its logAdd is a non-associative
// piecewise cubic (impl/pairwiseAligner.c:279-299): only the same operations
// in the same order reproduce its posteriors bit for bit.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mb {

#define MB_NEG_INF (__longlong_as_double(0xfff0000000000000LL))

// Flattened model as the kernels see it (built from margin_b200_model on the host).
struct DevModel {
    double t[9];       // MB_T_* order of include/margin_b200.h
    double match[16];  // ln substitution matrix
    double gap_x[5];   // [4] = ambiguous-symbol constant (impl/stateMachine.c:363-368)
    double gap_y[5];
    double ex_t[5][3];  // gap_x[b] + t[{OPEN_X, EXTEND_X, SWITCH_TO_X}]   (the reference's (eP + tP))
    double ey_t[5][3];  // gap_y[b] + t[{OPEN_Y, EXTEND_Y, SWITCH_TO_Y}]
    double end_p[2][3]; // [ragged][state]: endStateProb / raggedEndStateProb (impl/stateMachine.c:532-560)
    int32_t rle;
    int32_t max_rl;
    const double *rsm_scaled;  // 2.3025 * rsm[b][under][obs], device pointer (4*51*51)
    // Folded match emission for the tile kernels, indexed by packed symbols p = min(base,4)*51 + runLength:
    // emit_packed[pX*255 + pY] = match[xb][yb] + 2.3025*rsm[...] rounded once, exactly the sum the reference forms
    // per cell (impl/stateMachine.c:733-738).  Non-RLE models only use the runLength = 0 entries.
    const double *emit_packed;
};

enum { T_MM = 0, T_XM = 1, T_YM = 2, T_MX = 3, T_XX = 4, T_YX = 5, T_MY = 6, T_YY = 7, T_XY = 8 };

__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }

// lookup(): impl/pairwiseAligner.c:282-293.  Float literals promoted to double, Horner form,
// every operation rounded.  Coefficients are selected (no divergence).
__device__ __forceinline__ double la_lookup(double x) {
    const bool r1 = x <= 1.0, r2 = x <= 2.5, r3 = x <= 4.5;
    double c3 = r1 ? (double) -0.009350833524763f
                   : (r2 ? (double) -0.014532321752540f : (r3 ? (double) -0.004605031767994f : (double) -0.000458661602210f));
    double c2 = r1 ? (double) 0.130659527668286f
                   : (r2 ? (double) 0.139942324101744f : (r3 ? (double) 0.063427417320019f : (double) 0.009695946122598f));
    double c1 = r1 ? (double) 0.498799810682272f
                   : (r2 ? (double) 0.495635523139337f : (r3 ? (double) 0.695956496475118f : (double) 0.930734667215156f));
    double c0 = r1 ? (double) 0.693203116424741f
                   : (r2 ? (double) 0.692140569840976f : (r3 ? (double) 0.514272634594009f : (double) 0.168037164329057f));
    double v = dadd(dmul(c3, x), c2);
    v = dadd(dmul(v, x), c1);
    v = dadd(dmul(v, x), c0);
    return v;
}

// logAdd(): impl/pairwiseAligner.c:295-299.  hi/lo formulation is operation-identical:
// the reference computes lookup(max - min) + min, or returns max when min is -inf or the gap >= 7.5.
__device__ __forceinline__ double logadd(double x, double y) {
    const bool lt = x < y;
    const double hi = lt ? y : x;
    const double lo = lt ? x : y;
    if (lo == MB_NEG_INF) return hi;
    const double d = dadd(hi, -lo);
    if (d >= 7.5) return hi;
    return dadd(la_lookup(d), lo);
}

// The four cubic pieces of lookup() as a [piece][c3,c2,c1,c0] table (copied to shared memory by the tile kernels).
__device__ const double LA_COEF[16] = {
    (double) -0.009350833524763f, (double) 0.130659527668286f, (double) 0.498799810682272f, (double) 0.693203116424741f,
    (double) -0.014532321752540f, (double) 0.139942324101744f, (double) 0.495635523139337f, (double) 0.692140569840976f,
    (double) -0.004605031767994f, (double) 0.063427417320019f, (double) 0.695956496475118f, (double) 0.514272634594009f,
    (double) -0.000458661602210f, (double) 0.009695946122598f, (double) 0.930734667215156f, (double) 0.168037164329057f};

// Branch-free logAdd for the tile kernels; operation-for-operation the same arithmetic as logadd() above.
//   lo == -inf            -> hi - lo = +inf (or NaN when both are -inf) -> "d < 7.5" is false -> hi
//   d >= 7.5              -> hi
//   otherwise             -> lookup(d) + lo with the piece chosen by d <= 1, <= 2.5, <= 4.5
__device__ __forceinline__ double logadd_tab(double x, double y, const double *lac) {
    const bool lt = x < y;
    const double hi = lt ? y : x;
    const double lo = lt ? x : y;
    const double d = dadd(hi, -lo);
    const int idx = (int) (d > 1.0) + (int) (d > 2.5) + (int) (d > 4.5);
    const double2 c32 = *reinterpret_cast<const double2 *>(lac + 4 * idx);
    const double2 c10 = *reinterpret_cast<const double2 *>(lac + 4 * idx + 2);
    double v = dadd(dmul(c32.x, d), c32.y);
    v = dadd(dmul(v, d), c10.x);
    v = dadd(dmul(v, d), c10.y);
    v = dadd(v, lo);
    return (d < 7.5) ? v : hi;
}

// Symbol decoding (impl/stateMachine.c:116-122): base = low 8 bits, run length = high 8 bits.
__device__ __forceinline__ int gap_index(const DevModel &m, unsigned c) {
    unsigned b = m.rle ? (c & 255u) : c;
    return b >= 4u ? 4 : (int) b;
}

// Match emission (impl/stateMachine.c:378-383 and :733-738 with impl/repeatSubMatrix.c:11-35).
__device__ __forceinline__ double emit_match(const DevModel &m, unsigned cx, unsigned cy) {
    if (!m.rle) {
        return (cx >= 4u || cy >= 4u) ? -2.772588722 : m.match[cx * 4u + cy];
    }
    unsigned xb = cx & 255u, yb = cy & 255u;
    double sub = (xb >= 4u || yb >= 4u) ? -2.772588722 : m.match[xb * 4u + yb];
    unsigned tb = xb >= 4u ? 0u : xb;
    unsigned under = cx >> 8, obs = cy >> 8;
    double r = __ldg(m.rsm_scaled + ((size_t) tb * m.max_rl + under) * m.max_rl + obs);
    return dadd(sub, r);
}

struct Cell {
    double m, x, y;
};

__device__ __forceinline__ Cell cell_neg_inf() {
    Cell c;
    c.m = c.x = c.y = MB_NEG_INF;
    return c;
}

// cell_dotProduct (impl/pairwiseAligner.c:333-339)
__device__ __forceinline__ double cell_dot(const Cell &a, const Cell &b) {
    double tp = dadd(a.m, b.m);
    tp = logadd(tp, dadd(a.x, b.x));
    tp = logadd(tp, dadd(a.y, b.y));
    return tp;
}

}  // namespace mb
