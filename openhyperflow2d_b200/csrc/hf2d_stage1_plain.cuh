// hf2d_stage1_plain.cuh -- kernel K1a: Stage 1 (deeps2d_core.cpp:853-1026) and the residual / locally adaptive blending
// factor (:1062-1154) of the PLAIN tiles, one EQUATION at a time.
//
// For an unflagged interior node both pieces are independent per equation k: NextNode->S[k] needs S[k], A[k], B[k],
// beta[k] (and Src[k]) of the node and its four neighbours; the new beta[k] needs S[k], NextNode->S[k] and beta[k] of
// the node alone.  A work item is therefore (tile, k): a warp owns TI columns x 32 rows of ONE plane set, which makes it
// a pure streaming stencil -- every load of the item is issued before the first use (36-60 independent 256-byte
// requests per warp in flight instead of one equation's worth), the left/right neighbours of column c are the own
// values of columns c-1 / c+1 (each plane column is fetched once per item), the up/down neighbours come from the
// adjacent lanes by shuffle (lanes 0 / 31 fetch the row below / above the tile), and the 2 divisions + 1 square root of
// the blending factor run under the memory time.  The shadow S goes to the SN planes for K1b (k_point on plain tiles).
#pragma once
#include "hf2d_point.cuh"

#define K1A_COLS 4                // columns of a tile whose loads are issued together (8: 168 registers + spills; 4: one more latency per tile, 16 warps/SM)
#ifndef K1A_WARPS
#define K1A_WARPS 4
#endif
#define K1A_THREADS (32 * K1A_WARPS)
#ifndef K1A_MINB
#define K1A_MINB 4
#endif

struct K1aAcc { double rms, div; unsigned cnt; unsigned long long dd; };

// one work item with the equation index as a compile-time constant: the optional operands (Src for k >= 7, the dy2
// smoothing of the eps equation, F on axisymmetric grids) cost registers only where they exist
template <class KC, int K>
__device__ __forceinline__ void k1a_item(const Params& P, const StepState& st, int lane, int j, long long n_first, int TI, int want_rms, K1aAcc& acc) {
    constexpr uint32_t ct = CT_NODE_IS_SET, tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD;
    const long long ny = P.ny, N = P.N;
    const double dt = st.dt;
    const double dtdx = ddiv(dt, P.dx), dtdy = ddiv(dt, P.dy);                  // :843-844
    const double dyy = ddiv(P.dx, P.dx + P.dy), dxx = ddiv(P.dy, P.dx + P.dy);   // :643-644
    const double* __restrict__ Sk = P.S_in + (long long)K * N;
    const double* __restrict__ Ak = P.A + (long long)K * N;
    const double* __restrict__ Bk = P.B + (long long)K * N;
    double* __restrict__ betak = P.beta + (long long)K * N;
    double* __restrict__ SNk = P.SN + (long long)K * N;
    const unsigned f = eq_flags(ct, tt, K, KC::ns);     // compile-time (the eps equation carries the dy2 smoothing flag, hf2d_dev.cuh)
    constexpr bool SRC_ALWAYS = K >= 7;
    const bool has_src = SRC_ALWAYS || P.has_src;
    const bool trk_x = (P.need_dsx >> K) & 1u, trk_y = (P.need_dsy >> K) & 1u;
    const bool edge = lane == 0 || lane == 31;
    const long long eo = lane == 0 ? -1 : 1;                  // lane 0: the row below the tile, lane 31: the row above
    for (int cb = 0; cb < TI; cb += K1A_COLS) {
        const int nc = min(K1A_COLS, TI - cb);        // (= K1A_COLS for the default tile width: the predicates below are uniform)
        const long long n0 = n_first + (long long)cb * ny;        // node of column 0 of the batch
        // ---- every load of the batch, before the first use (plain pointer walks: one 64-bit add per column and plane) ----
        double s[K1A_COLS + 2], a[K1A_COLS + 2], b[K1A_COLS], be[K1A_COLS], se[K1A_COLS], bedge[K1A_COLS];
        double sr[K1A_COLS], dsu[K1A_COLS], bm2[K1A_COLS], fax[K1A_COLS];
        {
            const double* __restrict__ ps = Sk + (n0 - ny);
            const double* __restrict__ pa = Ak + (n0 - ny);
#pragma unroll
            for (int c = 0; c < K1A_COLS + 2; c++) {
                s[c] = 0.; a[c] = 0.;
                if (c < nc + 2) { s[c] = *ps; a[c] = *pa; }
                ps += ny; pa += ny;
            }
            const double* __restrict__ pb = Bk + n0;
            const double* __restrict__ pbe = betak + n0;
            const double* __restrict__ pse = Sk + (n0 + eo);
            const double* __restrict__ pbd = Bk + (n0 + eo);
            const double* __restrict__ psr = has_src ? P.Src + ((long long)K * N + n0) : nullptr;
            const double* __restrict__ pds = (f & 16) ? P.dSy_in[K] + (n0 + 1) : nullptr;
            const double* __restrict__ pfx = KC::axi ? P.F + ((long long)K * N + n0) : nullptr;
#pragma unroll
            for (int c = 0; c < K1A_COLS; c++) {
                b[c] = be[c] = sr[c] = se[c] = bedge[c] = dsu[c] = bm2[c] = fax[c] = 0.;
                if (c < nc) {
                    b[c] = *pb; be[c] = *pbe;
                    if (has_src) sr[c] = *psr;
                    if (edge) { se[c] = *pse; bedge[c] = *pbd; }
                    if (f & 16) { dsu[c] = *pds; bm2[c] = pb[-2]; }
                    if (KC::axi) fax[c] = *pfx;
                }
                pb += ny; pbe += ny; pse += ny; pbd += ny;
                if (has_src) psr += ny;
                if (f & 16) pds += ny;
                if (KC::axi) pfx += ny;
            }
        }
        // ---- Stage 1 + blending factor, column by column ----
        long long n = n0 - ny;
#pragma unroll
        for (int c = 0; c < K1A_COLS; c++) {
            if (c >= nc) break;
            n += ny;
            const double sc = s[c + 1], sl = s[c], srr = s[c + 2];
            double su = __shfl_down_sync(0xffffffffu, sc, 1), sd = __shfl_up_sync(0xffffffffu, sc, 1);
            double bu = __shfl_down_sync(0xffffffffu, b[c], 1), bd = __shfl_up_sync(0xffffffffu, b[c], 1);
            if (lane == 0) { sd = se[c]; bd = bedge[c]; }
            if (lane == 31) { su = se[c]; bu = bedge[c]; }
            const double beta = be[c], _beta = 1. - beta;
            double dXX = 0., dYY = 0.;
            if (f & 2) dXX = (a[c + 2] - a[c]) * 0.5;
            if (f & 4) dYY = (bu - bd) * 0.5;
            if (trk_x) P.dSx_out[K][n] = dXX;
            if (trk_y) P.dSy_out[K][n] = dYY;
            if (f & 16) {   // :1011-1013: up = last step's dSdy, down = this sweep's (the row below is an unflagged interior node too)
                const double dd = (b[c] - bm2[c]) * 0.5;
                dYY = (dsu[c] + dd) * 0.5;
            }
            double v;
            if (KC::axi) v = sc * beta + _beta * (dxx * (sl + srr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * (dYY + ddiv(fax[c], (double)(j + 1)))) + sr[c] * dt + 0.;
            else v = sc * beta + _beta * (dxx * (sl + srr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * dYY) + sr[c] * dt + 0.;
            SNk[n] = v;
            // residual + blending factor (:1062-1154)
            const bool res = sc != 0.;
            if (__any_sync(0xffffffffu, res)) {
                const double beta_min = OMIN(P.beta0, st.beta_scn);
                const double bb = beta_min * beta_min;
                double DD = 0., newb = 0.;
                auto ph = [&](auto& dv) {
                    double sqrt_RES;
                    DD = residual_dd(dv, sc, v, res, sqrt_RES, P.bFF);
                    const double xx = (P.bFF >= 4) ? sqrt_RES : (P.bFF >= 2) ? DD * DD : DD;
                    const double cap = (P.bFF & 1) ? (beta_min + beta) * 0.5 : beta_min;
                    const double quo = dv.div(bb, beta_min + xx, res);
                    newb = OMIN(cap, quo);
                };
                if (want_rms) { DivExact e; ph(e); }
                else { DivFast fdv; ph(fdv); if (__any_sync(0xffffffffu, fdv.is_bad())) { DivExact e; ph(e); } }
                if (res) betak[n] = newb;
                if (want_rms && res) {
                    if (P.isAlternateRMS) { acc.div += sc * sc; acc.rms += v - sc; } else { acc.rms += DD * DD; }
                    acc.cnt++;
                    if (DD != DD) atomicMax(&P.st->nan_last[K], (unsigned long long)n + 1ull);
                    else { const unsigned long long d = d2bits(DD); acc.dd = d > acc.dd ? d : acc.dd; }
                }
            }
        }
    }
}

template <class KC>
__global__ void __launch_bounds__(K1A_THREADS, K1A_MINB)
k_stage1_plain(Params P, const int* __restrict__ tiles, int n_tiles, int n_sb, int TI, int want_rms) {
    static_assert(KC::plain, "plain tiles only");
    constexpr int NE_C = !KC::ns ? 7 : (KC::ttc & TCT_Prandtl_Model) ? 7 : (KC::ttc & TCT_k_eps_Model) ? 9 : (KC::ttc & TCT_SA_Model) ? 8 : 7;
    const StepState st = *P.st;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // item = k * n_tiles + t: the warps of a block work on vertically adjacent tiles of the same plane set (contiguous memory)
    const long long item = (long long)blockIdx.x * K1A_WARPS + warp;
    const bool in_range = item < (long long)n_tiles * NE_C;
    const int k = in_range ? (int)(item / n_tiles) : 0;
    const int tile = in_range ? tiles[(int)(item - (long long)k * n_tiles)] : 0;
    const int tx = tile / n_sb, sb = tile - tx * n_sb;
    const int j = sb * TILE_H + lane;                 // plain tiles lie inside the grid with their whole ring: always a valid row
    const long long n_first = (long long)(P.i_lo + tx * TI) * P.ny + j;
    K1aAcc acc; acc.rms = 0.; acc.div = 0.; acc.cnt = 0u; acc.dd = 0ull;
    if (in_range) {
        switch (k) {
        case 0: k1a_item<KC, 0>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 1: k1a_item<KC, 1>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 2: k1a_item<KC, 2>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 3: k1a_item<KC, 3>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 4: k1a_item<KC, 4>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 5: k1a_item<KC, 5>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 6: k1a_item<KC, 6>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 7: if constexpr (NE_C > 7) k1a_item<KC, 7>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        case 8: if constexpr (NE_C > 8) k1a_item<KC, 8>(P, st, lane, j, n_first, TI, want_rms, acc); break;
        }
    }
    if (want_rms) {      // uniform per launch
        const double r0 = warp_sum(acc.rms), r1 = warp_sum(acc.div);
        const unsigned c = __reduce_add_sync(0xffffffffu, acc.cnt);
        const unsigned long long d = warp_max_u64(acc.dd);
        if (lane == 0 && c) {
            if (r0 != 0.) atomicAdd(&P.st->RMS[k], r0);
            if (r1 != 0.) atomicAdd(&P.st->sumDiv[k], r1);
            atomicAdd(&P.st->iRMS[k], (unsigned long long)c);
            if (d) atomicMax(&P.st->DDmax_bits[k], d);
        }
    }
}
