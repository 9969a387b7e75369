// hf2d_stage1_tma.cuh -- kernel K1a with TMA tile staging: Stage 1 (deeps2d_core.cpp:853-1026) and the residual /
// locally adaptive blending factor (:1062-1154) of the PLAIN tiles, one (tile, equation) work item per warp.
//
// Same arithmetic as k_stage1_plain (hf2d_stage1_plain.cuh, the fallback for grids TMA cannot address: odd ny, tile
// width != 8).  What changes is how the operands arrive.  The per-equation planes are 3-D tensors (row, column, equation)
// to the TMA unit; one elected lane issues 4 to 7 `cp.async.bulk.tensor.3d` box loads per item --
//     S    36 rows x 10 columns  (the tile with its halo ring: left/right columns, rows below/above; box starts on an even row)
//     A    32 x 10               (flux differences in x need columns c-1, c+1)
//     B    36 x 8                (rows j0-2 .. j0+33: up / down neighbours, and row j-2 for the dy2 smoothing of the eps equation)
//     beta 32 x 8, Src 32 x 8 (k >= 7 or decks with sources), dSdy 34 x 8 (eps; read at row j+1), F 32 x 8 (axisymmetric)
// -- into the warp's shared-memory stage and arms an mbarrier with the byte count; the warp waits on the barrier and then
// reads every operand, halo included, with LDS at compile-time offsets: no address arithmetic, no shuffles or edge-lane
// special cases, no load predicates.  Bytes in flight cost shared memory, not registers: 16 warps per SM each with a
// 9.6-13.7 KB stage keep ~150 KB per SM in flight, several times what HBM latency x bandwidth needs.
#pragma once
#include <cuda.h>
#include "hf2d_stage1_plain.cuh"

// (kept in device global memory: indexing a __grid_constant__ copy with a run-time index made the descriptor address illegal)
struct K1aMaps {                 // tensor maps of one grid (hf2d_gpu_upload_aos): (row, column, equation) unless noted
    CUtensorMap S[2];            // the two halves of the S ping-pong
    CUtensorMap A, B, beta, Src, F;
    CUtensorMap dSy[2];          // (row, column) of the tracked dSdy plane of the eps equation, both halves
};

#define K1T_TI 8
#define K1T_S_ROWS 36                 // rows j0-2 .. j0+33: the TMA unit wants the inner box coordinate 16-byte aligned (even row), see tools/probe
#define K1T_B_ROWS 36
// box offsets inside a stage: the TMA unit wants every destination 128-byte aligned
#define K1T_ALIGN128(x) (((x) + 127) / 128 * 128)
#define K1T_OFF_S 0
#define K1T_OFF_A K1T_ALIGN128(K1T_OFF_S + K1T_S_ROWS * (K1T_TI + 2) * 8)
#define K1T_OFF_B K1T_ALIGN128(K1T_OFF_A + 32 * (K1T_TI + 2) * 8)
#define K1T_OFF_BETA K1T_ALIGN128(K1T_OFF_B + K1T_B_ROWS * K1T_TI * 8)
#define K1T_OFF_X0 K1T_ALIGN128(K1T_OFF_BETA + 32 * K1T_TI * 8)                // Src
#define K1T_DS_ROWS 34                // dSdy of the eps equation: rows j0 .. j0+33, read at row j+1
#define K1T_OFF_X1 K1T_ALIGN128(K1T_OFF_X0 + 32 * K1T_TI * 8)                  // dSdy (eps) or F
#define K1T_OFF_X2 K1T_ALIGN128(K1T_OFF_X1 + K1T_DS_ROWS * K1T_TI * 8)          // F when dSdy is there too
#define K1T_STAGE_BYTES K1T_ALIGN128(K1T_OFF_X2 + 32 * K1T_TI * 8)             // 16128: every optional box present
#define K1T_STAGE_BYTES_SMALL K1T_OFF_X0                                       // 9856: S, A, B, beta only (equations 0..6 of a flat grid without sources)
#ifndef K1T_WARPS
#define K1T_WARPS 4
#endif
#define K1T_THREADS (32 * K1T_WARPS)
// full stage: flat grids have no F box (14208 B), axisymmetric ones do (16128 B); four-box stage: 9856 B flat, + the F box
// in the Src slot on axisymmetric grids (11904 B)
#define K1T_STAGE_OF(big, axi) ((big) ? ((axi) ? K1T_STAGE_BYTES : K1T_OFF_X2) : ((axi) ? K1T_OFF_X1 : K1T_STAGE_BYTES_SMALL))
#define K1T_SMEM_OF(big, axi) (K1T_WARPS * K1T_STAGE_OF(big, axi) + 64)
// Two launch shapes.  The stage of a warp is what limits residency (bytes in flight AND warps to cover the FP64 dependent-issue
// latency of the blending-factor divisions): 7 of the 9 equations of a flat k-eps grid need only the four basic boxes (9.9 KB),
// (plus the F box on axisymmetric grids: 11.9 KB, 4 blocks), so they run in a launch of their own with 5 blocks = 20 warps per SM and ONE code path for all seven (the equation index
// is a run-time plane offset: a third of the instruction footprint, the per-k switch had 13 % instruction-cache stalls);
// the equations with Src / dSdy / F boxes keep the larger stage (14 KB flat: 4 blocks = 16 warps per SM; 16 KB axisymmetric: 3 blocks) and the compile-time switch.
#ifndef K1T_MINB_BIG
#define K1T_MINB_BIG(axi) ((axi) ? 3 : 4)
#endif
#ifndef K1T_MINB_SMALL
#define K1T_MINB_SMALL(axi) ((axi) ? 4 : 5)
#endif
#ifndef K1T_BATCH
#define K1T_BATCH 4               // columns whose blending-factor divisions run as one branch-free block
#endif

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(map)), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

// the arithmetic of one staged item, equation index compile-time (as k1a_item, operands from shared memory)
// (KT = -1: equations 0..6 of a grid without Src / F planes -- identical flags, the index kk is a run-time plane offset)
template <class KC, int KT>
__device__ __forceinline__ void k1t_item(const Params& P, const StepState& st, const unsigned char* __restrict__ stage, int lane, int j, long long n_first,
                                         int want_rms, K1aAcc& acc, int kk) {
    constexpr int KF = KT < 0 ? 0 : KT;             // flags of equations 0..6 of an unflagged node are all "advance | ddx | ddy"
    const int K = KT < 0 ? kk : KT;
    constexpr uint32_t ct = CT_NODE_IS_SET, tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD;
    const long long ny = P.ny, N = P.N;
    const double dt = st.dt;
    const double dtdx = ddiv(dt, P.dx), dtdy = ddiv(dt, P.dy);                  // :843-844
    const double dyy = ddiv(P.dx, P.dx + P.dy), dxx = ddiv(P.dy, P.dx + P.dy);   // :643-644
    double* __restrict__ betak = P.beta + (long long)K * N;
    double* __restrict__ SNk = P.SN + (long long)K * N;
    constexpr unsigned f = eq_flags_c(ct, tt, KF, KC::ns);
    static_assert(KT >= 0 || f == 7u, "run-time equation index: plain flags");
    const bool has_src = KT < 0 ? false : (KT >= 7 || P.has_src);
    // (the run-time-index launch is only used when no dS plane of equations 0..6 is tracked: hf2d_step.inc)
    const bool trk_x = KT < 0 ? false : ((P.need_dsx >> KF) & 1u) != 0, trk_y = KT < 0 ? false : ((P.need_dsy >> KF) & 1u) != 0;
    constexpr bool DY2 = (f & 16) != 0;
    const double* __restrict__ sS = reinterpret_cast<const double*>(stage + K1T_OFF_S) + lane;      // row j-2 of column i0-1
    const double* __restrict__ sA = reinterpret_cast<const double*>(stage + K1T_OFF_A) + lane;      // row j   of column i0-1
    const double* __restrict__ sB = reinterpret_cast<const double*>(stage + K1T_OFF_B) + lane;      // row j-2 of column i0
    const double* __restrict__ sBe = reinterpret_cast<const double*>(stage + K1T_OFF_BETA) + lane;
    const double* __restrict__ sSrc = reinterpret_cast<const double*>(stage + K1T_OFF_X0) + lane;
    const double* __restrict__ sDs = reinterpret_cast<const double*>(stage + K1T_OFF_X1) + lane;
    const double* __restrict__ sF = reinterpret_cast<const double*>(stage + (KT < 0 ? K1T_OFF_X0 : DY2 ? K1T_OFF_X2 : K1T_OFF_X1)) + lane;   // (four-box stage: F sits in the Src slot)
    // Columns in batches of K1T_BATCH: first every column's Stage-1 value (straight-line, independent), then the residual /
    // blending-factor divisions of the whole batch as ONE branch-free block with a single validity vote behind it, so that
    // the 2 divisions + 1 square root of K1T_BATCH columns overlap (each is a dependent chain of ~12 FP64 instructions;
    // with 3 warps per scheduler resident a column-by-column vote left the FP64 pipe waiting on one chain at a time).
    const double beta_min = OMIN(P.beta0, st.beta_scn);
    const double bb = beta_min * beta_min;
#pragma unroll
    for (int cb = 0; cb < K1T_TI; cb += K1T_BATCH) {
        double scv[K1T_BATCH], vv[K1T_BATCH], bev[K1T_BATCH];
        bool any_res = false;
#pragma unroll
        for (int q = 0; q < K1T_BATCH; q++) {
            const int c = cb + q;
            const long long n = n_first + (long long)c * ny;
            const double sc = sS[(c + 1) * K1T_S_ROWS + 2], sl = sS[c * K1T_S_ROWS + 2], srr = sS[(c + 2) * K1T_S_ROWS + 2];
            const double su = sS[(c + 1) * K1T_S_ROWS + 3], sd = sS[(c + 1) * K1T_S_ROWS + 1];
            const double bu = sB[c * K1T_B_ROWS + 3], bd = sB[c * K1T_B_ROWS + 1];
            const double beta = sBe[c * 32], _beta = 1. - beta;
            double dXX = 0., dYY = 0.;
            if (f & 2) dXX = (sA[(c + 2) * 32] - sA[c * 32]) * 0.5;
            if (f & 4) dYY = (bu - bd) * 0.5;
            if (trk_x) P.dSx_out[KF][n] = dXX;
            if (trk_y) P.dSy_out[KF][n] = dYY;
            if (DY2) {   // :1011-1013: up = last step's dSdy, down = this sweep's (the row below is an unflagged interior node too)
                const double dd = (sB[c * K1T_B_ROWS + 2] - sB[c * K1T_B_ROWS]) * 0.5;
                dYY = (sDs[c * K1T_DS_ROWS + 1] + dd) * 0.5;
            }
            const double src = has_src ? sSrc[c * 32] : 0.;
            double v;
            if (KC::axi) v = sc * beta + _beta * (dxx * (sl + srr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * (dYY + ddiv(sF[c * 32], (double)(j + 1)))) + src * dt + 0.;
            else v = sc * beta + _beta * (dxx * (sl + srr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * dYY) + src * dt + 0.;
            SNk[n] = v;
            scv[q] = sc; vv[q] = v; bev[q] = beta;
            any_res |= sc != 0.;
        }
        // residual + blending factor (:1062-1154)
        if (!__any_sync(0xffffffffu, any_res)) continue;
        double DDv[K1T_BATCH], newb[K1T_BATCH];
        auto ph = [&](auto& dv) {
#pragma unroll
            for (int q = 0; q < K1T_BATCH; q++) {
                const bool res = scv[q] != 0.;
                double sqrt_RES;
                DDv[q] = residual_dd(dv, scv[q], vv[q], res, sqrt_RES, P.bFF);
                const double xx = (P.bFF >= 4) ? sqrt_RES : (P.bFF >= 2) ? DDv[q] * DDv[q] : DDv[q];
                const double cap = (P.bFF & 1) ? (beta_min + bev[q]) * 0.5 : beta_min;
                const double quo = dv.div(bb, beta_min + xx, res);
                newb[q] = OMIN(cap, quo);
            }
        };
        if (want_rms) { DivExact e; ph(e); }
        else { DivFast fdv; ph(fdv); if (__any_sync(0xffffffffu, fdv.is_bad())) { DivExact e; ph(e); } }
#pragma unroll
        for (int q = 0; q < K1T_BATCH; q++) {
            const bool res = scv[q] != 0.;
            const long long n = n_first + (long long)(cb + q) * ny;
            if (res) betak[n] = newb[q];
            if (want_rms && res) {
                const double sc = scv[q], v = vv[q], DD = DDv[q];
                if (P.isAlternateRMS) { acc.div += sc * sc; acc.rms += v - sc; } else { acc.rms += DD * DD; }
                acc.cnt++;
                if (DD != DD) atomicMax(&P.st->nan_last[K], (unsigned long long)n + 1ull);
                else { const unsigned long long d = d2bits(DD); acc.dd = d > acc.dd ? d : acc.dd; }
            }
        }
    }
}

// BIG = false: equations [k_first, k_first + k_count) within 0..6, four-box stage, run-time equation index;
// BIG = true:  any equations, full stage, compile-time switch.
template <class KC, bool BIG>
__global__ void __launch_bounds__(K1T_THREADS, BIG ? K1T_MINB_BIG(KC::axi) : K1T_MINB_SMALL(KC::axi))
k_stage1_tma(Params P, const K1aMaps* __restrict__ Mp, const int* __restrict__ tiles, int n_tiles, int n_sb, int s_half, int ds_half, int want_rms,
             int k_first, int k_count) {
    static_assert(KC::plain, "plain tiles only");
    constexpr int NE_C = !KC::ns ? 7 : (KC::ttc & TCT_Prandtl_Model) ? 7 : (KC::ttc & TCT_k_eps_Model) ? 9 : (KC::ttc & TCT_SA_Model) ? 8 : 7;
    constexpr uint32_t ct = CT_NODE_IS_SET, tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD;
    constexpr int STAGE = K1T_STAGE_OF(BIG, KC::axi);
    extern __shared__ __align__(128) unsigned char k1t_smem[];
    const StepState st = *P.st;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* stage = k1t_smem + (size_t)warp * STAGE;
    const unsigned bar = smem_u32(k1t_smem + (size_t)K1T_WARPS * STAGE + 8 * warp);
    // item = (k - k_first) * n_tiles + t: the warps of a block work on vertically adjacent tiles of the same plane set
    const long long item = (long long)blockIdx.x * K1T_WARPS + warp;
    const bool in_range = item < (long long)n_tiles * k_count;
    const int kq = in_range ? (int)(item / n_tiles) : 0;
    const int k = k_first + kq;
    const int tile = in_range ? tiles[(int)(item - (long long)kq * n_tiles)] : 0;
    const int tx = tile / n_sb, sb = tile - tx * n_sb;
    const int j0 = sb * TILE_H, j = j0 + lane;
    const int i0 = P.i_lo + tx * K1T_TI;
    const long long n_first = (long long)i0 * P.ny + j;
    K1aAcc acc; acc.rms = 0.; acc.div = 0.; acc.cnt = 0u; acc.dd = 0ull;
    if (in_range) {
        if (lane == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            const bool has_src = BIG && (k >= 7 || P.has_src);
            const bool dy2 = BIG && (eq_flags(ct, tt, k, KC::ns) & 16) != 0;
            unsigned bytes = (K1T_S_ROWS * (K1T_TI + 2) + 32 * (K1T_TI + 2) + K1T_B_ROWS * K1T_TI + 32 * K1T_TI) * 8;
            if (has_src) bytes += 32 * K1T_TI * 8;
            if (dy2) bytes += K1T_DS_ROWS * K1T_TI * 8;
            if (KC::axi) bytes += 32 * K1T_TI * 8;
            mbar_expect_tx(bar, bytes);
            const unsigned sbase = smem_u32(stage);
            tma_load_3d(sbase + K1T_OFF_S, &Mp->S[s_half], bar, j0 - 2, i0 - 1, k);
            tma_load_3d(sbase + K1T_OFF_A, &Mp->A, bar, j0, i0 - 1, k);
            tma_load_3d(sbase + K1T_OFF_B, &Mp->B, bar, j0 - 2, i0, k);
            tma_load_3d(sbase + K1T_OFF_BETA, &Mp->beta, bar, j0, i0, k);
            if constexpr (BIG) {
                if (has_src) tma_load_3d(sbase + K1T_OFF_X0, &Mp->Src, bar, j0, i0, k);
                if (dy2) tma_load_2d(sbase + K1T_OFF_X1, &Mp->dSy[ds_half], bar, j0, i0);
                if (KC::axi) tma_load_3d(sbase + (dy2 ? K1T_OFF_X2 : K1T_OFF_X1), &Mp->F, bar, j0, i0, k);
            } else if (KC::axi) tma_load_3d(sbase + K1T_OFF_X0, &Mp->F, bar, j0, i0, k);
        }
        __syncwarp();
        mbar_wait(bar, 0);
        if constexpr (!BIG) {
            k1t_item<KC, -1>(P, st, stage, lane, j, n_first, want_rms, acc, k);
        } else {
            switch (k) {
            case 0: k1t_item<KC, 0>(P, st, stage, lane, j, n_first, want_rms, acc, 0); break;
            case 1: k1t_item<KC, 1>(P, st, stage, lane, j, n_first, want_rms, acc, 1); break;
            case 2: k1t_item<KC, 2>(P, st, stage, lane, j, n_first, want_rms, acc, 2); break;
            case 3: k1t_item<KC, 3>(P, st, stage, lane, j, n_first, want_rms, acc, 3); break;
            case 4: k1t_item<KC, 4>(P, st, stage, lane, j, n_first, want_rms, acc, 4); break;
            case 5: k1t_item<KC, 5>(P, st, stage, lane, j, n_first, want_rms, acc, 5); break;
            case 6: k1t_item<KC, 6>(P, st, stage, lane, j, n_first, want_rms, acc, 6); break;
            case 7: if constexpr (NE_C > 7) k1t_item<KC, 7>(P, st, stage, lane, j, n_first, want_rms, acc, 7); break;
            case 8: if constexpr (NE_C > 8) k1t_item<KC, 8>(P, st, stage, lane, j, n_first, want_rms, acc, 8); break;
            }
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
