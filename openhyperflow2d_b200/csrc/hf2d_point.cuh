// hf2d_point.cuh -- kernel K1 of the DEEPS2D step: everything of one step that needs only the node's own NEW
// conserved state plus last step's fields of its four neighbours:
//   Stage 1 (deeps2d_core.cpp:853-1026: flux differences, soft BCs, beta-blended update into NextNode->S),
//   residual + locally adaptive blending factor (:1062-1154), commit (:1156-1164),
//   the own-node half of FillNode2D (hyper_flow_node.hpp:382-492: k, U, V, wall rewrites, p, Tg, the S-rewriting part
//   of TurbModRANS2D), CalcChemicalReactions (deeps2d_core.cpp:4697-4780) and the CFL dt (:1318-1326).
//
// Why this cut.  In the reference's in-place sweep node C reads its LEFT and DOWN neighbours after they went through
// commit + FillNode2D + chemistry in this step and its RIGHT and UP neighbours before.  What a neighbour
// exposes (S[0], S[4..8], U, V, Tg) depends only on that neighbour's own shadow S, flags and lagged CP/R -- never on
// gradients.  So the step splits into two embarrassingly parallel passes over ping-pong planes:
//   K1 (here):       every node's new state  -> *_out planes       (reads only *_in planes, A, B, beta)
//   K2 (hf2d_flux):  gradients from *_out (left, down) and *_in (right, up), then fluxes / turbulence sources.
// No thread waits for another, nothing is recomputed, and the division-heavy pointwise arithmetic (blending factor,
// equation of state, chemistry, 12 property tables, dt) sits in the same kernel as Stage 1's memory traffic, under
// whose latency it hides.
//
// Work decomposition: a warp owns a tile of TI columns x 32 rows (y is the contiguous index: one 256-byte segment per
// plane and column) and walks the columns in +x.  `plain` tiles (every node of the tile and of its neighbour ring is an
// unflagged interior node) run instantiations in which the flag words are compile-time constants.
#pragma once
#include <type_traits>
#include "hf2d_turb.cuh"
#include "hf2d_stage1.cuh"

#define TILE_H 32                 // rows per tile = lanes per warp
// Block shape.  The general-tile launch runs on the side stream BESIDE K1a (k_stage1_tma), whose blocks hold 61 K of the
// SM's 64 K registers: what the general blocks cost is the K1a blocks they displace.  Measured on the bench grid
// (ms/step; K1a's share): 4 warps x 3 blocks, 156 registers: 2.010 (0.858); 8 warps, plain 2 blocks / general 3 blocks
// at 80 registers: 1.935 (0.777); general 4 blocks at 64 registers (spills to L1): 1.929 (0.767); 16 warps: 1.960.
#ifndef K1_WARPS
#define K1_WARPS 8
#endif
#define K1_THREADS (32 * K1_WARPS)
#ifndef K1_MINB_PLAIN
#define K1_MINB_PLAIN 2
#endif
#ifndef K1_MINB_GEN
#define K1_MINB_GEN 4
#endif

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
// max / min over the warp of the bit patterns of non-negative doubles (order preserving): two 32-bit REDUX steps
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
    const unsigned hi = (unsigned)(v >> 32), mh = __reduce_max_sync(0xffffffffu, hi);
    const unsigned lo = (hi == mh) ? (unsigned)v : 0u, ml = __reduce_max_sync(0xffffffffu, lo);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
    const unsigned hi = (unsigned)(v >> 32), mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned lo = (hi == mh) ? (unsigned)v : 0xffffffffu, ml = __reduce_min_sync(0xffffffffu, lo);
    return ((unsigned long long)mh << 32) | ml;
}

// residual of one (node, equation) pair: DD_local (:1089-1100).  Shared with k_ddmax_fix.
template <class D>
__device__ __forceinline__ double residual_dd(D& dv, double s_old, double s_next, bool need, double& sqrt_res, int bFF) {
    const double absDD = s_next - s_old;
    const bool big = fabs(s_old) > 1.e-15;
    const double q = dv.div(absDD, need && big ? s_old : 1.0, need && big);
    const double DD = big ? fabs(q) : 1.0;
    sqrt_res = 0.;
    if (bFF >= 4) { const double r_ = dv.sqrt(DD, need && big); sqrt_res = big ? r_ : 0.; }
    return DD;
}

template <class KC>
__global__ void __launch_bounds__(K1_THREADS, KC::plain ? K1_MINB_PLAIN : K1_MINB_GEN)
k_point(Params P, const int* __restrict__ tiles, int n_tiles, int n_sb, int TI, int want_rms) {
    __shared__ double sh_acc[K1_WARPS][2 * NEQ];               // RMS, sumDiv partial sums of the warp (want_rms steps)
    __shared__ unsigned long long sh_u[K1_WARPS][2 * NEQ + 2]; // counts, max DD, dt-min, first error
    __shared__ SmemTables sh_tab;
    if (S2_SMEM_TABLES) smem_tables_fill(sh_tab, P.tab);
    const StepState st = *P.st;   // snapshot (the accumulators are only touched through atomics)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // plain tiles: a warp walks the TI columns of its tile; general tiles (few, latency-bound): one COLUMN per warp, so that
    // the launch is TI times wider and its critical path TI times shorter (nothing is carried from column to column)
    const int w_idx = blockIdx.x * K1_WARPS + warp;
    const int t_idx = KC::plain ? w_idx : w_idx / TI;
    const int tile = t_idx < n_tiles ? tiles[t_idx] : 0;
    const int tx = tile / n_sb, sb = tile - tx * n_sb;
    const int j = sb * TILE_H + lane;
    const int i0 = P.i_lo + tx * TI + (KC::plain ? 0 : w_idx - t_idx * TI);
    const int i1 = t_idx < n_tiles ? min(KC::plain ? i0 + TI : i0 + 1, P.i_hi) : i0;
    const bool valid_j = j < P.ny;
    const long long ny = P.ny, N = P.N;
    if (threadIdx.x < K1_WARPS * 2 * NEQ) (&sh_acc[0][0])[threadIdx.x] = 0.;
    if (lane < 2 * NEQ) sh_u[warp][lane] = 0ull;
    __syncthreads();
    unsigned long long r_dt = ~0ull, r_err = ~0ull;
    int trow[12];                  // row of the last look-up in each property table (tab_get_c)
#pragma unroll
    for (int q = 0; q < 12; q++) trow[q] = 1;
    const double dt = st.dt;
    const double dtdx = ddiv(dt, P.dx), dtdy = ddiv(dt, P.dy);                  // :843-844
    const double dyy = ddiv(P.dx, P.dx + P.dy), dxx = ddiv(P.dy, P.dx + P.dy);   // :643-644

    // A phase is run with the branch-free DivFast policy on plain tiles and re-run exactly when any lane asks for it
    auto run = [&](auto&& ph) {
        if constexpr (KC::plain) {
            DivFast f; ph(f);
            if (__any_sync(0xffffffffu, f.is_bad())) { DivExact e; ph(e); }
        } else { DivExact e; ph(e); }
    };
    constexpr int NE_C = KC::plain ? (!KC::ns ? 7 : (KC::ttc & TCT_Prandtl_Model) ? 7 : (KC::ttc & TCT_k_eps_Model) ? 9 : (KC::ttc & TCT_SA_Model) ? 8 : 7) : 0;
    constexpr bool TURB_C = KC::plain && KC::ns && (KC::ttc & (TCT_k_eps_Model | TCT_SA_Model)) != 0;

    bool tile_same = true;         // plain tiles: chemistry left S[4..6] of every node visited so far unchanged (Params::tile_inert)
    for (int i = i0; i < i1; i++) {
        const long long n = valid_j ? (long long)i * ny + j : 0;
        uint32_t ct, tt;
        if (KC::plain) { ct = CT_NODE_IS_SET; tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD; }
        else { ct = 0; tt = 0; if (valid_j) { ct = P.ct[n]; tt = P.tt[n]; } }
        const bool act = valid_j && node_active(ct);
        const bool fc = valid_j && !act && is_set(ct, NT_FC) && !is_set(ct, CT_SOLID);
        const bool upd = act || fc;
        const int ne = act ? num_eq(tt) : 0;
        const bool turb2 = KC::ns && (is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model));
        const bool hasturb = KC::ns && (tt & (TCT_k_eps_Model | TCT_SA_Model));
        SpecNode* sp = (!KC::plain && valid_j && (tt & TT_SPEC)) ? &P.spec[P.spec_idx[n]] : nullptr;

        // ================= Stage 1 (:853-1026): NextNode->S of this node =================
        double sn[NEQ];
        unsigned adv_mask = 0;          // bit k: equation k advanced by Stage 1 this step (sn[k] holds NextNode->S[k])
#pragma unroll
        for (int k = 0; k < NEQ; k++) sn[k] = 0.;
        // tracked dS planes must be carried over for nodes / equations that are not advanced
        if (!KC::plain && valid_j && (P.need_dsx | P.need_dsy)) {
            for (int k = 0; k < NEQ; k++) {
                const bool adv = act && k < ne && (eq_flags(ct, tt, k, KC::ns) & 1);
                if (!adv) {
                    if (P.need_dsx & (1u << k)) P.dSx_out[k][n] = P.dSx_in[k][n];
                    if (P.need_dsy & (1u << k)) P.dSy_out[k][n] = P.dSy_in[k][n];
                }
            }
        }
        if constexpr (KC::plain) {
            // plain tiles: K1a (k_stage1_plain) has written NextNode->S and the new blending factors, equation by equation
#pragma unroll
            for (int k = 0; k < NE_C; k++) if (valid_j) sn[k] = P.SN[(long long)k * N + n];
#if K_PREFETCH
            if (i + 1 < i1) {          // the next column's operands, requested under this column's arithmetic
                const long long m = n + ny;
#pragma unroll
                for (int k = 0; k < NE_C; k++) pf(&P.SN[(long long)k * N + m]);
                if (NE_C < 9 && TURB_C) pf(&P.S_in[8 * N + m]);
                pf(&P.CP_in[m]); pf(&P.R_in[m]); pf(&P.kg_in[m]);
                if (!K1B_LAZY) { pf(&P.U_in[m]); pf(&P.V_in[m]); pf(&P.Tg_in[m]); pf(&P.p_in[m]); }
                if (KC::ns) { pf(&P.mu_in[m]); if (KC::ttc & TCT_k_eps_Model) pf(&P.l_min[m]); }
            }
#endif
        } else if (act) {
            const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0, n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
            const long long nL = n - n1 * ny, nR = n + n2 * ny, nU = n + n3, nD = n - n4;
            const double n_n_1 = (n1 + n2 == 2) ? 0.5 : 1., m_m_1 = (n3 + n4 == 2) ? 0.5 : 1.;   // 1/max(n1+n2,1), 1/max(n3+n4,1)
            uint32_t ttR = 0u, ttU = 0u;
            if (!KC::plain) { ttR = n2 ? P.tt[nR] : 0u; ttU = n3 ? P.tt[nU] : 0u; }
#pragma unroll
            for (int k = 0; k < NEQ; k++) {
                if (k >= ne) continue;
                const unsigned f = eq_flags(ct, tt, k, KC::ns);
                if (!(f & 1)) continue;
                const long long o = (long long)k * N;
                const double beta = P.beta[o + n], _beta = 1. - beta;
                const double sc = P.S_in[o + n];
                const double sl = P.S_in[o + nL], sd = P.S_in[o + nD];
                const double sr = KC::plain ? P.S_in[o + nR] : (n2 ? s_old(P, nR, ttR, k) : sc);
                const double su = KC::plain ? P.S_in[o + nU] : (n3 ? s_old(P, nU, ttU, k) : sc);
                double dXX, dYY;
                if (f & 2) dXX = (P.A[o + nR] - P.A[o + nL]) * n_n_1; else dXX = 0.;
                if (f & 4) dYY = (P.B[o + nU] - P.B[o + nD]) * m_m_1; else dYY = 0.;
                if (P.need_dsx & (1u << k)) P.dSx_out[k][n] = dXX;
                if (P.need_dsy & (1u << k)) P.dSy_out[k][n] = dYY;
                if (f & 8) {    // :1008-1010  left: this sweep's value, right: last step's (own value if aliased)
                    const double dl = n1 ? ds_new<KC>(P, nL, k, 0) : dXX;
                    const double dr = n2 ? P.dSx_in[k][nR] : dXX;
                    dXX = (dl + dr) * 0.5;
                }
                if (f & 16) {   // :1011-1013
                    const double du = n3 ? P.dSy_in[k][nU] : dYY;
                    const double dd = n4 ? ds_new<KC>(P, nD, k, 1) : dYY;
                    dYY = (du + dd) * 0.5;
                }
                double src = 0.;
                if (k >= 7 || P.has_src) src = P.Src[o + n];
                const double srcadd = sp ? sp->SrcAdd[k] : 0.;
                double v;
                if (KC::axi) {
                    v = sc * beta + _beta * (dxx * (sl + sr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * (dYY + ddiv(P.F[o + n], (double)(j + 1)))) + src * dt + srcadd;
                } else {
                    v = sc * beta + _beta * (dxx * (sl + sr) + dyy * (su + sd)) * 0.5 - (dtdx * dXX + dtdy * dYY) + src * dt + srcadd;
                }
                sn[k] = v; adv_mask |= 1u << k;
                P.SN[o + n] = v;
            }
        }

        // ====== residual + blending factor (:1062-1154) fused with the commit (:1156-1164) ======
        Local L;
#pragma unroll
        for (int k = 0; k < NEQ; k++) L.S[k] = 0.;
        if constexpr (KC::plain) {
            // every equation k < NE_C is advanced and committed (K1a did the residual and the blending factor);
            // S[7], S[8] of a laminar / algebraic-model node stay what they are
#pragma unroll
            for (int k = 0; k < NEQ; k++) L.S[k] = !valid_j ? 0. : (k < NE_C) ? sn[k] : (k < 7 || TURB_C) ? P.S_in[(long long)k * N + n] : 0.;
        } else {
            DivExact dv;
#pragma unroll
            for (int k = 0; k < NEQ; k++) {
                double s = 0., c_rms = 0., c_div = 0.;
                unsigned long long c_dd = 0ull;
                unsigned c_cnt = 0u;
                bool isnan_dd = false;
                if (valid_j && (k < 7 || hasturb)) {
                    const long long o = (long long)k * N + n;
                    s = P.S_in[o];
                    if (act && k < ne) {
                        uint32_t c_flag = 0;
                        if (k < 4) c_flag = CT_Rho_CONST << k;
                        else if (k < 4 + NCOMP) c_flag = CT_Y_CONST;
                        else if (turb2) c_flag = TCT_k_CONST << (k - 4 - NCOMP);
                        const bool res = !is_set(ct, c_flag) && s != 0.;   // quirk kept: CT word even for k >= 7 (:1075)
                        const bool com = eq_committed(ct, tt, k, KC::ns);
                        double snk = sn[k];
                        if ((res || com) && !(adv_mask & (1u << k))) snk = P.SN[o];   // the shadow as an earlier step left it
                        if (res) {
                            const double Tmp = s;
                            const double absDD = snk - s;
                            double DD, sqrt_RES = 0;
                            if (fabs(Tmp) > 1.e-15) { DD = fabs(dv.div(absDD, Tmp)); if (P.bFF >= 4) sqrt_RES = dv.sqrt(DD); } else { DD = 1.0; }
                            double beta_min = OMIN(P.beta0, st.beta_scn);
                            if (is_set(ct, CT_NONREFLECTED)) beta_min = P.nrbc_beta0;
                            double* bp = &P.beta[o];
                            const double bb = beta_min * beta_min;
                            // the six blending-factor functions (deeps2d_core.cpp:1104-1121): min(cap, bb / (beta_min + x))
                            const double xx = (P.bFF >= 4) ? sqrt_RES : (P.bFF >= 2) ? DD * DD : DD;
                            const double cap = (P.bFF & 1) ? (beta_min + *bp) * 0.5 : beta_min;
                            const double quo = dv.div(bb, beta_min + xx);
                            *bp = OMIN(cap, quo);
                            if (P.isAlternateRMS) { c_div = Tmp * Tmp; c_rms = absDD; } else { c_rms = DD * DD; }
                            c_cnt = 1u;
                            isnan_dd = DD != DD;
                            c_dd = isnan_dd ? 0ull : d2bits(DD);
                        }
                        if (com) s = snk;
                    }
                }
                L.S[k] = s;
                if (want_rms) {   // uniform: residual norms are only read on output steps (deeps2d_core.cpp:1602-1637)
                    const double a = warp_sum(c_rms), b = warp_sum(c_div);
                    const unsigned c = __reduce_add_sync(0xffffffffu, c_cnt);
                    const unsigned long long d = warp_max_u64(c_dd);
                    if (isnan_dd) atomicMax(&P.st->nan_last[k], (unsigned long long)n + 1ull);
                    if (lane == 0 && c) {
                        sh_acc[warp][k] += a; sh_acc[warp][NEQ + k] += b; sh_u[warp][k] += c;
                        if (d > sh_u[warp][NEQ + k]) sh_u[warp][NEQ + k] = d;
                    }
                }
            }
        }

        // ====== own-node half of FillNode2D + CalcChemicalReactions + dt: the node's new state ======
        const GlobalLoad ld{P, n};
        const unsigned long long r_dt0 = r_dt, r_err0 = r_err;
        bool col_same = true;
        run([&](auto& dv) {
            r_dt = r_dt0; r_err = r_err0;
            if (!upd) return;
            local_rest<KC, true>(dv, ld, P, st, n, ct, tt, L);
            // ---- FillNode2D's wall handling of the virtual sources (hyper_flow_node.hpp:447-488): local data only ----
            if (!KC::plain && sp && L.filled) {
                if (is_set(ct, CT_WALL_LAW)) {
                } else if (is_set(ct, CT_WALL_NO_SLIP)) {
                    double sa0 = sp->SrcAdd[iRho];
                    if (st.is_src_add) {
                        sa0 = sp->BGX * (L.Upre - sp->Uw) * L.S[iRho] / P.dx + sp->BGY * (L.Vpre - sp->Vw) * L.S[iRho] / P.dy;
                        sp->SrcAdd[iRho] = sa0;
                        sp->SrcAdd[iRhoU] = sp->BGX * (L.Upre - sp->Uw) * L.S[iRho];
                        sp->SrcAdd[iRhoV] = sp->BGY * (L.Vpre - sp->Vw) * L.S[iRho];
                        sp->SrcAdd[iRhoE] = 0.;
                    } else {
                        sp->SrcAdd[iRhoU] = 0.; sp->SrcAdd[iRhoV] = 0.; sp->SrcAdd[iRhoE] = 0.;
                    }
                    for (int q = 0; q < NCOMP; q++) sp->SrcAdd[4 + q] = st.is_src_add ? sa0 * sp->Yw[q] : 0.;   // Y of LAST step's chemistry
                } else {
                    for (int k = 0; k < NEQ; k++) sp->SrcAdd[k] = 0.;
                }
            }
            // ---- primitives / conserved state into the *_out planes ----
            double R_new = L.R, CP_new = L.CP, mu_new = L.mu;
            double lam_new = 0.; bool have_lam = false;
            double s4 = L.S[4], s5 = L.S[5], s6 = L.S[6];
            if (act) {
                if (L.Tg < 0.) {
                    { const unsigned long long e_ = (unsigned long long)n; if (e_ < r_err) r_err = e_; }                      // :1246  (the reference aborts here)
                } else {
                    const double CFL_min = OMIN(P.CFL, st.CFL_scn);
                    const double AAA = dv.sqrt(L.kg * L.R * L.Tg);
                    const double dtl = CFL_min * OMIN(dv.div(P.dx, AAA + fabs(L.U)), dv.div(P.dy, AAA + fabs(L.V)));
                    { const unsigned long long b_ = d2bits(dtl); if (b_ < r_dt) r_dt = b_; }
                    // CalcChemicalReactions: mixture properties at the new temperature (:4730-4746)
                    const DevTable* tb = P.tab;
                    R_new = P.Rgas[0] * L.Yfu + P.Rgas[1] * L.Yox + P.Rgas[2] * L.Ycp + P.Rgas[3] * L.Yair;
                    // A species whose mass fraction is exactly 0 contributes (+-)0 to the three sums below whatever its table
                    // says, so its lookups are skipped (identical result unless the whole sum is zero or Tg is not finite).
                    const bool h0 = L.Yfu != 0., h1 = L.Yox != 0., h2 = L.Ycp != 0., h3 = L.Yair != 0.;
                    CP_new = (h0 ? tab_get_c(dv, sh_tab, tb, 0, L.Tg, trow) : 0.) * L.Yfu + (h1 ? tab_get_c(dv, sh_tab, tb, 1, L.Tg, trow) : 0.) * L.Yox +
                             (h2 ? tab_get_c(dv, sh_tab, tb, 2, L.Tg, trow) : 0.) * L.Ycp + (h3 ? tab_get_c(dv, sh_tab, tb, 3, L.Tg, trow) : 0.) * L.Yair;
                    if (KC::ns) {
                        lam_new = (h0 ? tab_get_c(dv, sh_tab, tb, 4, L.Tg, trow) : 0.) * L.Yfu + (h1 ? tab_get_c(dv, sh_tab, tb, 5, L.Tg, trow) : 0.) * L.Yox +
                                  (h2 ? tab_get_c(dv, sh_tab, tb, 6, L.Tg, trow) : 0.) * L.Ycp + (h3 ? tab_get_c(dv, sh_tab, tb, 7, L.Tg, trow) : 0.) * L.Yair;
                        have_lam = true;
                        mu_new = (h0 ? tab_get_c(dv, sh_tab, tb, 8, L.Tg, trow) : 0.) * L.Yfu + (h1 ? tab_get_c(dv, sh_tab, tb, 9, L.Tg, trow) : 0.) * L.Yox +
                                 (h2 ? tab_get_c(dv, sh_tab, tb, 10, L.Tg, trow) : 0.) * L.Ycp + (h3 ? tab_get_c(dv, sh_tab, tb, 11, L.Tg, trow) : 0.) * L.Yair;
                    }
                    s4 = L.y4; s5 = L.y5; s6 = L.y6;
                    if (sp) { for (int q = 0; q < 4; q++) sp->Yw[q] = L.Y[q]; }
                }
            }
            P.S_out[0 * N + n] = L.S[0]; P.S_out[1 * N + n] = L.S[1]; P.S_out[2 * N + n] = L.S[2]; P.S_out[3 * N + n] = L.S[3];
            P.S_out[4 * N + n] = s4; P.S_out[5 * N + n] = s5; P.S_out[6 * N + n] = s6;
            if (KC::plain) col_same = d2bits(s4) == d2bits(sn[4]) && d2bits(s5) == d2bits(sn[5]) && d2bits(s6) == d2bits(sn[6]);
            if (KC::plain ? TURB_C : hasturb) { P.S_out[7 * N + n] = L.S[7]; P.S_out[8 * N + n] = L.S[8]; }
            P.U_out[n] = L.U; P.V_out[n] = L.V; P.Tg_out[n] = L.Tg;
            P.kg_out[n] = L.kg; P.p_out[n] = L.p;
            P.R_out[n] = R_new; P.CP_out[n] = CP_new;
            if (KC::ns) { P.mu_out[n] = mu_new; P.lam_out[n] = have_lam ? lam_new : P.lam_in[n]; }
        });
        tile_same = tile_same && col_same;
    }
    if constexpr (KC::plain) {
        const bool all_same = __all_sync(0xffffffffu, tile_same);
        if (lane == 0 && t_idx < n_tiles && P.tile_inert) P.tile_inert[tile] = all_same ? 1 : 0;
    }

    // ---- block reduction of the step's scalars: warp REDUX/shuffles -> shared memory -> at most one atomic per block
    // and quantity.  dt-min / max residual / first-error use integer atomics on the bit patterns of non-negative doubles
    // (order preserving), so they are exact; the residual sums are floating-point atomics (summation order differs from
    // the reference's sweep: monitors only).
    {
        const unsigned long long a = warp_min_u64(r_dt), b = warp_min_u64(r_err);
        if (lane == 0) { sh_u[warp][2 * NEQ] = a; sh_u[warp][2 * NEQ + 1] = b; }
    }
    __syncthreads();
    const int tid = threadIdx.x;
    if (tid == 0) {
        unsigned long long s = ~0ull;
        for (int w = 0; w < K1_WARPS; w++) s = sh_u[w][2 * NEQ] < s ? sh_u[w][2 * NEQ] : s;
        // st.dt_min_bits is the value at kernel start: under the serial rule the minimum only ever decreases, so a
        // block that cannot lower it skips the atomic
        if (s < st.dt_min_bits || (P.dt_rule == 1 && s != ~0ull)) atomicMin(&P.st->dt_min_bits, s);
    } else if (tid == 32) {
        unsigned long long s = ~0ull;
        for (int w = 0; w < K1_WARPS; w++) s = sh_u[w][2 * NEQ + 1] < s ? sh_u[w][2 * NEQ + 1] : s;
        if (s != ~0ull) atomicMin(&P.st->err_node, s);
    }
    if (want_rms) {
        if (tid < 2 * NEQ) {             // threads 0..8: RMS sums, 9..17: sumDiv
            double s = 0.;
            for (int w = 0; w < K1_WARPS; w++) s += sh_acc[w][tid];
            if (s != 0.) atomicAdd(tid < NEQ ? &P.st->RMS[tid] : &P.st->sumDiv[tid - NEQ], s);
        } else if (tid >= 32 && tid < 32 + NEQ) {
            const int k = tid - 32;
            unsigned long long s = 0;
            for (int w = 0; w < K1_WARPS; w++) s += sh_u[w][k];
            if (s) atomicAdd(&P.st->iRMS[k], s);
        } else if (tid >= 64 && tid < 64 + NEQ) {
            const int k = tid - 64;
            unsigned long long s = 0;
            for (int w = 0; w < K1_WARPS; w++) s = sh_u[w][NEQ + k] > s ? sh_u[w][NEQ + k] : s;
            if (s) atomicMax(&P.st->DDmax_bits[k], s);
        }
    }
}

// The reference keeps the largest residual with  DD_max = max(DD_max, DD_local)  where max(a,b) = (a > b) ? a : b
// (deeps2d_core.cpp:1147, hyper_flow_node.hpp:17-23): a NaN DD_local REPLACES the running maximum and the next node
// replaces the NaN again, so after a sweep that met NaNs the value is the maximum over the nodes visited AFTER the last
// NaN.  k_point keeps NaNs out of its maximum and records the last NaN node per equation (StepState::nan_last); this
// kernel, launched after k_point on the steps that reduce residuals, does nothing unless there was one, and then redoes
// the maximum of that equation over the nodes behind it (S_in and the shadow SN are both still in place).
__global__ void k_ddmax_fix(Params P) {
    __shared__ unsigned long long sh[32];
    DivExact dv;
    for (int k = 0; k < NEQ; k++) {
        const unsigned long long last = P.st->nan_last[k];
        if (last == 0ull) continue;
        unsigned long long m = 0ull;
        for (long long n = (long long)last + (long long)blockIdx.x * blockDim.x + threadIdx.x; n < P.N; n += (long long)gridDim.x * blockDim.x) {
            const int i = (int)(n / P.ny);
            if (i < P.i_lo || i >= P.i_hi) continue;
            const uint32_t ct = P.ct[n], tt = P.tt[n];
            if (!node_active(ct) || k >= num_eq(tt)) continue;
            uint32_t c_flag = 0;
            if (k < 4) c_flag = CT_Rho_CONST << k;
            else if (k < 4 + NCOMP) c_flag = CT_Y_CONST;
            else if (P.ns && (is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model))) c_flag = TCT_k_CONST << (k - 4 - NCOMP);
            const double s = P.S_in[(long long)k * P.N + n];
            if (is_set(ct, c_flag) || s == 0.) continue;
            double DD = 1.0;
            if (fabs(s) > 1.e-15) DD = fabs(dv.div(P.SN[(long long)k * P.N + n] - s, s));
            if (DD == DD) { const unsigned long long b = d2bits(DD) + 1ull; m = b > m ? b : m; }
        }
        m = warp_max_u64(m);
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long s = 0ull;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) s = sh[w] > s ? sh[w] : s;
            atomicMax(&P.st->DDfix_bits[k], s);
        }
        __syncthreads();
    }
}
