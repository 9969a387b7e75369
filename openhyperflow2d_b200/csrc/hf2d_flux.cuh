// hf2d_flux.cuh -- kernel K2 of the DEEPS2D step: the gradient-dependent half of Stage 2 (deeps2d_core.cpp:1166-1244):
// mixed old/new gradients (:1169-1238), TurbModRANS2D (hyper_flow_node.hpp:602-957), the viscous / inviscid fluxes
// A, B, F and the turbulence sources of FillNode2D (:423-599).
//
// K1 (hf2d_point.cuh) has already written every node's new state into the *_out planes.  In the reference's in-place
// sweep node C sees its LEFT and DOWN neighbours after their update of this step and its RIGHT and UP neighbours
// before:        left, down  <- *_out planes          right, up  <- *_in planes
// (S_in of a Neumann node holds the Stage-1 in-place overwrite, which is what the reference's Stage 2 reads there).
// The node's own inputs: committed S after FillNode2D's rewrites = S_out, except the species densities, which FillNode2D
// uses BEFORE CalcChemicalReactions rewrote them (= the shadow SN where the equation was committed); U, V, Tg, p, k =
// *_out; the lagged properties CP, R, mu, lam (last step's chemistry) = *_in.
//
// Same tiles as K1: a warp owns TI columns x 32 rows and walks +x.  The DOWN neighbour's new state is what lane-1 just
// loaded as its own (one shuffle; lane 0 reads the row below the tile), the LEFT neighbour's new state is what this
// thread loaded as its own one column earlier (registers; the first column reads it).
#pragma once
#include "hf2d_point.cuh"

#ifndef K2_WARPS
#define K2_WARPS 4
#endif
#define K2_THREADS (32 * K2_WARPS)
#ifndef K2_MINB_PLAIN
#define K2_MINB_PLAIN 3
#endif
#ifndef K2_INERT
#define K2_INERT 1                // 0: always read the species densities before chemistry from the shadow planes
#endif
#ifndef K2_MINB_GEN
#define K2_MINB_GEN 3
#endif

struct NbState { double rho, y4, y5, y6, s7, s8, U, V, Tg; };

__device__ __forceinline__ void nb_load(const double* __restrict__ S, const double* __restrict__ U, const double* __restrict__ V,
                                        const double* __restrict__ Tg, long long N, long long m, bool turb, NbState& s) {
    s.rho = S[m]; s.y4 = S[4 * N + m]; s.y5 = S[5 * N + m]; s.y6 = S[6 * N + m];
    s.s7 = turb ? S[7 * N + m] : 0.; s.s8 = turb ? S[8 * N + m] : 0.;
    s.U = U[m]; s.V = V[m]; s.Tg = Tg[m];
}
__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

// gradient planes kept on request (hf2d_gpu_keep_gradients): slot order = SpecNode::grad
enum { G_droYdx = 0, G_droYdy = 4, G_dUdx = 8, G_dUdy, G_dVdx, G_dVdy, G_dTdx, G_dTdy, G_dkdx, G_dkdy, G_depsdx, G_depsdy, G_COUNT };

template <class KC>
__global__ void __launch_bounds__(K2_THREADS, KC::plain ? K2_MINB_PLAIN : K2_MINB_GEN)
k_flux(Params P, const int* __restrict__ tiles, int n_tiles, int n_sb, int TI, int want_grads) {
    const StepState st0 = *P.st;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // plain tiles: a warp walks the TI columns of its tile; general tiles: one column per warp (see k_point)
    const int w_idx = blockIdx.x * K2_WARPS + warp;
    const int t_idx = KC::plain ? w_idx : w_idx / TI;
    if (t_idx >= n_tiles) return;
    const int tile = tiles[t_idx];
    const int tx = tile / n_sb, sb = tile - tx * n_sb;
    const int j = sb * TILE_H + lane;
    const int i0 = P.i_lo + tx * TI + (KC::plain ? 0 : w_idx - t_idx * TI);
    const int i1 = min(KC::plain ? i0 + TI : i0 + 1, P.i_hi);
    const bool valid_j = j < P.ny;
    const long long ny = P.ny, N = P.N;
    const double yj = (KC::axi && valid_j) ? P.yrow[j] : 1.0;
    constexpr bool TURB_C = KC::plain && KC::ns && (KC::ttc & (TCT_k_eps_Model | TCT_SA_Model)) != 0;
    constexpr int NE_C = KC::plain ? (!KC::ns ? 7 : (KC::ttc & TCT_Prandtl_Model) ? 7 : (KC::ttc & TCT_k_eps_Model) ? 9 : (KC::ttc & TCT_SA_Model) ? 8 : 7) : 0;
    (void)NE_C;

    auto run = [&](auto&& ph) {
        if constexpr (KC::plain) {
            DivFast f; ph(f);
            if (__any_sync(0xffffffffu, f.is_bad())) { DivExact e; ph(e); }
        } else { DivExact e; ph(e); }
    };

    const bool inert = K2_INERT && KC::plain && P.tile_inert && P.tile_inert[tile] != 0;      // warp-uniform (Params::tile_inert)
    NbState carry;   // new state of the left neighbour = this thread's own node of the previous column
    carry.rho = carry.y4 = carry.y5 = carry.y6 = carry.s7 = carry.s8 = carry.U = carry.V = carry.Tg = 0.;
    bool have_carry = false;

    for (int i = i0; i < i1; i++) {
        const long long n = valid_j ? (long long)i * ny + j : 0;
        uint32_t ct, tt;
        if (KC::plain) { ct = CT_NODE_IS_SET; tt = (uint32_t)KC::ttc | TT_IDXL | TT_IDXR | TT_IDYU | TT_IDYD; }
        else { ct = 0; tt = 0; if (valid_j) { ct = P.ct[n]; tt = P.tt[n]; } }
        const bool act = valid_j && node_active(ct);
        const bool fc = valid_j && !act && is_set(ct, NT_FC) && !is_set(ct, CT_SOLID);
        const bool upd = act || fc;
        const bool hasturb = KC::ns && (tt & (TCT_k_eps_Model | TCT_SA_Model));
        const bool turbN = KC::plain ? TURB_C : (KC::ns != 0);     // neighbours' S[7], S[8] are read
        // ---- the node's own new state (K1's output) ----
        NbState mine;
        mine.rho = mine.y4 = mine.y5 = mine.y6 = mine.s7 = mine.s8 = mine.U = mine.V = mine.Tg = 0.;
        if (valid_j) nb_load(P.S_out, P.U_out, P.V_out, P.Tg_out, N, n, turbN, mine);
#if K_PREFETCH
        if constexpr (KC::plain) {
            if (i + 1 < i1) {      // the next column's operands, requested under this column's arithmetic
                const long long m = n + ny;
#pragma unroll
                for (int k = 0; k < 7; k++) pf(&P.S_out[(long long)k * N + m]);
                if (TURB_C) { pf(&P.S_out[7 * N + m]); pf(&P.S_out[8 * N + m]); }
                pf(&P.U_out[m]); pf(&P.V_out[m]); pf(&P.Tg_out[m]); pf(&P.p_out[m]); pf(&P.kg_out[m]);
                if (!inert) { pf(&P.SN[4 * N + m]); pf(&P.SN[5 * N + m]); pf(&P.SN[6 * N + m]); }
                if (KC::ns) {
                    pf(&P.CP_in[m]); pf(&P.R_in[m]); pf(&P.mu_in[m]); pf(&P.kg_in[m]); pf(&P.mu_t[m]); pf(&P.lam_in[m]);
                    if (KC::ttc != 0) pf(&P.l_min[m]);
                    // its right neighbour's old state (its up neighbour's lies in the lines this column touches as `right`)
                    const long long r = m + ny;
                    pf(&P.S_in[r]); pf(&P.S_in[4 * N + r]); pf(&P.S_in[5 * N + r]); pf(&P.S_in[6 * N + r]);
                    if (TURB_C) { pf(&P.S_in[7 * N + r]); pf(&P.S_in[8 * N + r]); }
                    pf(&P.U_in[r]); pf(&P.V_in[r]); pf(&P.Tg_in[r]);
                }
            }
        }
#endif
        // down neighbour's new state: lane-1 holds it as its own; lane 0 reads the row below the tile
        NbState nD;
        nD.rho = shfl_up1(mine.rho); nD.y4 = shfl_up1(mine.y4); nD.y5 = shfl_up1(mine.y5); nD.y6 = shfl_up1(mine.y6);
        nD.U = shfl_up1(mine.U); nD.V = shfl_up1(mine.V); nD.Tg = shfl_up1(mine.Tg);
        if (turbN) { nD.s7 = shfl_up1(mine.s7); nD.s8 = shfl_up1(mine.s8); } else { nD.s7 = 0.; nD.s8 = 0.; }
        if (lane == 0 && valid_j && j > 0) nb_load(P.S_out, P.U_out, P.V_out, P.Tg_out, N, n - 1, turbN, nD);
        NbState nL = carry;
        if (!have_carry && valid_j && i > 0) nb_load(P.S_out, P.U_out, P.V_out, P.Tg_out, N, n - ny, turbN, nL);

        if (KC::ns && upd) {
            // ---- rebuild what the own-node half left in registers in the reference's single visit ----
            Local L;
            const bool turb2 = is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model);
#pragma unroll
            for (int k = 0; k < 4; k++) L.S[k] = P.S_out[(long long)k * N + n];
            L.S[7] = mine.s7; L.S[8] = mine.s8;
            if (!turbN || !(KC::plain || hasturb)) { L.S[7] = 0.; L.S[8] = 0.; }
            // species densities as FillNode2D sees them: committed, before chemistry
            if (inert) { L.S[4] = mine.y4; L.S[5] = mine.y5; L.S[6] = mine.y6; }      // the same bits as the shadow planes hold
            else
#pragma unroll
            for (int k = 4; k < 7; k++) {
                const bool com = KC::plain ? true : (act && eq_committed(ct, tt, k, KC::ns));
                L.S[k] = com ? P.SN[(long long)k * N + n] : P.S_out[(long long)k * N + n];
            }
            // S[k], S[eps] as committed, before TurbModRANS2D's rewrites (they differ only where local_rest rewrote them)
            L.s7c = L.S[7]; L.s8c = L.S[8];
            if (KC::plain) {
                if (TURB_C && st0.is_init) { L.s7c = P.SN[7 * N + n]; if (NE_C > 8) L.s8c = P.SN[8 * N + n]; }
            } else if (hasturb) {
                const int ne = num_eq(tt);
                L.s7c = (act && turb2 && 7 < ne && eq_committed(ct, tt, 7, KC::ns)) ? P.SN[7 * N + n] : P.S_in[7 * N + n];
                L.s8c = (act && turb2 && 8 < ne && eq_committed(ct, tt, 8, KC::ns)) ? P.SN[8 * N + n] : P.S_in[8 * N + n];
            }
            L.U = mine.U; L.V = mine.V; L.Tg = mine.Tg;
            L.p = P.p_out[n]; L.kg = P.kg_out[n];
            L.CP = P.CP_in[n]; L.R = P.R_in[n]; L.mu = P.mu_in[n];
            L.Uo = L.Vo = L.Tgo = L.po = 0.;
            if (!KC::plain) { L.Uo = P.U_in[n]; L.Vo = P.V_in[n]; }
            if (!KC::plain || (KC::ttc & TCT_SA_Model)) L.Tgo = P.Tg_in[n];
            if (!KC::plain || (KC::ttc & TCT_k_eps_Model)) { if (P.tem == TEM_k_eps_Chien) L.po = P.p_in[n]; }
            L.active = act; L.fc = fc;
            L.filled = !(L.S[iRho] == 0) && !(P.kg_in[n] < 1);
            // U, V as TurbModRANS2D sees them: before the wall-law / no-slip rewrites (hyper_flow_node.hpp:413-435 vs :447-488)
            L.Ut = L.U; L.Vt = L.V;
            if (!KC::plain && L.filled && (is_set(ct, CT_WALL_LAW) || is_set(ct, CT_WALL_NO_SLIP))) {
                DivExact dv0;
                const bool c1 = act && eq_committed(ct, tt, 1, KC::ns), c2 = act && eq_committed(ct, tt, 2, KC::ns);
                L.Ut = is_set(ct, CT_U_CONST) ? L.Uo : dv0.div(c1 ? P.SN[1 * N + n] : P.S_in[1 * N + n], L.S[iRho]);
                L.Vt = is_set(ct, CT_V_CONST) ? L.Vo : dv0.div(c2 ? P.SN[2 * N + n] : P.S_in[2 * N + n], L.S[iRho]);
            }
            const GlobalLoad ld{P, n};
            // values Stage 2 both reads and overwrites in place are fetched before the phase so that it can be re-run
            const double mu_t0 = P.mu_t[n], lam0 = P.lam_in[n];
            run([&](auto& dv) {
                StepState st = st0;
                if (fc) { st.is_init = 0; st.is_mu_t = 1; }   // NT_FC nodes: FillNode2D(1,0,...) regardless of TurbStartIter (:1331)
                const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0, n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
                SpecNode* sp = (!KC::plain && (tt & TT_SPEC)) ? &P.spec[P.spec_idx[n]] : nullptr;
                // ---- gradients (:1169-1238): left/down neighbours already updated this sweep, right/up not yet ----
                Grads g;
                if (act) {
                    NbState own, nR, nU, l_ = nL, d_ = nD;   // `own`: a missing neighbour aliases the node itself (committed S, old U,V,Tg)
                    own.rho = L.S[iRho]; own.y4 = L.S[4]; own.y5 = L.S[5]; own.y6 = L.S[6]; own.s7 = L.s7c; own.s8 = L.s8c;
                    own.U = L.Uo; own.V = L.Vo; own.Tg = L.Tgo;
                    if (KC::plain || n2) nb_load(P.S_in, P.U_in, P.V_in, P.Tg_in, N, n + ny, turbN && (KC::plain || hasturb), nR); else nR = own;
                    if (KC::plain || n3) nb_load(P.S_in, P.U_in, P.V_in, P.Tg_in, N, n + 1, turbN && (KC::plain || hasturb), nU); else nU = own;
                    if (!KC::plain) { if (!n1) l_ = own; if (!n4) d_ = own; if (!hasturb) { l_.s7 = l_.s8 = d_.s7 = d_.s8 = 0.; } }
                    const double dxn = (n1 + n2 == 2) ? P.dxn[1] : P.dxn[0], dym = (n3 + n4 == 2) ? P.dyn[1] : P.dyn[0];   // dx_1/n_n, dy_1/m_m (:1169)
                    double aR = nR.rho, aL = l_.rho, aU = nU.rho, aD = d_.rho;
                    const bool gx = !is_set(ct, CT_dYdx_NULL), gy = !is_set(ct, CT_dYdy_NULL);
                    // droYdx/droYdy keep their stored value when the species Neumann flag is set; nothing reads the
                    // species entries in that case except FillNode2D below -> they are kept in the special-node table
                    if (sp) { for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; } }
                    else { for (int q = 0; q < 4; q++) { g.droYdx[q] = 0.; g.droYdy[q] = 0.; } }
                    g.droYdx[3] = 0.; g.droYdy[3] = 0.;
                    if (gx) { g.droYdx[0] = (nR.y4 - l_.y4) * dxn; aR -= nR.y4; aL -= l_.y4; }
                    if (gy) { g.droYdy[0] = (nU.y4 - d_.y4) * dym; aU -= nU.y4; aD -= d_.y4; }
                    if (gx) { g.droYdx[1] = (nR.y5 - l_.y5) * dxn; aR -= nR.y5; aL -= l_.y5; }
                    if (gy) { g.droYdy[1] = (nU.y5 - d_.y5) * dym; aU -= nU.y5; aD -= d_.y5; }
                    if (gx) { g.droYdx[2] = (nR.y6 - l_.y6) * dxn; aR -= nR.y6; aL -= l_.y6; }
                    if (gy) { g.droYdy[2] = (nU.y6 - d_.y6) * dym; aU -= nU.y6; aD -= d_.y6; }
                    if (gx) g.droYdx[3] = (aR - aL) * dxn;
                    if (gy) g.droYdy[3] = (aU - aD) * dym;
                    g.dkdx = g.dkdy = g.depsdx = g.depsdy = 0.;
                    if (sp) { g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17]; }
                    if (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW)) {
                        g.dUdx = (nR.U * n1 - l_.U * n2) * dxn; g.dVdx = (nR.V * n1 - l_.V * n2) * dxn;
                        g.dUdy = (nU.U * n3 - d_.U * n4) * dym; g.dVdy = (nU.V * n3 - d_.V * n4) * dym;
                        if (is_set(tt, TCT_k_eps_Model)) {
                            g.dkdx = dv.div((nR.s7 * n1 - l_.s7 * n2) * dxn, L.S[iRho]); g.depsdx = dv.div((nR.s8 * n1 - l_.s8 * n2) * dxn, L.S[iRho]);
                            g.dkdy = dv.div((nU.s7 * n3 - d_.s7 * n4) * dym, L.S[iRho]); g.depsdy = dv.div((nU.s8 * n3 - d_.s8 * n4) * dym, L.S[iRho]);
                        } else if (is_set(tt, TCT_SA_Model)) {
                            g.dkdx = dv.div((nR.s7 * n1 - l_.s7 * n2) * dxn, L.S[iRho]); g.dkdy = dv.div((nU.s7 * n3 - d_.s7 * n4) * dym, L.S[iRho]);
                        }
                    } else {
                        g.dUdx = (nR.U - l_.U) * dxn; g.dVdx = (nR.V - l_.V) * dxn;
                        g.dUdy = (nU.U - d_.U) * dym; g.dVdy = (nU.V - d_.V) * dym;
                        if (is_set(tt, TCT_k_eps_Model)) {
                            g.dkdx = dv.div((nR.s7 - l_.s7) * dxn, L.S[iRho]); g.depsdx = dv.div((nR.s8 - l_.s8) * dxn, L.S[iRho]);
                            g.dkdy = dv.div((nU.s7 - d_.s7) * dym, L.S[iRho]); g.depsdy = dv.div((nU.s8 - d_.s8) * dym, L.S[iRho]);
                        } else if (is_set(tt, TCT_SA_Model)) {
                            g.dkdx = dv.div((nR.s7 - l_.s7) * dxn, L.S[iRho]); g.dkdy = dv.div((nU.s7 - d_.s7) * dym, L.S[iRho]);
                        }
                    }
                    g.dTdx = (nR.Tg - l_.Tg) * dxn; g.dTdy = (nU.Tg - d_.Tg) * dym;
                    // dUdy, dVdx of the wall nodes: read by Recalc_y_plus (deeps2d_core.cpp:2364-2386; hf2d_gpu_recalc_y_plus) ->
                    // kept in the special-node table; these two slots are read back by nobody on an active node
                    if constexpr (!KC::plain) {
                        if (sp && (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW))) { sp->grad[9] = g.dUdy; sp->grad[10] = g.dVdx; }
                    }
                    // every gradient member of the record, on request (the last step before a download): a run restarted from
                    // the swap file re-runs FillNode2D(0,1) on every node with them (InitDEEPS2D, deeps2d_core.cpp:4565; ScanArea :2204)
                    if (want_grads && P.grad) {
                        double* G = P.grad + n;
#pragma unroll
                        for (int q = 0; q < 4; q++) { G[(G_droYdx + q) * N] = g.droYdx[q]; G[(G_droYdy + q) * N] = g.droYdy[q]; }
                        G[G_dUdx * N] = g.dUdx; G[G_dUdy * N] = g.dUdy; G[G_dVdx * N] = g.dVdx; G[G_dVdy * N] = g.dVdy;
                        G[G_dTdx * N] = g.dTdx; G[G_dTdy * N] = g.dTdy;
                        G[G_dkdx * N] = g.dkdx; G[G_dkdy * N] = g.dkdy; G[G_depsdx * N] = g.depsdx; G[G_depsdy * N] = g.depsdy;
                    }
                } else {   // NT_FC node: FillNode2D runs on the gradients stored in the record (never refreshed)
                    for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; }
                    g.dUdx = sp->grad[8]; g.dUdy = sp->grad[9]; g.dVdx = sp->grad[10]; g.dVdy = sp->grad[11];
                    g.dTdx = sp->grad[12]; g.dTdy = sp->grad[13];
                    g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17];
                }
#include "hf2d_flux_fill.inc"
            });
        } else if (!KC::ns && upd) {
            // Euler: FillNode2D's inviscid fluxes only (no gradients)
            Local L;
#pragma unroll
            for (int k = 0; k < 4; k++) L.S[k] = P.S_out[(long long)k * N + n];
            if (inert) { L.S[4] = mine.y4; L.S[5] = mine.y5; L.S[6] = mine.y6; }
            else
#pragma unroll
            for (int k = 4; k < 7; k++) {
                const bool com = KC::plain ? true : (act && eq_committed(ct, tt, k, KC::ns));
                L.S[k] = com ? P.SN[(long long)k * N + n] : P.S_out[(long long)k * N + n];
            }
            L.S[7] = L.S[8] = 0.; L.s7c = L.s8c = 0.;
            L.U = mine.U; L.V = mine.V; L.Tg = mine.Tg; L.p = P.p_out[n]; L.kg = P.kg_out[n];
            L.CP = L.R = L.mu = L.Uo = L.Vo = L.Tgo = L.po = L.Ut = L.Vt = 0.;
            L.active = act; L.fc = fc;
            L.filled = !(L.S[iRho] == 0) && !(P.kg_in[n] < 1);
            const GlobalLoad ld{P, n};
            const double mu_t0 = 0., lam0 = 0.;
            DivExact dv;
            StepState st = st0;
            SpecNode* sp = nullptr;
            Grads g;
            (void)ld; (void)mu_t0; (void)lam0; (void)g; (void)sp;
#include "hf2d_flux_fill.inc"
        }
        carry = mine; have_carry = true;
    }
}
