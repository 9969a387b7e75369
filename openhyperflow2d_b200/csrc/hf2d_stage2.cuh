// hf2d_stage2.cuh -- Stage 2 of the DEEPS2D step (deeps2d_core.cpp:1028-1334) as ONE kernel:
// residual + blending factor, commit, mixed old/new gradients, FillNode2D, TurbModRANS2D, CFL dt, chemistry,
// with warp-shuffle block reductions for dt-min and the residual sums.
//
// Reads the *_in planes, writes the *_out planes (ping-pong) because neighbours read both a node's last-step
// state (right/up in the sweep) and its recomputed new state (left/down), see hf2d_local.cuh.
#pragma once
#include "hf2d_turb.cuh"

struct NbState { double rho, y4, y5, y6, s7, s8, U, V, Tg; };

__device__ __forceinline__ void nb_old(const Params& P, long long m, bool turb, NbState& s) {
    const long long N = P.N;
    s.rho = P.S_in[m]; s.y4 = P.S_in[4 * N + m]; s.y5 = P.S_in[5 * N + m]; s.y6 = P.S_in[6 * N + m];
    s.s7 = turb ? P.S_in[7 * N + m] : 0.; s.s8 = turb ? P.S_in[8 * N + m] : 0.;
    s.U = P.U_in[m]; s.V = P.V_in[m]; s.Tg = P.Tg_in[m];
}
__device__ __forceinline__ void nb_new(const Params& P, const StepState& st, long long m, NbState& s) {
    Local L;
    local_update<false>(P, st, m, P.ct[m], P.tt[m], L);
    s.rho = L.S[iRho]; s.y4 = L.y4; s.y5 = L.y5; s.y6 = L.y6; s.s7 = L.S[7]; s.s8 = L.S[8];
    s.U = L.U; s.V = L.V; s.Tg = L.Tg;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long w = __shfl_down_sync(0xffffffffu, v, o); v = w > v ? w : v; }
    return v;
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long w = __shfl_down_sync(0xffffffffu, v, o); v = w < v ? w : v; }
    return v;
}

#define S2_BX 64
#define S2_BY 4

__global__ void __launch_bounds__(S2_BX * S2_BY) k_stage2(Params P) {
    __shared__ double sh_d[2 * NEQ][S2_BX * S2_BY / 32];
    __shared__ unsigned long long sh_u[2 * NEQ + 2][S2_BX * S2_BY / 32];
    StepState st = *P.st;         // snapshot (the accumulators are only touched through atomics)
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = P.i_lo + blockIdx.y * blockDim.y + threadIdx.y;
    const bool inb = (j < P.ny && i < P.i_hi);
    const long long ny = P.ny, N = P.N, n = inb ? (long long)i * ny + j : 0;
    // per-thread reduction inputs
    double r_rms[NEQ], r_div[NEQ];
    unsigned long long r_cnt[NEQ], r_dd[NEQ];
#pragma unroll
    for (int k = 0; k < NEQ; k++) { r_rms[k] = 0.; r_div[k] = 0.; r_cnt[k] = 0ull; r_dd[k] = 0ull; }
    unsigned long long r_dt = ~0ull, r_err = ~0ull;

    uint32_t ct = 0, tt = 0;
    if (inb) { ct = P.ct[n]; tt = P.tt[n]; }
    const bool act = inb && node_active(ct);
    const bool fc = inb && !act && is_set(ct, NT_FC) && !is_set(ct, CT_SOLID);
    if (act || fc) {
        if (fc) { st.is_init = 0; st.is_mu_t = 1; }   // NT_FC nodes: FillNode2D(1,0,...) regardless of TurbStartIter (:1331)
        const int ne = num_eq(tt);
        const bool turb2 = P.ns && (is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model));
        const bool hasturb = P.ns && (tt & (TCT_k_eps_Model | TCT_SA_Model));
        const int n1 = (tt & TT_IDXL) ? 1 : 0, n2 = (tt & TT_IDXR) ? 1 : 0, n3 = (tt & TT_IDYU) ? 1 : 0, n4 = (tt & TT_IDYD) ? 1 : 0;
        SpecNode* sp = (tt & TT_SPEC) ? &P.spec[P.spec_idx[n]] : nullptr;
        // ---- residual and locally adaptive blending factor (:1062-1154); uses S after Stage-1 overwrite ----
        if (act) {
            for (int k = 0; k < ne; k++) {
                uint32_t c_flag = 0;
                if (k < 4) c_flag = CT_Rho_CONST << k;
                else if (k < 4 + NCOMP) c_flag = CT_Y_CONST;
                else if (turb2) c_flag = TCT_k_CONST << (k - 4 - NCOMP);
                const double s1 = P.S_in[(long long)k * N + n];
                // quirk kept: the gate tests the CT word even for k >= 7 (:1075)
                if (!is_set(ct, c_flag) && s1 != 0.) {
                    const double Tmp = s1;
                    const double absDD = P.SN[(long long)k * N + n] - s1;
                    double DD, sqrt_RES = 0;
                    if (fabs(Tmp) > 1.e-15) { DD = fabs(absDD / Tmp); sqrt_RES = sqrt(DD); } else { DD = 1.0; }
                    double beta_min = OMIN(P.beta0, st.beta_scn);
                    if (is_set(ct, CT_NONREFLECTED)) beta_min = P.nrbc_beta0;
                    double* bp = &P.beta[(long long)k * N + n];
                    const double bb = beta_min * beta_min;
                    double bnew;
                    switch (P.bFF) {
                    case 0: bnew = OMIN(beta_min, bb / (beta_min + DD)); break;
                    case 1: bnew = OMIN((beta_min + *bp) * 0.5, bb / (beta_min + DD)); break;
                    case 2: bnew = OMIN(beta_min, bb / (beta_min + DD * DD)); break;
                    case 3: bnew = OMIN((beta_min + *bp) * 0.5, bb / (beta_min + DD * DD)); break;
                    case 4: bnew = OMIN(beta_min, bb / (beta_min + sqrt_RES)); break;
                    case 5: bnew = OMIN((beta_min + *bp) * 0.5, bb / (beta_min + sqrt_RES)); break;
                    default: bnew = *bp; break;
                    }
                    *bp = bnew;
                    if (P.isAlternateRMS) { r_div[k] = Tmp * Tmp; r_rms[k] = absDD; } else { r_rms[k] = DD * DD; }
                    r_cnt[k] = 1ull;
                    r_dd[k] = d2bits(DD);
                }
            }
        }
        // ---- commit + the local half of FillNode2D + species chemistry ----
        Local L;
        local_update<true>(P, st, n, ct, tt, L);
        // ---- gradients (:1169-1238): left/down neighbours already updated this sweep, right/up not yet ----
        Grads g;
        const double yj = P.axi ? P.yrow[j] : 1.0;
        if (P.ns) {
            if (act) {
                NbState own, nR, nL, nU, nD;   // `own`: a missing neighbour aliases the node itself (committed S, old U,V,Tg)
                own.rho = L.S[iRho]; own.y4 = L.S[4]; own.y5 = L.S[5]; own.y6 = L.S[6]; own.s7 = L.s7c; own.s8 = L.s8c;
                own.U = L.Uo; own.V = L.Vo; own.Tg = L.Tgo;
                if (n2) nb_old(P, n + ny, hasturb, nR); else nR = own;
                if (n3) nb_old(P, n + 1, hasturb, nU); else nU = own;
                if (n1) nb_new(P, st, n - ny, nL); else nL = own;
                if (n4) nb_new(P, st, n - 1, nD); else nD = own;
                const double n_n = OMAX(n1 + n2, 1), m_m = OMAX(n3 + n4, 1);
                const double dx_1 = 1.0 / P.dx, dy_1 = 1.0 / P.dy;
                const double dxn = dx_1 / n_n, dym = dy_1 / m_m;
                double aR = nR.rho, aL = nL.rho, aU = nU.rho, aD = nD.rho;
                const bool gx = !is_set(ct, CT_dYdx_NULL), gy = !is_set(ct, CT_dYdy_NULL);
                // droYdx/droYdy keep their stored value when the species Neumann flag is set; nothing reads the
                // species entries in that case except FillNode2D below -> they are kept in the special-node table
                if (sp) { for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; } }
                else { for (int q = 0; q < 4; q++) { g.droYdx[q] = 0.; g.droYdy[q] = 0.; } }
                g.droYdx[3] = 0.; g.droYdy[3] = 0.;
                if (gx) {
                    g.droYdx[0] = (nR.y4 - nL.y4) * dxn; aR -= nR.y4; aL -= nL.y4;
                }
                if (gy) { g.droYdy[0] = (nU.y4 - nD.y4) * dym; aU -= nU.y4; aD -= nD.y4; }
                if (gx) { g.droYdx[1] = (nR.y5 - nL.y5) * dxn; aR -= nR.y5; aL -= nL.y5; }
                if (gy) { g.droYdy[1] = (nU.y5 - nD.y5) * dym; aU -= nU.y5; aD -= nD.y5; }
                if (gx) { g.droYdx[2] = (nR.y6 - nL.y6) * dxn; aR -= nR.y6; aL -= nL.y6; }
                if (gy) { g.droYdy[2] = (nU.y6 - nD.y6) * dym; aU -= nU.y6; aD -= nD.y6; }
                if (gx) g.droYdx[3] = (aR - aL) * dxn;
                if (gy) g.droYdy[3] = (aU - aD) * dym;
                g.dkdx = g.dkdy = g.depsdx = g.depsdy = 0.;
                if (sp) { g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17]; }
                if (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW)) {
                    g.dUdx = (nR.U * n1 - nL.U * n2) * dxn; g.dVdx = (nR.V * n1 - nL.V * n2) * dxn;
                    g.dUdy = (nU.U * n3 - nD.U * n4) * dym; g.dVdy = (nU.V * n3 - nD.V * n4) * dym;
                    if (is_set(tt, TCT_k_eps_Model)) {
                        g.dkdx = (nR.s7 * n1 - nL.s7 * n2) * dxn / L.S[iRho]; g.depsdx = (nR.s8 * n1 - nL.s8 * n2) * dxn / L.S[iRho];
                        g.dkdy = (nU.s7 * n3 - nD.s7 * n4) * dym / L.S[iRho]; g.depsdy = (nU.s8 * n3 - nD.s8 * n4) * dym / L.S[iRho];
                    } else if (is_set(tt, TCT_SA_Model)) {
                        g.dkdx = (nR.s7 * n1 - nL.s7 * n2) * dxn / L.S[iRho]; g.dkdy = (nU.s7 * n3 - nD.s7 * n4) * dym / L.S[iRho];
                    }
                } else {
                    g.dUdx = (nR.U - nL.U) * dxn; g.dVdx = (nR.V - nL.V) * dxn;
                    g.dUdy = (nU.U - nD.U) * dym; g.dVdy = (nU.V - nD.V) * dym;
                    if (is_set(tt, TCT_k_eps_Model)) {
                        g.dkdx = (nR.s7 - nL.s7) * dxn / L.S[iRho]; g.depsdx = (nR.s8 - nL.s8) * dxn / L.S[iRho];
                        g.dkdy = (nU.s7 - nD.s7) * dym / L.S[iRho]; g.depsdy = (nU.s8 - nD.s8) * dym / L.S[iRho];
                    } else if (is_set(tt, TCT_SA_Model)) {
                        g.dkdx = (nR.s7 - nL.s7) * dxn / L.S[iRho]; g.dkdy = (nU.s7 - nD.s7) * dym / L.S[iRho];
                    }
                }
                g.dTdx = (nR.Tg - nL.Tg) * dxn; g.dTdy = (nU.Tg - nD.Tg) * dym;
            } else {   // NT_FC node: FillNode2D runs on the gradients stored in the record (never refreshed)
                for (int q = 0; q < 4; q++) { g.droYdx[q] = sp->grad[q]; g.droYdy[q] = sp->grad[4 + q]; }
                g.dUdx = sp->grad[8]; g.dUdy = sp->grad[9]; g.dVdx = sp->grad[10]; g.dVdy = sp->grad[11];
                g.dTdx = sp->grad[12]; g.dTdy = sp->grad[13];
                g.dkdx = sp->grad[14]; g.dkdy = sp->grad[15]; g.depsdx = sp->grad[16]; g.depsdy = sp->grad[17];
            }
        }
#include "hf2d_stage2_fill.inc"
    }
#include "hf2d_stage2_reduce.inc"
}
