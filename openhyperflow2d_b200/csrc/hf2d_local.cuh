// hf2d_local.cuh -- the part of Stage 2 that depends on a node's OWN data only.
//
// In the reference's in-place sweep (deeps2d_core.cpp:1028-1334) node C reads its left/down neighbours AFTER
// they went through commit + FillNode2D + CalcChemicalReactions in this sweep and its right/up neighbours BEFORE
// (SURVEY.md F4).  What a neighbour exposes (S[0], S[4..8], U, V, Tg) depends only on that neighbour's own
// shadow S, flags and lagged CP/R -- never on gradients -- so any thread can recompute a neighbour's "new state"
// with local_update() instead of waiting for it.  The same function is the first half of the node's own update.
#pragma once
#include "hf2d_dev.cuh"

// pow(x, 3/2) of the k-eps boundary rewrites (hyper_flow_node.hpp:760,775,779)
// The reference calls glibc's pow, which is correctly rounded in all but ~0.08 % of its results (measured against the
// double-double evaluation below over 2e7 random arguments); CUDA's pow is a 1-2 ulp function.  x^1.5 = x * sqrt(x) with
// sqrt(x) = s + e (s the rounded root, e = (x - s*s) / 2s, the residual exact by FMA) and the product x*s split exactly
// by FMA, rounded ONCE: the correctly rounded result, i.e. glibc's bit for bit in 99.92 % of the calls, in ~8 FP64
// instructions instead of the ~150 of the library routine.  Every last-bit difference here is a seed that the k-eps
// system amplifies step after step (tests/test_zz_r2_parity_holes_gpu.py, drift cases), so fewer seeds = less drift.
__device__ __forceinline__ double pow15(double x) {
    if (!(x > 0.) || !(x < __longlong_as_double(0x7ff0000000000000ll))) return pow(x, 1.5);   // 0, negative, inf, NaN: the library's answers
    const double s = sqrt(x);
    const double r = fma(-s, s, x);
    const double e = r / (2.0 * s);
    const double p = x * s;
    const double pe = fma(x, s, -p);
    return p + (pe + x * e);
}

struct Local {
    double S[NEQ];          // committed S after FillNode2D's local rewrites, BEFORE chemistry
    double s7c, s8c;        // committed S[7], S[8] before TurbModRANS2D rewrote them
    double U, V, Tg, p, kg; // after FillNode2D
    double Uo, Vo, Tgo, po; // before (last step's)
    double CP, R, mu;       // lagged properties (last step's chemistry)
    double y4, y5, y6;      // S[4..6] after chemistry (what later nodes of the sweep read)
    double Yfu, Yox, Ycp, Yair;   // mass fractions inside chemistry at the point where R/CP/lam/mu are formed
    double Y[4];            // FlowNode2D::Y after chemistry
    double Upre, Vpre;      // S[RhoU]/S[Rho], S[RhoV]/S[Rho] of a no-slip wall node before the wall reset (virtual sources)
    double Ut, Vt;          // U, V as TurbModRANS2D sees them: after the U_CONST/V_CONST block, BEFORE the wall-law /
                            // no-slip rewrites (hyper_flow_node.hpp:413-435 run before :447-488)
    bool   filled;          // FillNode2D ran past its early returns
    bool   active, fc;
};

// commit test of equation k (deeps2d_core.cpp:1156-1164)
__device__ __forceinline__ bool eq_committed(uint32_t ct, uint32_t tt, int k, int ns) {
    if (k < 4) return !is_set(ct, CT_Rho_CONST << k);
    if (k < 4 + NCOMP) return !is_set(ct, CT_Y_CONST);
    if (ns && (is_set(tt, TCT_k_eps_Model) || is_set(tt, TCT_SA_Model))) return !is_set(tt, TCT_k_CONST << (k - 4 - NCOMP));
    return false;
}

// Kernel configuration: solver mode and flow type are compile-time; TTC >= 0 marks a "plain" instantiation whose
// nodes are all unflagged interior nodes (CT == CT_NODE_IS_SET, TurbType == TTC, four neighbours present), so that
// every flag test below constant-folds; TTC < 0 reads the flag words of each node.
template <int NS, int AXI, int TTC> struct KCfg {
    static constexpr int ns = NS, axi = AXI, ttc = TTC;
    static constexpr bool plain = TTC >= 0;
};

#ifndef K1B_LAZY
#define K1B_LAZY 2                // plain tiles fetch last step's U, V, Tg, p only where FillNode2D returns early (S[Rho] == 0 or k < 1: never on
                                  // a sane grid).  1 = decided per lane: SLOWER than fetching them always (0.444 vs 0.416 ms: divergent branch
                                  // in front of four dependent loads); 2 = one warp-uniform decision (__any_sync): 0.429 -> 0.422 ms.  0 = always.
#endif
// Where a node's own inputs come from: last step's planes.
struct GlobalLoad {
    const Params& P; long long m;
    __device__ __forceinline__ double S(int k) const { return P.S_in[(long long)k * P.N + m]; }
    __device__ __forceinline__ double SN(int k) const { return P.SN[(long long)k * P.N + m]; }
    __device__ __forceinline__ double U() const { return P.U_in[m]; }
    __device__ __forceinline__ double V() const { return P.V_in[m]; }
    __device__ __forceinline__ double Tg() const { return P.Tg_in[m]; }
    __device__ __forceinline__ double CP() const { return P.CP_in[m]; }
    __device__ __forceinline__ double R() const { return P.R_in[m]; }
    __device__ __forceinline__ double kg() const { return P.kg_in[m]; }
    __device__ __forceinline__ double p() const { return P.p_in[m]; }
    __device__ __forceinline__ double mu() const { return P.mu_in[m]; }
    __device__ __forceinline__ double mu_t() const { return P.mu_t[m]; }
    __device__ __forceinline__ double lam() const { return P.lam_in[m]; }
    __device__ __forceinline__ double lmin() const { return P.l_min[m]; }
};
// The part of local_update that follows the commit loop: L.S[] holds the committed S of node m.
// FULL=false computes only what neighbours read (S[0], S[4..8] post-chemistry, U, V, Tg).
template <class KC, bool FULL, class D, class LD>
__device__ __forceinline__ void local_rest(D& dv, const LD& ld, const Params& P, const StepState& st, long long m, uint32_t ct, uint32_t tt, Local& L) {
    const bool act = node_active(ct);
    const bool fc = !act && is_set(ct, NT_FC) && !is_set(ct, CT_SOLID);
    L.active = act; L.fc = fc; L.filled = false;
    // NT_FC nodes always run FillNode2D(1,0,...) (deeps2d_core.cpp:1331), active nodes follow TurbStartIter (:1240-1244)
    const int is_init = fc ? 0 : st.is_init;
    L.s7c = L.S[7]; L.s8c = L.S[8];
    L.CP = ld.CP(); L.R = ld.R();
    L.mu = KC::ns ? ld.mu() : 0.;
    const double kg_old = ld.kg();
    L.kg = kg_old;
    // Last step's U, V, Tg, p survive only where FillNode2D returns early (S[Rho] == 0 or k < 1) or a CONST flag pins them:
    // an unflagged interior node overwrites all four, so the plain-tile kernels fetch them only in the early-return case
    // (4 planes = 32 B per cell that the bulk of the grid never reads).
    if (!KC::plain || !K1B_LAZY || (K1B_LAZY == 2 ? __any_sync(0xffffffffu, (L.S[iRho] == 0) || (kg_old < 1)) : ((L.S[iRho] == 0) || (kg_old < 1)))) {   // (2: one warp-uniform decision)
        L.Uo = ld.U(); L.Vo = ld.V(); L.Tgo = ld.Tg();
        if (FULL) L.po = ld.p();
    } else { L.Uo = L.Vo = L.Tgo = 0.; if (FULL) L.po = 0.; }
    L.U = L.Uo; L.V = L.Vo; L.Tg = L.Tgo;
    L.Ut = L.Uo; L.Vt = L.Vo;
    if (FULL) L.p = L.po;
    if (!act && !fc) { L.y4 = L.S[4]; L.y5 = L.S[5]; L.y6 = L.S[6]; return; }
    // ---- FillNode2D, local part (hyper_flow_node.hpp:382-492) ----
    if (!(L.S[iRho] == 0) && !(kg_old < 1)) {
        L.filled = true;
        L.kg = dv.div(L.CP, L.CP - L.R);
        if (is_set(ct, CT_U_CONST)) L.S[iRhoU] = L.U * L.S[iRho]; else L.U = dv.div(L.S[iRhoU], L.S[iRho]);
        if (is_set(ct, CT_V_CONST)) L.S[iRhoV] = L.V * L.S[iRho]; else L.V = dv.div(L.S[iRhoV], L.S[iRho]);
        L.Ut = L.U; L.Vt = L.V;
        if (KC::ns && (tt & TT_MASK) != 0) {
            // the S-rewriting part of TurbModRANS2D (:753-779 k-eps, :856-862 SA)
            if (is_set(tt, TCT_Prandtl_Model)) {
            } else if (is_set(tt, TCT_k_eps_Model)) {
                const double l = OMAX(ld.lmin(), OMIN(P.dy, P.dx)) * 0.41;
                const double p34 = (P.tem == TEM_k_eps_RNG) ? P.pow_cmu_34[1] : P.pow_cmu_34[0];
                if (is_init) {
                    const double TmpI = P.turb_I * sqrt(L.U * L.U + L.V * L.V + 1.e-30);
                    L.S[iK] = 1.5 * TmpI * TmpI * L.S[iRho];
                    L.S[iEps] = p34 * pow15(L.S[iK] / L.S[iRho]) / l;
                }
                if (is_set(tt, TCT_k_CONST)) {
                    const double TmpI = P.turb_I * sqrt(L.U * L.U + L.V * L.V + 1.e-30);
                    L.S[iK] = 1.5 * TmpI * TmpI * L.S[iRho];
                }
                if (is_set(tt, TCT_eps_CONST)) L.S[iEps] = p34 * pow15(L.S[iK] / L.S[iRho]) / l;
                if (is_set(tt, TCT_eps_Cmk2kXn_WALL)) L.S[iEps] = p34 * pow15(L.S[iK] / L.S[iRho]) / l;
            } else if (is_set(tt, TCT_SA_Model)) {
                if (is_init) L.S[iNut] = L.mu / L.S[iRho] / 100.0;
                else if (is_set(ct, CT_WALL_NO_SLIP) || is_set(ct, CT_WALL_LAW) || is_set(tt, TCT_k_CONST)) L.S[iNut] = 0.;
                else if (is_set(ct, NT_FC)) L.S[iNut] = L.mu / L.S[iRho] * P.turb_I;
            }
        }
        double Tmp1 = L.S[iRho], Tmp3 = 0.;
#pragma unroll
        for (int i = 0; i < NCOMP; i++) { Tmp3 += P.Hu[i] * L.S[i + 4]; Tmp1 -= L.S[i + 4]; }
        Tmp3 += P.Hu[NCOMP] * Tmp1;
        if (is_set(ct, CT_WALL_LAW)) {
            const SpecNode& sp = P.spec[P.spec_idx[m]];
            Tmp1 = sqrt(L.U * L.U + L.V * L.V + 1.e-30);
            L.S[iRhoU] = Tmp1 * sp.BGX;
            L.S[iRhoV] = Tmp1 * sp.BGY;
            L.U = dv.div(L.S[iRhoU], L.S[iRho]);
            L.V = dv.div(L.S[iRhoV], L.S[iRho]);
        } else if (is_set(ct, CT_WALL_NO_SLIP)) {
            const SpecNode& sp = P.spec[P.spec_idx[m]];
            if (FULL) { L.Upre = L.S[iRhoU] / L.S[iRho]; L.Vpre = L.S[iRhoV] / L.S[iRho]; }
            L.U = sp.Uw; L.V = sp.Vw;
            L.S[iRhoU] = L.U * L.S[iRho];
            L.S[iRhoV] = L.V * L.S[iRho];
        }
        const double pp = (L.kg - 1.) * (L.S[iRhoE] - L.S[iRho] * (L.U * L.U + L.V * L.V) * 0.5 - Tmp3);
        L.Tg = dv.div(dv.div(pp, L.R), L.S[iRho]);
        if (FULL) L.p = pp;
    }
    // ---- species part of CalcChemicalReactions (deeps2d_core.cpp:4702-4728, 4748-4777); active nodes only ----
    L.y4 = L.S[4]; L.y5 = L.S[5]; L.y6 = L.S[6];
    if (act) {
        double Yfu = dv.div(L.S[iYfu], L.S[0]), Yox = dv.div(L.S[iYox], L.S[0]), Ycp = dv.div(L.S[iYcp], L.S[0]);
        double Yair = 1. - (Yfu + Yox + Ycp);
        const bool yconst = is_set(ct, CT_Y_CONST);
        if (!yconst) {
            double Y0 = dv.div(1., Yfu + Yox + Ycp + Yair);
            Yfu = Yfu * Y0; Yox = Yox * Y0; Ycp = Ycp * Y0;
            const double Tf = P.tf_uniform ? P.Tf0 : P.Tf[m];
            if (L.Tg > Tf) {
                if (Yox > Yfu * P.K0) { Yox = Yox - Yfu * P.K0; Yfu = 0.; Ycp = 1. - Yox - Yair; }
                else { Yfu = Yfu - dv.div(Yox, P.K0); Yox = 0.; Ycp = 1. - Yfu - Yair; }
            }
        }
        if (FULL) { L.Yfu = Yfu; L.Yox = Yox; L.Ycp = Ycp; L.Yair = Yair; }
        if (Yair < 1.e-5) Yair = 0.;
        if (Ycp < 1.e-8) Ycp = 0.;
        if (Yox < 1.e-8) Yox = 0.;
        if (Yfu < 1.e-8) Yfu = 0.;
        const double Y0 = dv.div(1., Yfu + Yox + Ycp + Yair);
        Yfu = Yfu * Y0; Yox = Yox * Y0; Ycp = Ycp * Y0; Yair = Yair * Y0;
        if (FULL) { L.Y[0] = Yfu; L.Y[1] = Yox; L.Y[2] = Ycp; L.Y[3] = Yair; }
        if (!yconst) { L.y4 = fabs(Yfu * L.S[0]); L.y5 = fabs(Yox * L.S[0]); L.y6 = fabs(Ycp * L.S[0]); }
    }
}
